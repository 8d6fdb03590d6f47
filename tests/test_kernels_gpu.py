"""Kernel-level parity: each CUDA entry point (called through the C ABI) against the oracle's
restatement of the same op (oracle/gpt2_lora_ref.py) or plain fp32 torch on the CPU.

Tolerances: AdamW <= 1e-6 relative on identical inputs; cross-entropy loss
<= 1e-5 vs fp32, gradient within 1 bf16 ulp of the fp32 gradient of the bf16 logits; GEMM
<= 1e-2 relative to the fp32 product of the bf16-rounded operands.
"""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import gpt2_lora_ref as ref  # noqa: E402  (the checker; tests are allowed to import it)


@pytest.fixture(scope="module")
def ops():
    from b200ft import ops as _ops
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    return _ops


DEV = "cuda:0"


def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


# ----------------------------------------------------------------------------- AdamW
@pytest.mark.parametrize("n", [1, 7, 8, 1000, 294912 + 3])
@pytest.mark.parametrize("g_bf16", [False, True])
def test_adamw_matches_torch_optim(ops, n, g_bf16):
    g = torch.Generator().manual_seed(n)
    o = ref.AdamWConfig(lr=3e-3, beta1=0.9, beta2=0.95, eps=1e-8, weight_decay=0.1)
    p0 = torch.randn(n, generator=g)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.AdamW([p_ref], lr=o.lr, betas=(o.beta1, o.beta2), eps=o.eps, weight_decay=o.weight_decay,
                            foreach=False, fused=False)
    p = p0.to(DEV)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    shadow = torch.empty(n, dtype=torch.bfloat16, device=DEV)
    scale = 0.5
    for step in range(1, 4):
        grad = torch.randn(n, generator=g) * 0.1
        if g_bf16:
            grad = bf16_round(grad)
        p_ref.grad = grad * scale
        opt.step()
        gd = grad.to(DEV, torch.bfloat16 if g_bf16 else torch.float32)
        ops.adamw_step(p, m, v, gd, step=step, lr=o.lr, beta1=o.beta1, beta2=o.beta2, eps=o.eps,
                       weight_decay=o.weight_decay, shadow=shadow, grad_scale=scale)
        torch.testing.assert_close(p.cpu(), p_ref.detach(), rtol=1e-6, atol=1e-7)
        # and against the written-out formula of the oracle
    st = opt.state[p_ref]
    torch.testing.assert_close(m.cpu(), st["exp_avg"], rtol=1e-6, atol=1e-8)
    torch.testing.assert_close(v.cpu(), st["exp_avg_sq"], rtol=1e-6, atol=1e-10)
    assert torch.equal(shadow.cpu(), p.cpu().to(torch.bfloat16))


def test_adamw_device_scale_and_gradnorm(ops):
    n = 5000
    g = torch.Generator().manual_seed(0)
    grad = torch.randn(n, generator=g)
    gd = grad.to(DEV)
    ws = ops.gradnorm_workspace(n, DEV)
    nsq = torch.zeros(1, device=DEV)
    coef = torch.zeros(1, device=DEV)
    for _ in range(2):  # workspace must be reusable
        ops.gradnorm_sq(gd, nsq, ws)
        torch.testing.assert_close(nsq.cpu()[0], (grad.double() ** 2).sum().float(), rtol=1e-5, atol=0)
    ops.clip_coef(nsq, coef, max_norm=1.0, pre_scale=0.5)
    expect = min(1.0, 1.0 / (0.5 * float(grad.norm()) + 1e-6))
    assert abs(float(coef.cpu()) - expect) < 1e-6
    o = ref.AdamWConfig()
    p = torch.ones(n, device=DEV)
    m = torch.zeros(n, device=DEV)
    v = torch.zeros(n, device=DEV)
    ops.adamw_step(p, m, v, gd, step=1, lr=o.lr, beta1=o.beta1, beta2=o.beta2, eps=o.eps,
                   weight_decay=o.weight_decay, grad_scale=0.5, grad_scale_dev=coef)
    pr, mr, vr = ref.adamw_ref(torch.ones(n), grad, torch.zeros(n), torch.zeros(n), 1, o, grad_scale=0.5 * expect)
    torch.testing.assert_close(p.cpu(), pr, rtol=1e-6, atol=1e-7)
    torch.testing.assert_close(m.cpu(), mr, rtol=1e-5, atol=1e-8)


# ----------------------------------------------------------------------------- softmax-CE
@pytest.mark.parametrize("rows,vocab,ld", [(4, 503, 512), (64, 50257, 50304), (3, 8, 8), (5, 1000, 1024)])
def test_softmax_xent_matches_oracle(ops, rows, vocab, ld):
    g = torch.Generator().manual_seed(rows * 1000 + vocab)
    z = bf16_round(torch.randn(rows, vocab, generator=g) * 3.0)
    targets = torch.randint(0, vocab, (rows,), generator=g)
    if rows > 2:
        targets[1] = -100  # an ignored row
    loss_ref, dz_ref = ref.softmax_xent_ref(z, targets)
    zz = torch.full((rows, ld), 7.0)  # garbage in the padding columns must not matter
    zz[:, :vocab] = z
    zd = zz.to(DEV, torch.bfloat16)
    loss = torch.zeros(1, device=DEV)
    ws = ops.softmax_xent_workspace(rows, DEV)
    ops.softmax_xent_fwd_bwd(zd, targets.to(DEV), loss, ws, vocab=vocab)
    assert abs(float(loss.cpu()) - float(loss_ref)) <= 1e-5 * max(1.0, abs(float(loss_ref)))
    dz = zd.cpu().float()
    assert torch.all(dz[:, vocab:] == 0)
    # 1 bf16 ulp of the fp32 gradient (relative 2^-8) plus a small absolute floor
    torch.testing.assert_close(dz[:, :vocab], dz_ref, rtol=2 ** -7, atol=1e-7)
    if rows > 2:
        assert torch.all(dz[1] == 0)


def test_softmax_xent_all_ignored_and_scale(ops):
    rows, vocab = 4, 64
    z = bf16_round(torch.randn(rows, vocab))
    zd = z.to(DEV, torch.bfloat16)
    loss = torch.ones(1, device=DEV)
    ws = ops.softmax_xent_workspace(rows, DEV)
    ops.softmax_xent_fwd_bwd(zd, torch.full((rows,), -100, device=DEV), loss, ws, vocab=vocab)
    assert float(loss.cpu()) == 0.0 and torch.all(zd == 0)
    targets = torch.randint(0, vocab, (rows,))
    _, dz_ref = ref.softmax_xent_ref(z, targets)
    zd = z.to(DEV, torch.bfloat16)
    ops.softmax_xent_fwd_bwd(zd, targets.to(DEV), loss, ws, vocab=vocab, grad_scale=0.25)
    torch.testing.assert_close(zd.cpu().float(), 0.25 * dz_ref, rtol=2 ** -7, atol=1e-7)


def test_softmax_xent_out_of_range_targets_are_ignored_and_counted(ops):
    """A target >= vocab or < 0 (other than ignore_index) gets no loss and a zero gradient row, is left out of the
    mean's denominator, and is counted for the host — loss and gradient stay consistent (torch raises here)."""
    rows, vocab = 6, 200
    g = torch.Generator().manual_seed(5)
    z = bf16_round(torch.randn(rows, vocab, generator=g) * 2)
    targets = torch.randint(0, vocab, (rows,), generator=g)
    bad = targets.clone()
    bad[0], bad[4] = vocab, -7
    as_ignored = targets.clone()
    as_ignored[0] = as_ignored[4] = -100
    loss_ref, dz_ref = ref.softmax_xent_ref(z, as_ignored)
    zd = z.to(DEV, torch.bfloat16)
    loss = torch.zeros(1, device=DEV)
    ws = ops.softmax_xent_workspace(rows, DEV)
    ops.softmax_xent_fwd_bwd(zd, bad.to(DEV), loss, ws, vocab=vocab)
    assert abs(float(loss.cpu()) - float(loss_ref)) <= 1e-5 * max(1.0, abs(float(loss_ref)))
    torch.testing.assert_close(zd.cpu().float(), dz_ref, rtol=2 ** -7, atol=1e-7)
    assert torch.all(zd[0] == 0) and torch.all(zd[4] == 0)
    assert ops.softmax_xent_bad_targets(ws) == 2
    zd = z.to(DEV, torch.bfloat16)
    ops.softmax_xent_fwd_bwd(zd, targets.to(DEV), loss, ws, vocab=vocab)
    assert ops.softmax_xent_bad_targets(ws) == 0


@pytest.mark.parametrize("rows,vocab,ld", [(700, 1000, 1024), (9, 40000, 40000), (5, 262144, 262144), (3, 300000, 300000)])
def test_softmax_xent_many_rows_and_every_kernel_choice(ops, rows, vocab, ld):
    """More rows than SMs (every persistent CTA loops, the 3-stage ring wraps), ignored rows in between, and row
    lengths that select the staged kernel (<= 32768 logits), the two-sweep kernel at 8 x 256 threads per SM and the
    two-sweep kernel with capped 512-thread CTAs (long rows)."""
    g = torch.Generator().manual_seed(rows + vocab)
    z = bf16_round(torch.randn(rows, vocab, generator=g) * 3.0)
    targets = torch.randint(0, vocab, (rows,), generator=g)
    targets[::7] = -100
    targets[-1] = vocab - 1
    loss_ref, dz_ref = ref.softmax_xent_ref(z, targets)
    zz = torch.full((rows, ld), 3.0)
    zz[:, :vocab] = z
    zd = zz.to(DEV, torch.bfloat16)
    loss = torch.zeros(1, device=DEV)
    ws = ops.softmax_xent_workspace(rows, DEV)
    for _ in range(2):      # the scratch words are reusable
        zd.copy_(zz)
        ops.softmax_xent_fwd_bwd(zd, targets.to(DEV), loss, ws, vocab=vocab)
        assert abs(float(loss.cpu()) - float(loss_ref)) <= 1e-5 * max(1.0, abs(float(loss_ref)))
        torch.testing.assert_close(zd.cpu().float()[:, :vocab], dz_ref, rtol=2 ** -7, atol=1e-7)


def test_shift_targets(ops):
    labels = torch.arange(24).view(3, 8)
    out = torch.empty(3, 8, dtype=torch.int64, device=DEV)
    ops.shift_targets(labels.to(DEV), out)
    exp = torch.full((3, 8), -100)
    exp[:, :-1] = labels[:, 1:]
    assert torch.equal(out.cpu(), exp)


# ----------------------------------------------------------------------------- LayerNorm
@pytest.mark.parametrize("rows,d", [(5, 64), (37, 768), (9, 2048), (3, 8), (4, 4096)])
def test_layernorm_fwd_bwd(ops, rows, d):
    g = torch.Generator().manual_seed(d)
    x = bf16_round(torch.randn(rows, d, generator=g) * 2 + 0.3)
    gamma = 1 + 0.1 * torch.randn(d, generator=g)
    beta = 0.1 * torch.randn(d, generator=g)
    dy = bf16_round(torch.randn(rows, d, generator=g))
    dres = bf16_round(torch.randn(rows, d, generator=g))
    xr = x.clone().requires_grad_(True)
    y_ref = torch.nn.functional.layer_norm(xr, (d,), gamma, beta, 1e-5)
    y_ref.backward(dy)
    xd = x.to(DEV, torch.bfloat16)
    y = torch.empty_like(xd)
    mean = torch.empty(rows, device=DEV)
    rstd = torch.empty(rows, device=DEV)
    ops.layernorm_fwd(xd, gamma.to(DEV), beta.to(DEV), y, mean, rstd, 1e-5)
    torch.testing.assert_close(y.cpu().float(), y_ref.detach(), rtol=2 ** -7, atol=2e-3)
    torch.testing.assert_close(mean.cpu(), x.mean(1), rtol=1e-5, atol=1e-5)
    dx = torch.empty_like(xd)
    ops.layernorm_bwd_dx(dy.to(DEV, torch.bfloat16), xd, gamma.to(DEV), mean, rstd, dx)
    torch.testing.assert_close(dx.cpu().float(), xr.grad, rtol=2 ** -6, atol=4e-3)
    ops.layernorm_bwd_dx(dy.to(DEV, torch.bfloat16), xd, gamma.to(DEV), mean, rstd, dx,
                         dres=dres.to(DEV, torch.bfloat16))
    torch.testing.assert_close(dx.cpu().float(), xr.grad + dres, rtol=2 ** -6, atol=8e-3)


# ----------------------------------------------------------------------------- small helpers
def test_embed_transpose_cast(ops):
    g = torch.Generator().manual_seed(3)
    V, ctx, d, B, T = 503, 64, 64, 2, 32
    wte = bf16_round(torch.randn(V, d, generator=g))
    wpe = bf16_round(torch.randn(ctx, d, generator=g))
    ids = torch.randint(0, V, (B, T), generator=g)
    x = torch.empty(B * T, d, dtype=torch.bfloat16, device=DEV)
    ops.embed_fwd(ids.to(DEV).view(-1), wte.to(DEV, torch.bfloat16), wpe.to(DEV, torch.bfloat16), x, seq=T, vocab=V)
    exp = wte[ids.view(-1)] + wpe[torch.arange(T).repeat(B)]
    torch.testing.assert_close(x.cpu().float(), exp, rtol=2 ** -8, atol=1e-6)

    a = torch.randn(37, 70, generator=g).to(torch.bfloat16)
    out = torch.zeros(70, 40, dtype=torch.bfloat16, device=DEV)
    ops.transpose(a.to(DEV), out, rows=37, cols=70)
    assert torch.equal(out.cpu()[:, :37], a.t())

    f = torch.randn(1003, generator=g)
    o = torch.empty(1003, dtype=torch.bfloat16, device=DEV)
    ops.cast_f32_to_bf16(f.to(DEV), o)
    assert torch.equal(o.cpu(), f.to(torch.bfloat16))


@pytest.mark.parametrize("M,r,cols", [(300, 16, 192), (1000, 8, 768), (130, 16, 2304), (4096, 32, 4096), (700, 32, 2048)])
def test_lora_grad(ops, M, r, cols):
    g = torch.Generator().manual_seed(M)
    S = bf16_round(torch.randn(M, max(16, r), generator=g))
    T = bf16_round(torch.randn(M, cols, generator=g))
    out = torch.zeros(r, cols, device=DEV)
    ops.lora_grad(S.to(DEV, torch.bfloat16), T.to(DEV, torch.bfloat16), out, r=r, cols=cols, alpha=2.0)
    exp = 2.0 * (S[:, :r].double().t() @ T.double()).float()
    torch.testing.assert_close(out.cpu(), exp, rtol=1e-4, atol=1e-3)
    ops.lora_grad(S.to(DEV, torch.bfloat16), T.to(DEV, torch.bfloat16), out, r=r, cols=cols, alpha=2.0)
    torch.testing.assert_close(out.cpu(), 2 * exp, rtol=1e-4, atol=2e-3)  # accumulates


# ----------------------------------------------------------------------------- tcgen05 GEMM
def _gemm_case(ops, M, N, K, a_mn=False, b_mn=False, K2=0, epilogue=0, bias=False, alpha=1.0, block_n=0, seed=0,
               cta_group=0):
    g = torch.Generator().manual_seed(seed + M + 3 * N + 7 * K)
    A = bf16_round(torch.randn(M, K, generator=g))
    Bm = bf16_round(torch.randn(N, K, generator=g))
    acc = A.double() @ Bm.double().t()
    Ad = (A.t().contiguous() if a_mn else A).to(DEV, torch.bfloat16)
    Bd = (Bm.t().contiguous() if b_mn else Bm).to(DEV, torch.bfloat16)
    kw = {}
    if K2:
        A2 = bf16_round(torch.randn(M, K2, generator=g))
        B2 = bf16_round(torch.randn(N, K2, generator=g))
        acc = acc + A2.double() @ B2.double().t()
        kw.update(a2=A2.to(DEV, torch.bfloat16), b2=B2.to(DEV, torch.bfloat16))
    v = alpha * acc
    if bias:
        bvec = torch.randn(N, generator=g)
        v = v + bvec.double()
        kw["bias"] = bvec.to(DEV)
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    out2 = None
    if epilogue == 1:
        aux = bf16_round(torch.randn(M, N, generator=g))
        exp = v + aux.double()
        kw["aux"] = aux.to(DEV, torch.bfloat16)
    elif epilogue == 2:
        exp = v
        out2 = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
        kw["out2"] = out2
    elif epilogue == 3:
        aux = bf16_round(torch.randn(M, N, generator=g))
        xa = aux.clone().requires_grad_(True)
        ref.gelu_new(xa).sum().backward()
        exp = v * xa.grad.double()
        kw["aux"] = aux.to(DEV, torch.bfloat16)
    else:
        exp = v
    ops.gemm(Ad, Bd, out, a_mn=a_mn, b_mn=b_mn, epilogue=epilogue, alpha=alpha, block_n=block_n,
             cta_group=cta_group, **kw)
    torch.cuda.synchronize()
    got = out.cpu().double()
    assert not torch.isnan(got).any(), "GEMM left output elements unwritten"
    scale = math.sqrt(K + K2)
    err = (got - exp).abs().max().item()
    assert err <= 1e-2 * scale + 2 ** -7 * exp.abs().max().item(), f"max abs err {err} (scale {scale})"
    if out2 is not None:
        exp2 = ref.gelu_new(bf16_round(exp.float())).double()
        err2 = (out2.cpu().double() - exp2).abs().max().item()
        assert err2 <= 2e-2 + 2 ** -7 * exp2.abs().max().item(), f"gelu output err {err2}"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (128, 256, 768), (256, 512, 128), (300, 200, 136),
                                   (64, 192, 64), (1000, 768, 3072), (129, 16, 768), (2048, 2304, 768)])
def test_gemm_kmajor(ops, M, N, K):
    _gemm_case(ops, M, N, K)


@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 512, 256), (200, 136, 328), (128, 16, 192)])
def test_gemm_mn_major(ops, M, N, K, a_mn, b_mn):
    _gemm_case(ops, M, N, K, a_mn=a_mn, b_mn=b_mn)


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("a_mn,b_mn", [(False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (512, 384, 520), (304, 200, 136), (2304, 768, 1000),
                                   (1024, 3072, 4096), (136, 128, 72)])
def test_gemm_mn_major_cta_groups(ops, M, N, K, a_mn, b_mn, cta_group):
    """The weight-gradient (both operands MN-major) and dX-without-transposed-copy (B MN-major) forms
    of the full-parameter step, on the single-CTA and the CTA-pair instantiation."""
    _gemm_case(ops, M, N, K, a_mn=a_mn, b_mn=b_mn, cta_group=cta_group)


@pytest.mark.parametrize("block_n", [128, 256])
def test_gemm_mn_major_pair_block_n_and_epilogue(ops, block_n):
    _gemm_case(ops, 512, 512, 256, a_mn=True, b_mn=True, cta_group=2, block_n=block_n)
    _gemm_case(ops, 512, 512, 256, b_mn=True, cta_group=2, block_n=block_n, epilogue=1)


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (256, 256, 768), (512, 512, 128), (300, 200, 136), (129, 256, 768),
                                   (1000, 768, 3072), (2048, 2304, 768), (4096, 768, 256), (200, 128, 72)])
def test_gemm_cta_groups(ops, M, N, K, cta_group):
    # cta_group 2 = CTA pairs on 256-row tiles (ragged M: second CTA's rows partly / wholly out of range)
    _gemm_case(ops, M, N, K, cta_group=cta_group)


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("epilogue", [0, 1, 2, 3])
def test_gemm_epilogues_bias_alpha(ops, epilogue, cta_group):
    _gemm_case(ops, 384, 512, 192, epilogue=epilogue, bias=True, alpha=0.5, cta_group=cta_group)


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("K2", [16, 8, 64, 80])
def test_gemm_second_segment(ops, K2, cta_group):
    # the LoRA slab: a rank-K2 product accumulated into the same TMEM tile
    _gemm_case(ops, 512, 768, 256, K2=K2, bias=True, cta_group=cta_group)


@pytest.mark.parametrize("cta_group", [1, 2])
def test_gemm_four_segments_with_strided_views(ops, cta_group):
    # backward of the fused qkv projection: dq.Wq^T + dk.Wk^T + dv.Wv^T + dU.A^T, the B operands
    # being column blocks (views) of one [d, 3d] weight
    g = torch.Generator().manual_seed(5)
    M, d, r = 640, 256, 16
    W = bf16_round(torch.randn(d, 3 * d, generator=g) * 0.1)
    As = [bf16_round(torch.randn(M, d, generator=g)) for _ in range(3)]
    dU = bf16_round(torch.randn(M, r, generator=g))
    Ad = bf16_round(torch.randn(d, 48, generator=g))      # adapter block lives inside a wider matrix
    exp = sum(As[i].double() @ W[:, i * d:(i + 1) * d].double().t() for i in range(3)) + dU.double() @ Ad[:, 16:32].double().t()
    Wd = W.to(DEV, torch.bfloat16)
    Add = Ad.to(DEV, torch.bfloat16)
    Asd = [a.to(DEV, torch.bfloat16) for a in As]
    out = torch.full((M, d), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.gemm(Asd[0], Wd[:, :d], out, a2=Asd[1], b2=Wd[:, d:2 * d],
             more=[(Asd[2], Wd[:, 2 * d:]), (dU.to(DEV, torch.bfloat16), Add[:, 16:32])], cta_group=cta_group)
    got = out.cpu().double()
    assert not torch.isnan(got).any()
    assert (got - exp).abs().max().item() <= 1e-2 * math.sqrt(3 * d) + 2 ** -7 * exp.abs().max().item()


@pytest.mark.parametrize("cta_group", [1, 2])
@pytest.mark.parametrize("block_n", [16, 64, 128, 192, 256])
def test_gemm_block_n(ops, block_n, cta_group):
    _gemm_case(ops, 640, 768, 320, block_n=block_n, cta_group=cta_group)


@pytest.mark.parametrize("cta_group", [1, 2])
def test_gemm_persistent_many_tiles(ops, cta_group):
    # more tiles than SMs so every CTA loops and both TMEM accumulator stages wrap
    _gemm_case(ops, 4096, 2304, 256, cta_group=cta_group)
    _gemm_case(ops, 8192, 3072, 768, epilogue=2, bias=True, cta_group=cta_group)


@pytest.mark.parametrize("M,N,K,split_k", [(256, 16, 4096, 0), (200, 8, 1000, 3), (768, 16, 32768, 0), (128, 64, 640, 5)])
def test_gemm_splitk_atomic_transposed(ops, M, N, K, split_k):
    # the adapter-gradient shape: both operands MN-major, long reduction, fp32 C^T accumulated
    g = torch.Generator().manual_seed(M + K)
    A = bf16_round(torch.randn(K, M, generator=g))       # [K, M]: M contiguous
    Bm = bf16_round(torch.randn(K, N, generator=g))      # [K, N]: N contiguous
    out = torch.ones(N, M, device=DEV)                   # accumulates on top of what is there
    ops.gemm(A.to(DEV, torch.bfloat16), Bm.to(DEV, torch.bfloat16), out, a_mn=True, b_mn=True, epilogue=4,
             alpha=0.5, split_k=split_k)
    exp = 1.0 + 0.5 * (Bm.double().t() @ A.double())
    torch.testing.assert_close(out.cpu().double(), exp, rtol=1e-4, atol=1e-3 * (K ** 0.5))


def test_gemm_rejects_bad_arguments(ops):
    from b200ft._lib import B200FTError
    a = torch.zeros(128, 60, dtype=torch.bfloat16, device=DEV)  # K pitch 60 is not a multiple of 8
    b = torch.zeros(128, 60, dtype=torch.bfloat16, device=DEV)
    out = torch.zeros(128, 128, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(B200FTError):
        ops.gemm(a, b, out)


def test_dependent_launch_switch_changes_nothing_but_the_launch_attribute(ops):
    """b200ft_set_dependent_launch: returns the previous setting; a chain of dependent kernels gives bit-identical
    results with programmatic dependent launch on and off (every kernel waits for its predecessor before its first
    global access)."""
    g = torch.Generator().manual_seed(11)
    x = torch.randn(4096, 768, generator=g).to(DEV, torch.bfloat16)
    w = torch.randn(768, 768, generator=g).to(DEV, torch.bfloat16) * 0.05
    gam, bet = torch.ones(768, device=DEV), torch.zeros(768, device=DEV)
    outs = []
    prev = ops.set_dependent_launch(True)
    try:
        for on in (True, False, True):
            assert ops.set_dependent_launch(on) in (True, False)
            y = torch.empty_like(x); z = torch.empty_like(x)
            mean = torch.empty(4096, device=DEV); rstd = torch.empty(4096, device=DEV)
            cur = x
            for _ in range(6):   # LayerNorm -> GEMM -> LayerNorm -> ... : each launch consumes its predecessor's output
                ops.layernorm_fwd(cur, gam, bet, y, mean, rstd, 1e-5)
                ops.gemm(y, w, z)
                cur = z
            outs.append(z.clone())
        assert ops.set_dependent_launch(False) is True
        assert ops.set_dependent_launch(True) is False
    finally:
        ops.set_dependent_launch(prev)
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    assert torch.isfinite(outs[0].float()).all()
