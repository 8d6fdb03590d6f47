"""Parity at the dimensions of BASELINE.json configs[2..4] that the small-model tests never touch
(VERDICT r1 item 1): Llama-3-8B geometry — d_model 4096, 32 query heads over 8 kv heads of d_head 128,
d_ff 14336, vocabulary 128256 — for the kernels and for one whole layer through both step engines, plus a
4-layer Llama-3.2-1B-width model and the loss at configs[1]'s full token count (32 x 1024).

The big GEMM expectations are fp64 products of the bf16-rounded operands evaluated with torch on the GPU
(a checker; the CPU would need minutes); everything else is checked against the CPU oracle.
Tolerances are the ones of the small tests: GEMM <= 1e-2*sqrt(K) abs, CE loss 1e-5 / gradient
1 bf16 ulp, step loss 2e-2 rel, gradient cosine >= 0.99."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import gpt2_lora_ref as gref  # noqa: E402
from oracle import llama_lora_ref as ref  # noqa: E402
from oracle.gpt2_lora_ref import AdamWConfig  # noqa: E402

DEV = "cuda:0"


@pytest.fixture(scope="module")
def ops():
    from b200ft import ops as _ops
    return _ops


def _cos(a, b):
    a, b = a.flatten().double(), b.flatten().double()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def _rel_l2(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))


# ------------------------------------------------------------------ GEMM shapes of the 8B model
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True)])
@pytest.mark.parametrize("M,N,K", [(512, 4096, 14336),      # down_proj forward / its dX
                                   (512, 28672, 4096),      # fused gate|up
                                   (512, 128256, 4096),     # lm_head
                                   (14336, 4096, 512)])     # dW of down_proj: token index is K
def test_gemm_at_llama_8b_shapes(ops, M, N, K, a_mn, b_mn):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    A = torch.randn(M, K, generator=g, device=DEV).to(torch.bfloat16)
    Bm = torch.randn(N, K, generator=g, device=DEV).to(torch.bfloat16)
    Ad = A.t().contiguous() if a_mn else A
    Bd = Bm.t().contiguous() if b_mn else Bm
    out = torch.full((M, N), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.gemm(Ad, Bd, out, a_mn=a_mn, b_mn=b_mn)
    torch.cuda.synchronize()
    assert not torch.isnan(out).any()
    worst, peak = 0.0, 0.0
    for c0 in range(0, N, 16384):                      # fp64 expectation in column blocks (memory)
        exp = A.double() @ Bm[c0: c0 + 16384].double().t()
        worst = max(worst, float((out[:, c0: c0 + 16384].double() - exp).abs().max()))
        peak = max(peak, float(exp.abs().max()))
    assert worst <= 1e-2 * math.sqrt(K) + 2 ** -7 * peak, f"max abs err {worst}"


def test_gemm_residual_epilogue_at_8b_width(ops):
    g = torch.Generator(device=DEV).manual_seed(3)
    M, N, K = 4096, 4096, 14336                         # down_proj + residual, one full micro-batch
    A = torch.randn(M, K, generator=g, device=DEV).to(torch.bfloat16)
    Bm = (torch.randn(N, K, generator=g, device=DEV) * 0.05).to(torch.bfloat16)
    aux = torch.randn(M, N, generator=g, device=DEV).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=DEV)
    ops.gemm(A, Bm, out, epilogue=1, aux=aux)
    exp = A.double() @ Bm.double().t() + aux.double()
    err = float((out.double() - exp).abs().max())
    assert err <= 1e-2 * math.sqrt(K) * 0.05 + 2 ** -7 * float(exp.abs().max()), err


# ------------------------------------------------------------------ fused CE at the Llama vocabulary
@pytest.mark.parametrize("rows,vocab,ld", [(256, 128256, 128256), (37, 128256, 128256), (300, 32000, 32000)])
def test_softmax_xent_at_llama_vocab(ops, rows, vocab, ld):
    g = torch.Generator().manual_seed(rows)
    z = (torch.randn(rows, vocab, generator=g) * 3.0).to(torch.bfloat16).float()
    targets = torch.randint(0, vocab, (rows,), generator=g)
    targets[1] = -100
    targets[2] = vocab - 1
    targets[3] = 0
    loss_ref, dz_ref = gref.softmax_xent_ref(z, targets)
    zd = z.to(DEV, torch.bfloat16)
    loss = torch.zeros(1, device=DEV)
    ws = ops.softmax_xent_workspace(rows, DEV)
    ops.softmax_xent_fwd_bwd(zd, targets.to(DEV), loss, ws, vocab=vocab)
    assert abs(float(loss.cpu()) - float(loss_ref)) <= 1e-5 * max(1.0, abs(float(loss_ref)))
    torch.testing.assert_close(zd.cpu().float(), dz_ref, rtol=2 ** -7, atol=1e-7)
    assert torch.all(zd[1] == 0)


# ------------------------------------------------------------------ one Llama-3-8B layer through both engines
CFG_8B_LAYER = ref.LlamaConfig(n_layer=1, d_model=4096, n_head=32, n_kv_head=8, d_ff=14336, vocab=4099, ctx=8192,
                               rope_theta=500000.0, rope_llama3=False, tie_embeddings=False)


def _dims(cfg):
    from b200ft import llama
    return llama.LlamaDims(n_layer=cfg.n_layer, d_model=cfg.d_model, n_head=cfg.n_head, n_kv_head=cfg.n_kv_head,
                           d_ff=cfg.d_ff, vocab=cfg.vocab, ctx=cfg.ctx, rms_eps=cfg.rms_eps, rope_theta=cfg.rope_theta,
                           rope_llama3=cfg.rope_llama3, rope_factor=cfg.rope_factor,
                           rope_low_freq_factor=cfg.rope_low_freq_factor, rope_high_freq_factor=cfg.rope_high_freq_factor,
                           rope_original_ctx=cfg.rope_original_ctx, tie_embeddings=cfg.tie_embeddings,
                           d_head_override=cfg.d_head_override)


def _opt(o):
    from b200ft import gpt2
    return gpt2.OptSpec(o.lr, o.beta1, o.beta2, o.eps, o.weight_decay, o.max_grad_norm)


def _lora_parity(cfg, batch, seq, r, steps, attention):
    from b200ft import gpt2, llama
    assert cfg.d_head == cfg.d_model // cfg.n_head
    lora, opt = ref.LoRAConfig(r=r, alpha=2.0 * r), AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    oracle = ref.OracleTrainer(cfg, lora, opt, base, ad)
    eng = llama.LlamaLoRAStep(_dims(cfg), gpt2.LoRASpec(r, 2.0 * r), _opt(opt), base, ad, batch=batch, seq=seq,
                              device=DEV, attention=attention)
    for s in range(steps):
        ids, labels = ref.synthetic_batch(cfg, batch, seq, seed=1234 + s)
        l_ref = oracle.step(ids, labels)
        l_gpu = float(eng.step(ids.cuda(), labels.cuda()).cpu())
        assert abs(l_gpu - l_ref) / abs(l_ref) <= 2e-2, f"step {s}: loss {l_gpu} vs oracle {l_ref}"
        if s == 0:
            g_gpu = eng.grads_state()
            for k, g_ref in oracle.last_grads.items():
                assert _cos(g_gpu[k], g_ref) >= 0.99, f"{k}: cos {_cos(g_gpu[k], g_ref)}"
                assert _rel_l2(g_gpu[k], g_ref) <= 5e-2, f"{k}: rel l2 {_rel_l2(g_gpu[k], g_ref)}"
    new = eng.adapters_state()
    for k, t in oracle.trainable.items():
        assert _cos(new[k] - ad[k], t.detach() - ad[k]) >= 0.99, k


def _full_parity(cfg, batch, seq, steps, attention):
    from b200ft import llama_full
    opt = AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    oracle = ref.FullOracleTrainer(cfg, opt, base)
    eng = llama_full.LlamaFullStep(_dims(cfg), _opt(opt), base, batch=batch, seq=seq, device=DEV, attention=attention)
    for s in range(steps):
        ids, labels = ref.synthetic_batch(cfg, batch, seq, seed=1234 + s)
        l_ref = oracle.step(ids, labels)
        l_gpu = float(eng.step(ids.to(DEV), labels.to(DEV)).cpu())
        assert abs(l_gpu - l_ref) / abs(l_ref) <= 2e-2, f"step {s}: loss {l_gpu} vs oracle {l_ref}"
        if s == 0:
            g_gpu = eng.grads_state()
            for k, g_ref in oracle.last_grads.items():
                assert _cos(g_gpu[k], g_ref) >= 0.99, f"{k}: cos {_cos(g_gpu[k], g_ref)}"
                assert _rel_l2(g_gpu[k], g_ref) <= 6e-2, f"{k}: rel l2 {_rel_l2(g_gpu[k], g_ref)}"
    new = eng.params_state()
    for k, t in oracle.trainable.items():
        assert _cos(new[k] - base[k], t.detach() - base[k]) >= 0.97, k


@pytest.mark.parametrize("attention", ["sdpa", "own"])
def test_llama_8b_layer_lora_step(attention):
    _lora_parity(CFG_8B_LAYER, batch=1, seq=512, r=16, steps=2, attention=attention)


@pytest.mark.parametrize("attention", ["sdpa", "own"])
def test_llama_8b_layer_full_parameter_step(attention):
    _full_parity(CFG_8B_LAYER, batch=1, seq=512, steps=1, attention=attention)


def test_llama_1b_width_four_layers():
    cfg = ref.LlamaConfig(n_layer=4, d_model=2048, n_head=32, n_kv_head=8, d_ff=8192, vocab=4099, ctx=131072)
    _lora_parity(cfg, batch=2, seq=256, r=16, steps=2, attention="own")


# ------------------------------------------------------------------ configs[1]'s token count: 32 x 1024, GPT-2 width
def test_gpt2_width_two_layers_at_c2_token_count():
    """batch 32 x seq 1024 (the benchmarked shape: 32768 rows through every kernel, real vocabulary), two layers.
    The oracle's loss is evaluated sequence by sequence without autograd (the fp32 logits of the whole batch
    would be 6.6 GB); the mean over B*(T-1) targets is the same number."""
    from b200ft import gpt2
    cfg = gref.GPT2Config(n_layer=2, d_model=768, n_head=12, d_ff=3072, vocab=50257, ctx=1024)
    lora = gref.LoRAConfig(r=8, alpha=16.0)
    base = gref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = gref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    B, T = 32, 1024
    ids, labels = gref.synthetic_batch(cfg, B, T, seed=1234)
    total = 0.0
    with torch.no_grad():
        for b0 in range(0, B, 4):
            total += float(gref.causal_lm_loss(gref.forward(base, ad, ids[b0:b0 + 4], cfg, lora), labels[b0:b0 + 4])) * 4
    l_ref = total / B
    dims = gpt2.GPT2Dims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.d_ff, cfg.vocab, cfg.ctx, cfg.ln_eps)
    for attention in ("sdpa", "own"):
        eng = gpt2.GPT2LoRAStep(dims, gpt2.LoRASpec(8, 16.0), gpt2.OptSpec(), base, ad, batch=B, seq=T, device=DEV,
                                attention=attention)
        l_gpu = float(eng.step(ids.cuda(), labels.cuda()).cpu())
        assert abs(l_gpu - l_ref) / l_ref <= 2e-2, f"{attention}: loss {l_gpu} vs oracle {l_ref}"
        g = eng.grads_state()
        assert all(torch.isfinite(t).all() and float(t.norm()) > 0 for t in g.values())
        del eng
        torch.cuda.empty_cache()
