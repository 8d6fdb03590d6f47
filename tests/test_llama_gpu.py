"""Llama-family kernels and the full Llama LoRA step against the CPU oracle
(oracle/llama_lora_ref.py).  Same tolerances as the GPT-2 tests."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import llama_lora_ref as ref  # noqa: E402
from oracle.gpt2_lora_ref import AdamWConfig  # noqa: E402

DEV = "cuda:0"


def bf16_round(t):
    return t.to(torch.bfloat16).to(torch.float32)


@pytest.fixture(scope="module")
def ops():
    from b200ft import ops as _ops
    return _ops


@pytest.mark.parametrize("rows,d", [(5, 64), (37, 2048), (3, 4096), (9, 8)])
def test_rmsnorm_fwd_bwd(ops, rows, d):
    g = torch.Generator().manual_seed(d)
    x = bf16_round(torch.randn(rows, d, generator=g) * 2 + 0.3)
    gamma = 1 + 0.1 * torch.randn(d, generator=g)
    dy, dres = bf16_round(torch.randn(rows, d, generator=g)), bf16_round(torch.randn(rows, d, generator=g))
    xr = x.clone().requires_grad_(True)
    y_ref = ref.rmsnorm(xr, gamma, 1e-5)
    y_ref.backward(dy)
    xd = x.to(DEV, torch.bfloat16)
    y, rstd = torch.empty_like(xd), torch.empty(rows, device=DEV)
    ops.rmsnorm_fwd(xd, gamma.to(DEV), y, rstd, 1e-5)
    torch.testing.assert_close(y.cpu().float(), y_ref.detach(), rtol=2 ** -7, atol=2e-3)
    torch.testing.assert_close(rstd.cpu(), torch.rsqrt(x.pow(2).mean(1) + 1e-5), rtol=1e-5, atol=1e-6)
    dx = torch.empty_like(xd)
    ops.rmsnorm_bwd_dx(dy.to(DEV, torch.bfloat16), xd, gamma.to(DEV), rstd, dx)
    torch.testing.assert_close(dx.cpu().float(), xr.grad, rtol=2 ** -6, atol=4e-3)
    ops.rmsnorm_bwd_dx(dy.to(DEV, torch.bfloat16), xd, gamma.to(DEV), rstd, dx, dres=dres.to(DEV, torch.bfloat16))
    torch.testing.assert_close(dx.cpu().float(), xr.grad + dres, rtol=2 ** -6, atol=8e-3)


@pytest.mark.parametrize("dh,heads,extra_cols", [(64, 5, 64), (16, 6, 0), (128, 2, 256)])
def test_rope_forward_and_inverse(ops, dh, heads, extra_cols):
    g = torch.Generator().manual_seed(dh)
    B, T = 2, 24
    cfg = ref.LlamaConfig(d_model=dh * 4, n_head=4, rope_theta=10000.0, rope_llama3=False, d_head_override=dh)
    cos, sin = ref.rope_tables(cfg, T)
    ld = heads * dh + extra_cols
    x = bf16_round(torch.randn(B * T, ld, generator=g))
    xh = x[:, : heads * dh].view(B, T, heads, dh).transpose(1, 2)
    exp = ref.apply_rope(xh, cos, sin).transpose(1, 2).reshape(B * T, heads * dh)
    xd = x.to(DEV, torch.bfloat16)
    ops.rope_inplace(xd, cos.to(DEV), sin.to(DEV), seq=T, n_heads=heads, d_head=dh)
    got = xd.cpu().float()
    torch.testing.assert_close(got[:, : heads * dh], exp, rtol=2 ** -7, atol=2e-2)
    assert torch.equal(got[:, heads * dh:], x[:, heads * dh:])          # columns past the heads untouched
    ops.rope_inplace(xd, cos.to(DEV), sin.to(DEV), seq=T, n_heads=heads, d_head=dh, backward=True)
    torch.testing.assert_close(xd.cpu().float(), x, rtol=2 ** -6, atol=4e-2)  # inverse rotation (two bf16 roundings)


@pytest.mark.parametrize("rows,ff", [(7, 128), (33, 8192)])
def test_swiglu_fwd_bwd(ops, rows, ff):
    g = torch.Generator().manual_seed(ff)
    gu = bf16_round(torch.randn(rows, 2 * ff, generator=g) * 2)
    dact = bf16_round(torch.randn(rows, ff, generator=g))
    gur = gu.clone().requires_grad_(True)
    act_ref = torch.nn.functional.silu(gur[:, :ff]) * gur[:, ff:]
    act_ref.backward(dact)
    gud = gu.to(DEV, torch.bfloat16)
    act = torch.empty(rows, ff, dtype=torch.bfloat16, device=DEV)
    ops.swiglu_fwd(gud, act)
    torch.testing.assert_close(act.cpu().float(), act_ref.detach(), rtol=2 ** -7, atol=1e-3)
    dgu = torch.empty_like(gud)
    ops.swiglu_bwd(dact.to(DEV, torch.bfloat16), gud, dgu)
    torch.testing.assert_close(dgu.cpu().float(), gur.grad, rtol=2 ** -6, atol=2e-3)
    # the recompute form used by the full-parameter step: same gradient, plus the forward product again
    dgu2, act2 = torch.empty_like(gud), torch.empty_like(act)
    ops.swiglu_bwd(dact.to(DEV, torch.bfloat16), gud, dgu2, act=act2)
    assert torch.equal(dgu2, dgu) and torch.equal(act2, act)


def _engine(cfg, lora, opt, base, ad, batch, seq, **kw):
    from b200ft import gpt2, llama
    dims = llama.LlamaDims(cfg.n_layer, cfg.d_model, cfg.n_head, cfg.n_kv_head, cfg.d_ff, cfg.vocab, cfg.ctx,
                           cfg.rms_eps, cfg.rope_theta, cfg.rope_llama3, cfg.rope_factor, cfg.rope_low_freq_factor,
                           cfg.rope_high_freq_factor, cfg.rope_original_ctx, cfg.tie_embeddings, cfg.d_head_override)
    return llama.LlamaLoRAStep(dims, gpt2.LoRASpec(lora.r, lora.alpha),
                               gpt2.OptSpec(opt.lr, opt.beta1, opt.beta2, opt.eps, opt.weight_decay, opt.max_grad_norm),
                               base, ad, batch=batch, seq=seq, device=DEV, **kw)


def _cos(a, b):
    a, b = a.flatten().double(), b.flatten().double()
    return float((a @ b) / (a.norm() * b.norm() + 1e-30))


def _parity(cfg, batch, seq, r, steps, **kw):
    lora = ref.LoRAConfig(r=r, alpha=2.0 * r)
    opt = AdamWConfig(lr=1e-3)
    base = ref.init_base_params(cfg, seed=0, randomize_all=True)
    ad = ref.init_lora_params(cfg, lora, seed=1, b_init="normal")
    oracle = ref.OracleTrainer(cfg, lora, opt, base, ad)
    eng = _engine(cfg, lora, opt, base, ad, batch, seq, **kw)
    for s in range(steps):
        ids, labels = ref.synthetic_batch(cfg, batch, seq, seed=1234 + s)
        l_ref = oracle.step(ids, labels)
        l_gpu = float(eng.step(ids.cuda(), labels.cuda()).cpu())
        assert abs(l_gpu - l_ref) / abs(l_ref) <= 2e-2, f"step {s}: loss {l_gpu} vs oracle {l_ref}"
        if s == 0:
            g_gpu = eng.grads_state()
            for k, g_ref in oracle.last_grads.items():
                assert _cos(g_gpu[k], g_ref) >= 0.99, f"{k}: cos {_cos(g_gpu[k], g_ref)}"
                rel = float((g_gpu[k].double() - g_ref.double()).norm() / g_ref.double().norm())
                assert rel <= 5e-2, f"{k}: rel l2 {rel}"
    new = eng.adapters_state()
    for k, t in oracle.trainable.items():
        assert _cos(new[k] - ad[k], t.detach() - ad[k]) >= 0.99, k


def test_llama_tiny_three_steps():
    _parity(ref.LLAMA_TINY, batch=2, seq=32, r=16, steps=3)


def test_llama_tiny_rank8_ragged():
    _parity(ref.LLAMA_TINY, batch=3, seq=40, r=8, steps=2)


def test_llama_one_layer_real_width():
    # Llama-3.2-1B width / heads / GQA ratio / llama3 rope, one layer, small vocab to keep the CPU oracle quick
    cfg = ref.LlamaConfig(n_layer=1, d_model=2048, n_head=32, n_kv_head=8, d_ff=8192, vocab=4099, ctx=131072)
    _parity(cfg, batch=1, seq=256, r=16, steps=2)


def test_llama_real_width_with_own_attention_kernels():
    """attention="own" at Llama-3.2-1B's head geometry (32 q heads over 8 kv heads of 64): the step with no
    library kernel in it."""
    cfg = ref.LlamaConfig(n_layer=1, d_model=2048, n_head=32, n_kv_head=8, d_ff=8192, vocab=4099, ctx=131072)
    _parity(cfg, batch=1, seq=256, r=16, steps=2, attention="own")


def test_own_attention_rejects_other_head_sizes():
    with pytest.raises(ValueError):
        _parity(ref.LLAMA_TINY, batch=2, seq=32, r=16, steps=1, attention="own")   # d_head 16: neither 64 nor 128
