"""Own causal flash-attention (tcgen05) against the oracle's explicit masked-softmax attention."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def ref_attention(qkv, B, T, H, KV, dh, q_off, k_off, v_off):
    """fp32, explicit: what oracle/gpt2_lora_ref.py:causal_attention and llama forward() compute."""
    x = qkv.float().view(B, T, -1)
    q = x[..., q_off:q_off + H * dh].view(B, T, H, dh).transpose(1, 2)
    k = x[..., k_off:k_off + KV * dh].view(B, T, KV, dh).transpose(1, 2).repeat_interleave(H // KV, dim=1)
    v = x[..., v_off:v_off + KV * dh].view(B, T, KV, dh).transpose(1, 2).repeat_interleave(H // KV, dim=1)
    s = (q @ k.transpose(-1, -2)) / math.sqrt(dh)
    mask = torch.ones(T, T, dtype=torch.bool).tril()
    s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)                                  # [B, H, T]
    o = torch.softmax(s, dim=-1) @ v
    return o.transpose(1, 2).reshape(B * T, H * dh), lse


@pytest.mark.parametrize("dh", [64, 128])
@pytest.mark.parametrize("B,T,H,KV", [(1, 128, 1, 1), (2, 256, 3, 3), (1, 1024, 2, 2), (2, 384, 8, 2), (3, 200, 4, 1),
                                      (1, 40, 2, 2), (3, 2048, 8, 2)])
def test_attn_fwd_matches_oracle(B, T, H, KV, dh):
    from b200ft import ops
    from oracle import gpt2_lora_ref as ref
    g = torch.Generator().manual_seed(T + H)
    ld = (H + 2 * KV) * dh
    qkv = (torch.randn(B * T, ld, generator=g) * 1.5).to(torch.bfloat16)
    q_off, k_off, v_off = 0, H * dh, (H + KV) * dh
    o_ref, lse_ref = ref_attention(qkv, B, T, H, KV, dh, q_off, k_off, v_off)
    if H == KV and dh == 64:  # the GPT-2 oracle's own function on the same buffer
        o2 = ref.causal_attention(qkv.float().view(B, T, ld), H).reshape(B * T, H * dh)
        torch.testing.assert_close(o2, o_ref, rtol=1e-5, atol=1e-5)
    qd = qkv.to(DEV)
    out = torch.full((B * T, H * dh), float("nan"), dtype=torch.bfloat16, device=DEV)
    lse2 = torch.full((B * H, T), float("nan"), device=DEV)
    ops.attn_fwd(qd, out, lse2, batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=q_off, k_off=k_off,
                 v_off=v_off, scale=1.0 / math.sqrt(dh))
    torch.cuda.synchronize()
    got = out.cpu().float()
    assert not torch.isnan(got).any()
    torch.testing.assert_close(got, o_ref, rtol=2e-2, atol=2e-2)          # bf16 P and bf16 output
    lse_got = lse2.cpu().view(B, H, T) * math.log(2.0)                    # log2 domain -> natural log
    torch.testing.assert_close(lse_got, lse_ref, rtol=1e-3, atol=2e-3)


def test_attn_fwd_rejects_other_head_sizes():
    from b200ft import ops
    from b200ft._lib import B200FTError
    x = torch.zeros(128, 3 * 96, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(B200FTError):
        ops.attn_fwd(x, torch.zeros(128, 96, dtype=torch.bfloat16, device=DEV), torch.zeros(1, 128, device=DEV),
                     batch=1, seq=128, n_head=1, n_kv_head=1, d_head=96, q_off=0, k_off=96, v_off=192, scale=0.1)


@pytest.mark.parametrize("dh,deterministic", [(64, False), (64, True), (128, False)])
@pytest.mark.parametrize("B,T,H,KV", [(1, 128, 1, 1), (2, 256, 3, 3), (1, 1024, 2, 2), (2, 384, 8, 2), (3, 200, 4, 1),
                                      (1, 40, 2, 2), (40, 512, 4, 4)])
def test_attn_bwd_matches_oracle_autograd(B, T, H, KV, dh, deterministic):
    """dq | dk | dv of the own backward (fed by the own forward's out and log-sum-exp) against autograd
    through the explicit fp32 attention; GQA sums the group's heads into dk / dv; ragged lengths.  d_head 128
    has one form (two kernels, no atomics); d_head 64 the fused default and the two-kernel deterministic one."""
    from b200ft import ops
    g = torch.Generator().manual_seed(7 * T + H)
    ld = (H + 2 * KV) * dh
    qkv = (torch.randn(B * T, ld, generator=g) * 1.2).to(torch.bfloat16)
    dout = torch.randn(B * T, H * dh, generator=g).to(torch.bfloat16)
    q_off, k_off, v_off = 0, H * dh, (H + KV) * dh
    x = qkv.float().clone().requires_grad_(True)
    o_ref, _ = ref_attention(x, B, T, H, KV, dh, q_off, k_off, v_off)
    o_ref.backward(dout.float())
    qd, dod = qkv.to(DEV), dout.to(DEV)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=DEV)
    lse2 = torch.empty(B * H, T, device=DEV)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=q_off, k_off=k_off, v_off=v_off,
              scale=1.0 / math.sqrt(dh))
    ops.attn_fwd(qd, out, lse2, **kw)
    dqkv = torch.full((B * T, ld), float("nan"), dtype=torch.bfloat16, device=DEV)
    ws = ops.attn_bwd_workspace(B, T, H, dh, DEV)
    ops.attn_bwd(qd, out, dod, lse2, ws, dqkv, deterministic=deterministic, **kw)
    torch.cuda.synchronize()
    got = dqkv.cpu().float()
    assert not torch.isnan(got).any(), "backward left gradient elements unwritten"
    ref_g = x.grad
    for name, lo, hi in (("dq", q_off, k_off), ("dk", k_off, v_off), ("dv", v_off, ld)):
        a, r = got[:, lo:hi], ref_g[:, lo:hi]
        cos = float((a.double().flatten() @ r.double().flatten()) / (a.double().norm() * r.double().norm()))
        rel = float((a - r).norm() / r.norm())
        assert cos >= 0.999 and rel <= 3e-2, f"{name}: cos {cos} rel l2 {rel}"
        torch.testing.assert_close(a, r, rtol=5e-2, atol=3e-2 * float(r.abs().max()))


def test_attn_bwd_d128_is_bit_reproducible_and_handles_long_gqa():
    """Llama-3-8B's head geometry (4 q heads per kv head, d_head 128) at T = 2048: two runs give identical bits, and
    the result matches autograd of the explicit attention."""
    from b200ft import ops
    B, T, H, KV, dh = 1, 2048, 8, 2, 128
    g = torch.Generator().manual_seed(11)
    ld = (H + 2 * KV) * dh
    qkv = (torch.randn(B * T, ld, generator=g) * 1.2).to(torch.bfloat16)
    dout = torch.randn(B * T, H * dh, generator=g).to(torch.bfloat16)
    q_off, k_off, v_off = 0, H * dh, (H + KV) * dh
    x = qkv.float().clone().requires_grad_(True)
    o_ref, _ = ref_attention(x, B, T, H, KV, dh, q_off, k_off, v_off)
    o_ref.backward(dout.float())
    qd, dod = qkv.to(DEV), dout.to(DEV)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=q_off, k_off=k_off, v_off=v_off,
              scale=1.0 / math.sqrt(dh))
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=DEV)
    lse2, ws = torch.empty(B * H, T, device=DEV), ops.attn_bwd_workspace(B, T, H, dh, DEV)
    ops.attn_fwd(qd, out, lse2, **kw)
    res = []
    for _ in range(2):
        dqkv = torch.full((B * T, ld), float("nan"), dtype=torch.bfloat16, device=DEV)
        ops.attn_bwd(qd, out, dod, lse2, ws, dqkv, **kw)
        res.append(dqkv.clone())
    assert torch.equal(res[0], res[1])
    got, ref_g = res[0].cpu().float(), x.grad
    for name, lo, hi in (("dq", q_off, k_off), ("dk", k_off, v_off), ("dv", v_off, ld)):
        a, r = got[:, lo:hi], ref_g[:, lo:hi]
        cos = float((a.double().flatten() @ r.double().flatten()) / (a.double().norm() * r.double().norm()))
        assert cos >= 0.999 and float((a - r).norm() / r.norm()) <= 3e-2, name


def test_attn_bwd_deterministic_path_is_bit_reproducible():
    from b200ft import ops
    B, T, H, KV, dh = 2, 384, 4, 2, 64
    g = torch.Generator().manual_seed(1)
    ld = (H + 2 * KV) * dh
    qd = torch.randn(B * T, ld, generator=g).to(torch.bfloat16).to(DEV)
    dod = torch.randn(B * T, H * dh, generator=g).to(torch.bfloat16).to(DEV)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh,
              scale=0.125)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=DEV)
    lse2, ws = torch.empty(B * H, T, device=DEV), ops.attn_bwd_workspace(B, T, H, dh, DEV)
    ops.attn_fwd(qd, out, lse2, **kw)
    res = []
    for _ in range(2):
        dqkv = torch.zeros(B * T, ld, dtype=torch.bfloat16, device=DEV)
        ops.attn_bwd(qd, out, dod, lse2, ws, dqkv, deterministic=True, **kw)
        res.append(dqkv.clone())
    assert torch.equal(res[0], res[1])
    # the fused default: dk, dv are exact too; dq is summed from bf16-rounded partials (one per visible kv tile, here
    # up to 3), in an order that varies from run to run
    fused = torch.full((B * T, ld), float("nan"), dtype=torch.bfloat16, device=DEV)
    ops.attn_bwd(qd, out, dod, lse2, ws, fused, **kw)
    assert torch.equal(fused[:, H * dh:], res[0][:, H * dh:])
    a, r = fused[:, : H * dh].float(), res[0][:, : H * dh].float()
    assert float((a - r).norm() / r.norm()) <= 2 ** -7
    torch.testing.assert_close(a, r, rtol=2 ** -5, atol=2e-3 * float(r.abs().max()))


@pytest.mark.parametrize("dh", [64, 128])
def test_attn_fwd_rescales_output_accumulator_when_the_row_max_grows(dh):
    """Keys whose magnitude rises tile after tile force the lazy-rescaling path on every kv tile: the reference max
    moves by far more than 2^8, so the output tile accumulated in TMEM has to be rescaled in place."""
    from b200ft import ops
    B, T, H, KV = 1, 640, 2, 2
    g = torch.Generator().manual_seed(3)
    ld = (H + 2 * KV) * dh
    x = torch.randn(B * T, ld, generator=g)
    ramp = (1.0 + 3.0 * (torch.arange(T) // 128).float())[:, None]       # x1, x4, x7, ... per 128-key tile
    x[:, H * dh: (H + KV) * dh] *= ramp
    qkv = x.to(torch.bfloat16)
    q_off, k_off, v_off = 0, H * dh, (H + KV) * dh
    o_ref, lse_ref = ref_attention(qkv, B, T, H, KV, dh, q_off, k_off, v_off)
    out = torch.full((B * T, H * dh), float("nan"), dtype=torch.bfloat16, device=DEV)
    lse2 = torch.empty(B * H, T, device=DEV)
    ops.attn_fwd(qkv.to(DEV), out, lse2, batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=q_off, k_off=k_off,
                 v_off=v_off, scale=1.0 / math.sqrt(dh))
    got = out.cpu().float()
    assert not torch.isnan(got).any()
    torch.testing.assert_close(got, o_ref, rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(lse2.cpu().view(B, H, T) * math.log(2.0), lse_ref, rtol=1e-3, atol=5e-3)
