import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from b200ft import ops
from test_attention_gpu import ref_attention
DEV = "cuda:0"
for (B, T, H, KV, dh) in [(9, 512, 4, 4, 128), (10, 512, 4, 4, 128), (40, 512, 4, 4, 128), (2, 4096, 8, 2, 128)]:
    g = torch.Generator().manual_seed(7 * T + H)
    ld = (H + 2 * KV) * dh
    qkv = (torch.randn(B * T, ld, generator=g) * 1.2).to(torch.bfloat16)
    dout = torch.randn(B * T, H * dh, generator=g).to(torch.bfloat16)
    q_off, k_off, v_off = 0, H * dh, (H + KV) * dh
    qd, dod = qkv.to(DEV), dout.to(DEV)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=DEV)
    lse2 = torch.empty(B * H, T, device=DEV)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=q_off, k_off=k_off, v_off=v_off, scale=1.0 / math.sqrt(dh))
    ops.attn_fwd(qd, out, lse2, **kw)
    dqkv = torch.full((B * T, ld), float("nan"), dtype=torch.bfloat16, device=DEV)
    ws = ops.attn_bwd_workspace(B, T, H, dh, DEV)
    ops.attn_bwd(qd, out, dod, lse2, ws, dqkv, **kw)
    torch.cuda.synchronize()
    got = dqkv.float().cpu()
    x = qkv.float().clone().requires_grad_(True)
    if B * T <= 8192:
        o_ref, _ = ref_attention(x, B, T, H, KV, dh, q_off, k_off, v_off)
        o_ref.backward(dout.float())
    for name, lo, hi in (("dq", q_off, k_off), ("dk", k_off, v_off), ("dv", v_off, ld)):
        a = got[:, lo:hi]
        nan = torch.isnan(a)
        msg = f"{(B,T,H,KV)} {name}: nan {int(nan.sum())}"
        if nan.any():
            rows = nan.any(1).nonzero().flatten()
            msg += f" rows {rows[:6].tolist()}..{rows[-3:].tolist()} (n={len(rows)}) tiles {sorted(set((rows // 128).tolist()))[:20]}"
        elif x.grad is not None:
            r = x.grad[:, lo:hi]
            msg += f" rel {float((a - r).norm() / r.norm()):.4f}"
        print(msg)
