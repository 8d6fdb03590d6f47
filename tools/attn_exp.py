"""Bottleneck experiments on the attention backward: times variant builds of the library
(tools/_exp/lib_<VARIANT>.so, compiled with -DB200FT_EXP_<VARIANT>; results are numerically wrong by design)."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
from b200ft import _lib
if len(sys.argv) > 1 and sys.argv[1] != "base":
    _lib.LIB_PATH = os.path.join(ROOT, "tools", "_exp", f"lib_{sys.argv[1]}.so")
from b200ft import ops
for (B, T, H, KV) in [(32, 1024, 12, 12), (8, 4096, 12, 12)]:
    dh = 64; ld = (H + 2 * KV) * dh
    qkv = torch.randn(B * T, ld, device="cuda").to(torch.bfloat16)
    dout = torch.randn(B * T, H * dh, device="cuda").to(torch.bfloat16)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B * H, T, device="cuda")
    ws = ops.attn_bwd_workspace(B, T, H, dh, "cuda"); dqkv = torch.empty(B * T, ld, dtype=torch.bfloat16, device="cuda")
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
    ops.attn_fwd(qkv, out, lse, **kw)
    for det in (False, True):
        for _ in range(3): ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, deterministic=det, **kw)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(10): ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, deterministic=det, **kw)
        e1.record(); torch.cuda.synchronize()
        print(f"{sys.argv[1] if len(sys.argv) > 1 else 'base'} B={B} T={T} {'two-kernel' if det else 'fused'}: {e0.elapsed_time(e1) * 100:.1f} us", flush=True)
