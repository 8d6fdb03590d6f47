"""Bottleneck experiments on the attention forward (see tools/attn_exp.py): variant builds in tools/_exp/."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
from b200ft import _lib
name = sys.argv[1] if len(sys.argv) > 1 else "base"
if name != "base":
    _lib.LIB_PATH = os.path.join(ROOT, "tools", "_exp", f"lib_{name}.so")
from b200ft import ops
for (B, T, H, KV) in [(32, 1024, 12, 12), (8, 4096, 12, 12)]:
    dh = 64; ld = (H + 2 * KV) * dh
    qkv = torch.randn(B * T, ld, device="cuda").to(torch.bfloat16)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B * H, T, device="cuda")
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
    for _ in range(3): ops.attn_fwd(qkv, out, lse, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(20): ops.attn_fwd(qkv, out, lse, **kw)
    e1.record(); torch.cuda.synchronize()
    print(f"{name} B={B} T={T} fwd: {e0.elapsed_time(e1) * 50:.1f} us", flush=True)
