"""A couple of launches of the own attention forward and backward at one shape: the short command ncu profiles."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
from b200ft import ops
B, T, H, KV = (int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "32,1024,12,12").split(","))
dh = 64
ld = (H + 2 * KV) * dh
qkv = torch.randn(B * T, ld, device="cuda").to(torch.bfloat16)
dout = torch.randn(B * T, H * dh, device="cuda").to(torch.bfloat16)
out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B * H, T, device="cuda")
ws = ops.attn_bwd_workspace(B, T, H, dh, "cuda"); dqkv = torch.empty(B * T, ld, dtype=torch.bfloat16, device="cuda")
kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
for _ in range(2):
    ops.attn_fwd(qkv, out, lse, **kw)
    ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw)
torch.cuda.synchronize(); print("ok")
