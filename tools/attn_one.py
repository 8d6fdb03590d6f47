import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
from b200ft import ops
B, T, H, KV, dh = 8, 4096, 12, 12, 64
qkv = torch.randn(B * T, (H + 2 * KV) * dh, device="cuda").to(torch.bfloat16)
out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device="cuda"); lse = torch.empty(B * H, T, device="cuda")
for _ in range(3):
    ops.attn_fwd(qkv, out, lse, batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
torch.cuda.synchronize(); print("ok")
