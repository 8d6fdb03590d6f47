"""Own attention forward (tcgen05) vs torch SDPA (cuDNN) forward, CUDA-event timed."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch, torch.nn.functional as F
from b200ft import ops, _lib
if os.environ.get("B200FT_PROBE_LIB"): _lib.LIB_PATH = os.path.join(ROOT, os.environ["B200FT_PROBE_LIB"])  # variant build, measurement only
dev = "cuda:0"
def timeit(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n
import ast
# shapes: (batch, seq, heads, kv heads[, d_head = 64]); defaults = C2, C3, a long d_head-64 case, C4/C5 (Llama-3-8B)
SHAPES = ast.literal_eval(sys.argv[1]) if len(sys.argv) > 1 else [(32, 1024, 12, 12), (4, 2048, 32, 8), (8, 4096, 12, 12),
                                                                  (1, 4096, 32, 8, 128), (2, 4096, 32, 8, 128)]
for shape in SHAPES:
    B, T, H, KV = shape[:4]
    dh = shape[4] if len(shape) > 4 else 64
    ld = (H + 2 * KV) * dh
    qkv = torch.randn(B * T, ld, device=dev).to(torch.bfloat16)
    out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=dev)
    lse = torch.empty(B * H, T, device=dev)
    flops = 4.0 * B * H * dh * T * T / 2
    us = timeit(lambda: ops.attn_fwd(qkv, out, lse, batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0,
                                     k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh)))
    v5 = qkv.view(B, T, ld)
    q = v5[..., : H * dh].view(B, T, H, dh).permute(0, 2, 1, 3)
    k = v5[..., H * dh: (H + KV) * dh].view(B, T, KV, dh).permute(0, 2, 1, 3)
    v = v5[..., (H + KV) * dh:].view(B, T, KV, dh).permute(0, 2, 1, 3)
    us_lib = timeit(lambda: F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(H != KV)))
    o_lib = F.scaled_dot_product_attention(q, k, v, is_causal=True, enable_gqa=(H != KV)).permute(0, 2, 1, 3).reshape(B * T, H * dh)
    err = (out.float() - o_lib.float()).abs().max().item()
    print(f"B={B} T={T} H={H} KV={KV} dh={dh}: own {us:8.1f} us ({flops/us/1e6:6.0f} TF)   SDPA {us_lib:8.1f} us ({flops/us_lib/1e6:6.0f} TF)   max|diff| {err:.4f}")
    # ---- backward: own (delta + dk|dv + dq kernels) vs SDPA's autograd (cuDNN) ----
    dout = torch.randn(B * T, H * dh, device=dev).to(torch.bfloat16)
    dqkv = torch.empty(B * T, ld, dtype=torch.bfloat16, device=dev)
    ws = ops.attn_bwd_workspace(B, T, H, dh, dev)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh,
              scale=1 / math.sqrt(dh))
    us_b = timeit(lambda: ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw))
    us_bd = timeit(lambda: ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, deterministic=True, **kw))
    qg, kg, vg = (t.detach().requires_grad_(True) for t in (q, k, v))
    og = F.scaled_dot_product_attention(qg, kg, vg, is_causal=True, enable_gqa=(H != KV))
    go = dout.view(B, T, H, dh).permute(0, 2, 1, 3)
    us_bl = timeit(lambda: torch.autograd.grad(og, (qg, kg, vg), go, retain_graph=True))
    fb = 2.5 * flops
    print(f"      backward: own {us_b:8.1f} us ({fb/us_b/1e6:6.0f} TF)  [two-kernel {us_bd:8.1f} us]   SDPA {us_bl:8.1f} us ({fb/us_bl/1e6:6.0f} TF)")
