"""LayerNorm / softmax-CE / AdamW micro-probe (HBM GB/s with CUDA events)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
from b200ft import ops
dev = "cuda:0"
M, d = 32768, 768
x = torch.randn(M, d, device=dev).to(torch.bfloat16); y = torch.empty_like(x); dy = torch.randn_like(x); dres = torch.randn_like(x); dx = torch.empty_like(x)
g = torch.ones(d, device=dev); b = torch.zeros(d, device=dev); mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
big = [torch.randn(M, d, device=dev).to(torch.bfloat16) for _ in range(8)]  # rotate inputs: working set > L2
def timeit(fn, n=40):
    for i in range(5): fn(i)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n
us = timeit(lambda i: ops.layernorm_fwd(big[i % 8], g, b, y, mean, rstd, 1e-5))
print(f"layernorm_fwd  {us:7.1f} us  {4.0*M*d/us/1e3:7.0f} GB/s")
ops.layernorm_fwd(x, g, b, y, mean, rstd, 1e-5)
us = timeit(lambda i: ops.layernorm_bwd_dx(big[i % 8], big[(i + 3) % 8], g, mean, rstd, dx, dres=big[(i + 5) % 8]))
print(f"layernorm_bwd  {us:7.1f} us  {8.0*M*d/us/1e3:7.0f} GB/s")
