"""Fused AdamW bandwidth against buffer size (bf16 grads + bf16 shadow, 28 B/param): is the
8 B-parameter launch of the full-parameter step slower per byte than the 256 Mi-element one?"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch  # noqa: E402
from b200ft import ops  # noqa: E402

dev = "cuda:0"
for n in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "268435456,1073741824,4294967296").split(",")]:
    p = torch.zeros(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    g = torch.full((n,), 0.01, dtype=torch.bfloat16, device=dev)
    sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
    kw = dict(lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, shadow=sh)
    for i in range(2):
        ops.adamw_step(p, m, v, g, step=i + 1, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    iters = 5
    for i in range(iters):
        ops.adamw_step(p, m, v, g, step=i + 3, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"n={n}: {ms:.3f} ms  {28.0 * n / ms / 1e6:.0f} GB/s", flush=True)
    del p, m, v, g, sh
    torch.cuda.empty_cache()
