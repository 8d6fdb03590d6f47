"""Short single-purpose commands for ncu: one launch class each, at the bench's sizes.
  python tools/hbm_probe.py xent | adamw | lnfwd | lnbwd | loragrad"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch
dev = "cuda:0"
from b200ft import ops, _lib
if os.environ.get("B200FT_PROBE_LIB"): _lib.LIB_PATH = os.path.join(ROOT, os.environ["B200FT_PROBE_LIB"])  # variant build, measurement only
what = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
def run(fn, bytes_):
    for _ in range(2): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    print(f"{what}: {us:.1f} us/launch, {bytes_/us/1e3:.0f} GB/s algorithmic")
if what == "xent":
    # python tools/hbm_probe.py xent [iters] [rows] [vocab]: the kernel alone, in place on a buffer far larger than L2
    # (after the first launch the buffer holds gradients: still a valid input, same traffic and instruction count)
    M = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
    V = int(sys.argv[4]) if len(sys.argv) > 4 else 50257
    ld = (V + 127) // 128 * 128
    z = (torch.randn(M, ld, device=dev) * 3).to(torch.bfloat16)
    t = torch.randint(0, V, (M,), device=dev); loss = torch.zeros(1, device=dev); ws = ops.softmax_xent_workspace(M, dev)
    def f():
        ops.softmax_xent_fwd_bwd(z, t, loss, ws, vocab=V)
    run(f, 2.0 * M * V * 2 + 12 * M)
elif what == "adamw":
    n = 256 * 1024 * 1024
    p = torch.zeros(n, device=dev); m = torch.zeros(n, device=dev); v = torch.zeros(n, device=dev)
    g = torch.full((n,), 0.01, dtype=torch.bfloat16, device=dev); sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
    st = [0]
    def f():
        st[0] += 1
        ops.adamw_step(p, m, v, g, step=st[0], lr=1e-4, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=0.01, shadow=sh)
    run(f, 28.0 * n)
elif what in ("lnfwd", "lnbwd"):
    M, d = 32768, 768
    xs = [torch.randn(M, d, device=dev).to(torch.bfloat16) for _ in range(8)]
    y = torch.empty_like(xs[0]); g = torch.ones(d, device=dev); b = torch.zeros(d, device=dev)
    mean = torch.empty(M, device=dev); rstd = torch.empty(M, device=dev)
    ops.layernorm_fwd(xs[0], g, b, y, mean, rstd, 1e-5)
    i = [0]
    if what == "lnfwd":
        def f():
            i[0] += 1; ops.layernorm_fwd(xs[i[0] % 8], g, b, y, mean, rstd, 1e-5)
        run(f, 4.0 * M * d)
    else:
        def f():
            i[0] += 1; ops.layernorm_bwd_dx(xs[i[0] % 8], xs[(i[0] + 3) % 8], g, mean, rstd, y, dres=xs[(i[0] + 5) % 8])
        run(f, 8.0 * M * d)
elif what == "loragrad":
    M, cols = 32768, 768
    S = torch.randn(M, 16, device=dev).to(torch.bfloat16); T = torch.randn(M, cols, device=dev).to(torch.bfloat16)
    out = torch.zeros(16, cols, device=dev)
    run(lambda: ops.lora_grad(S, T, out, r=16, cols=cols), 2.0 * M * (cols + 16))
