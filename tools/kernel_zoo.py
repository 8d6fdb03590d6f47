"""One launch of every kernel of the library that the step-level profiles do not already cover,
at the sizes the workloads use — the short command ncu profiles for profiles/*_zoo_*.  Prints
CUDA-event times (the numbers quoted; ncu's own are cold-cache)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch  # noqa: E402
from b200ft import ops  # noqa: E402

dev = "cuda:0"
bf = dict(dtype=torch.bfloat16, device=dev)
f32 = dict(dtype=torch.float32, device=dev)
HBM = 6549.8


ITERS = int(os.environ.get("ZOO_ITERS", "5"))  # 0 under ncu: one launch per kernel


def timed(name, bytes_, fn, iters=None):
    iters = ITERS if iters is None else iters
    fn()
    if iters == 0:
        return
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    gbs = bytes_ / us / 1e3
    print(f"{name:44s} {us:9.1f} us  {gbs:7.0f} GB/s  {100 * gbs / HBM:5.1f} % of measured HBM", flush=True)


def rnd(*shape):
    return torch.randn(*shape, device=dev).to(torch.bfloat16)


# LayerNorm at C2 (32768 x 768)
M, d = 32768, 768
x, y, dy, dx, dres = rnd(M, d), torch.empty(M, d, **bf), rnd(M, d), torch.empty(M, d, **bf), rnd(M, d)
g, b = torch.ones(d, **f32), torch.zeros(d, **f32)
mean, rstd = torch.empty(M, **f32), torch.empty(M, **f32)
timed("layernorm_fwd 32768x768", 4.0 * M * d, lambda: ops.layernorm_fwd(x, g, b, y, mean, rstd, 1e-5))
timed("layernorm_bwd_dx(+dres) 32768x768", 8.0 * M * d, lambda: ops.layernorm_bwd_dx(dy, x, g, mean, rstd, dx, dres=dres))
# RMSNorm at C3 (8192 x 2048) and C4/C5 (4096 x 4096), dgamma
for M, d in ((8192, 2048), (4096, 4096)):
    x, y, dy, dx, dres = rnd(M, d), torch.empty(M, d, **bf), rnd(M, d), torch.empty(M, d, **bf), rnd(M, d)
    g, rstd, dg = torch.ones(d, **f32), torch.empty(M, **f32), torch.zeros(d, **f32)
    timed(f"rmsnorm_fwd {M}x{d}", 4.0 * M * d, lambda: ops.rmsnorm_fwd(x, g, y, rstd, 1e-5))
    timed(f"rmsnorm_bwd_dx(+dres) {M}x{d}", 8.0 * M * d, lambda: ops.rmsnorm_bwd_dx(dy, x, g, rstd, dx, dres=dres))
    timed(f"rmsnorm_bwd_dgamma {M}x{d}", 4.0 * M * d, lambda: ops.rmsnorm_bwd_dgamma(dy, x, rstd, dg))
# RoPE + SwiGLU at C5 (4096 tokens, 32+8 heads of 128, ff 14336)
M, H, dh, ff = 4096, 40, 128, 14336
qkv = rnd(M, 6144)
from b200ft.llama import LLAMA_3_8B, rope_tables  # noqa: E402
cos, sin = rope_tables(LLAMA_3_8B, 4096, dev)
timed("rope_inplace 4096 x 40 heads x 128", 4.0 * M * H * dh, lambda: ops.rope_inplace(qkv, cos, sin, seq=4096, n_heads=H, d_head=dh))
gu, act, dact, dgu = rnd(M, 2 * ff), torch.empty(M, ff, **bf), rnd(M, ff), torch.empty(M, 2 * ff, **bf)
timed("swiglu_fwd 4096x14336", 6.0 * M * ff, lambda: ops.swiglu_fwd(gu, act))
timed("swiglu_bwd 4096x14336", 10.0 * M * ff, lambda: ops.swiglu_bwd(dact, gu, dgu))
timed("swiglu_bwd + act rewrite 4096x14336", 12.0 * M * ff, lambda: ops.swiglu_bwd(dact, gu, dgu, act=act))
# embedding gather / scatter-add at C5
V, d = 128256, 4096
wte, dwte = rnd(V, d), torch.zeros(V, d, **bf)
ids = torch.randint(0, V, (M,), device=dev)
xe, dxe = torch.empty(M, d, **bf), rnd(M, d)
timed("embed_fwd 4096 tokens x 4096", 4.0 * M * d, lambda: ops.embed_fwd(ids, wte, None, xe, seq=4096, vocab=V))
timed("embed_bwd 4096 tokens x 4096", 6.0 * M * d, lambda: ops.embed_bwd(ids, dxe, dwte, vocab=V))
del wte, dwte
# adapter gradient (split-K, atomic transposed epilogue) at C2
M, cols, r = 32768, 2304, 16
U, dY, out = rnd(M, r), rnd(M, cols), torch.zeros(r, cols, **f32)
timed("lora_grad 32768 x 16 x 2304", 2.0 * M * (cols + r), lambda: ops.lora_grad(U, dY, out, r=r, cols=cols))
# helpers
a, at = rnd(4096, 6144), torch.empty(6144, 4096, **bf)
timed("transpose_bf16 4096x6144", 4.0 * 4096 * 6144, lambda: ops.transpose(a, at))
n = 64 * 1024 * 1024
pf, pb = torch.randn(n, device=dev), torch.empty(n, **bf)
timed("cast_f32_to_bf16 64Mi", 6.0 * n, lambda: ops.cast_f32_to_bf16(pf, pb))
ws, sq = ops.gradnorm_workspace(n, dev), torch.zeros(1, **f32)
timed("gradnorm_sq bf16 64Mi", 2.0 * n, lambda: ops.gradnorm_sq(pb, sq, ws))
# MN-major GEMMs of the full-parameter step
for (m, n_, k, amn, bmn) in ((28672, 4096, 4096, True, True), (4096, 28672, 4096, False, True)):
    A = rnd(k, m) if amn else rnd(m, k)
    B = rnd(k, n_) if bmn else rnd(n_, k)
    C = torch.empty(m, n_, **bf)
    fl = 2.0 * m * n_ * k
    fn = lambda: ops.gemm(A, B, C, a_mn=amn, b_mn=bmn)
    fn()
    if ITERS == 0:
        continue
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(5):
        fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 5
    print(f"gemm {m}x{n_}x{k} a_mn={int(amn)} b_mn={int(bmn)}: {us:9.1f} us  {fl / us / 1e6:7.0f} TFLOP/s", flush=True)
