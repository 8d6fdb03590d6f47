"""GEMM micro-probe: times b200ft_gemm_bf16 on the step's shapes for each (cta_group, block_n)
with CUDA events.  Used to choose tile configurations and as the short command ncu profiles."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import torch  # noqa: E402
from b200ft import _lib, ops  # noqa: E402

if os.environ.get("B200FT_PROBE_LIB"):  # a variant build of the library (tools/_exp/), measurement only
    _lib.LIB_PATH = os.path.join(ROOT, os.environ["B200FT_PROBE_LIB"])


def run(M, N, K, cg, bn, epi, iters, k2=0, layout=0, max_ctas=0):
    dev = "cuda:0"
    a_mn, b_mn = bool(layout & 1), bool(layout & 2)
    a = torch.randn((K, M) if a_mn else (M, K), device=dev).to(torch.bfloat16)
    b = torch.randn((K, N) if b_mn else (N, K), device=dev).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    kw = {}
    if epi == 2:
        kw["out2"] = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    if epi in (1, 3):
        kw["aux"] = torch.randn(M, N, device=dev).to(torch.bfloat16)
    if k2:
        kw["a2"] = torch.randn(M, k2, device=dev).to(torch.bfloat16)
        kw["b2"] = torch.randn(N, k2, device=dev).to(torch.bfloat16)
    bias = torch.randn(N, device=dev)
    for _ in range(3):
        ops.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, epilogue=epi, bias=bias, block_n=bn, cta_group=cg, max_ctas=max_ctas, **kw)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        ops.gemm(a, b, out, a_mn=a_mn, b_mn=b_mn, epilogue=epi, bias=bias, block_n=bn, cta_group=cg, max_ctas=max_ctas, **kw)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / iters
    return us, 2.0 * M * N * (K + k2) / us / 1e6


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="32768x768x3072x0,32768x3072x768x2,32768x3072x768x3,32768x2304x768x0,"
                                        "32768x768x768x1,32768x768x50304x0,32768x50304x768x0")
    ap.add_argument("--cgs", default="1,2")
    ap.add_argument("--layout", type=int, default=0, help="1 = A MN-major, 2 = B MN-major, 3 = both")
    ap.add_argument("--bns", default="128,192,256")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--max-ctas", type=int, default=0, help="run on this many CTAs only (0 = all SMs)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    rows = []
    for sh in args.shapes.split(","):
        M, N, K, epi = (int(x) for x in sh.split("x"))
        for cg in (int(x) for x in args.cgs.split(",")):
            for bn in (int(x) for x in args.bns.split(",")):
                us, tf = run(M, N, K, cg, bn, epi, args.iters, layout=args.layout, max_ctas=args.max_ctas)
                rows.append({"M": M, "N": N, "K": K, "epi": epi, "cg": cg, "bn": bn, "layout": args.layout, "us": us, "tflops": tf})
                print(f"M={M} N={N} K={K} epi={epi} layout={args.layout} cg={cg} bn={bn:3d}: {us:9.1f} us  {tf:7.1f} TFLOP/s", flush=True)
    if args.out:
        json.dump(rows, open(args.out, "w"), indent=1)
