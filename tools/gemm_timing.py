"""Per-phase clock readings of CTA 0 of the tcgen05 GEMM (library built with -DB200FT_ATTN_TIMING; debug only).
  python tools/gemm_timing.py M N K epilogue"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import numpy as np, torch
from b200ft import ops
from b200ft._lib import lib
M, N, K, epi = (int(x) for x in sys.argv[1:5])
dev = "cuda:0"
a = torch.randn(M, K, device=dev).to(torch.bfloat16); b = torch.randn(N, K, device=dev).to(torch.bfloat16)
out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
kw = {}
if epi == 2: kw["out2"] = torch.empty_like(out)
if epi in (1, 3): kw["aux"] = torch.randn(M, N, device=dev).to(torch.bfloat16)
run = lambda: ops.gemm(a, b, out, epilogue=epi, **kw)
for _ in range(3): run()
L = lib(); fn = L.b200ft_debug_gemm_timing; fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
log = np.zeros((2, 8192, 2), dtype=np.int64); cnt = np.zeros(2, dtype=np.int32)
fn(None, None, 1); run(); fn(log.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p), 0)
names = {20: "tile start", 21: "accumulator free", 22: "k-block start", 23: "operands landed", 33: "chunk start", 34: "aux landed", 35: "acc in regs", 36: "math + sts done", 37: "store issued", 38: "older store read", 39: "next aux issued", 30: "epi: wait acc", 31: "epi: acc ready", 32: "epi: tile drained"}
for role in (1, 0):
    n = min(int(cnt[role]), 8192); ev = log[role, :n]
    print(f"--- role {role} ({'issuer' if role else 'epilogue warp 0'}): {n} events, {int(ev[-1,1]-ev[0,1])} clk in all")
    lo, hi = n // 6, n - n // 10
    acc = {}
    for i in range(max(lo, 1), hi):
        acc.setdefault((int(ev[i - 1, 0]), int(ev[i, 0])), []).append(int(ev[i, 1] - ev[i - 1, 1]))
    for (x, y), v in sorted(acc.items()):
        print(f"  {names.get(x, x):>18} -> {names.get(y, y):<18} mean {np.mean(v):7.0f}  min {np.min(v):6d}  max {np.max(v):6d}  n {len(v)}  sum {np.sum(v)}")
    per = [int(ev[i, 1]) for i in range(lo, hi) if ev[i, 0] == (20 if role else 30)]
    if len(per) > 2: print(f"  tile period: {np.mean(np.diff(per)):.0f} clk")
