"""Per-phase clock readings of CTA 0 of the attention forward (library built with -DB200FT_ATTN_TIMING; debug only).
  python tools/attn_timing.py B T H KV dh"""
import ctypes, math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "finetune-controller_b200"))
import numpy as np, torch
from b200ft import ops
from b200ft._lib import lib
B, T, H, KV, dh = (int(x) for x in sys.argv[1:6])
dev = "cuda:0"
ld = (H + 2 * KV) * dh
qkv = torch.randn(B * T, ld, device=dev).to(torch.bfloat16)
out = torch.empty(B * T, H * dh, dtype=torch.bfloat16, device=dev)
lse = torch.empty(B * H, T, device=dev)
run = lambda: ops.attn_fwd(qkv, out, lse, batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh,
                           v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
for _ in range(3): run()
L = lib()
log = np.zeros((2, 4096, 2), dtype=np.int64); cnt = np.zeros(2, dtype=np.int32)
L.b200ft_debug_attn_timing.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
L.b200ft_debug_attn_timing(None, None, 1)
run()
L.b200ft_debug_attn_timing(log.ctypes.data_as(ctypes.c_void_p), cnt.ctypes.data_as(ctypes.c_void_p), 0)
names = {0: "wait S", 1: "S ready", 2: "scores loaded", 3: "max found", 4: "exchanged", 5: "exps issued", 6: "P handed over",
         10: "wait P_A", 11: "P_A ready", 12: "PV_A issued, K/S_A free", 13: "S_A issued; wait P_B", 14: "P_B ready",
         15: "PV_B issued, S_B free", 16: "S_B issued"}
for role in (0, 1):
    n = min(int(cnt[role]), 4096)
    ev = log[role, :n]
    if n == 0:
        continue
    print(f"--- role {role}: {n} events")
    # per-tag mean delta to the previous event, over the steady state (skip first 20 %, last 10 %)
    lo, hi = n // 5, n - n // 10
    acc = {}
    for i in range(max(lo, 1), hi):
        acc.setdefault((int(ev[i - 1, 0]), int(ev[i, 0])), []).append(int(ev[i, 1] - ev[i - 1, 1]))
    for (a, b), v in sorted(acc.items()):
        print(f"  {names.get(a, a):>26} -> {names.get(b, b):<26} mean {np.mean(v):7.0f}  min {np.min(v):6d}  max {np.max(v):6d}  n {len(v)}")
    per = [int(ev[i, 1]) for i in range(lo, hi) if ev[i, 0] == (0 if role == 0 else 10)]
    if len(per) > 2:
        print(f"  period: {np.mean(np.diff(per)):.0f} clk")

# ---- backward, d_head 128 (two kernels: MODE 0 = dK|dV, MODE 1 = dQ) ----
if dh == 128 and hasattr(L, "b200ft_debug_attn_bwd128_timing"):
    dout = torch.randn(B * T, H * dh, device=dev).to(torch.bfloat16)
    dqkv = torch.empty(B * T, ld, dtype=torch.bfloat16, device=dev)
    ws = ops.attn_bwd_workspace(B, T, H, dh, dev)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
    for _ in range(3): ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw)
    fn = L.b200ft_debug_attn_bwd128_timing
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    blog = np.zeros((2, 2, 4096, 2), dtype=np.int64); bcnt = np.zeros((2, 2), dtype=np.int32)
    fn(None, None, 1)
    ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw)
    fn(blog.ctypes.data_as(ctypes.c_void_p), bcnt.ctypes.data_as(ctypes.c_void_p), 0)
    bn = {0: "wait S,dP", 1: "S,dP ready", 2: "P computed", 3: "dS computed", 4: "handed over",
          10: "wait P,dS", 11: "P,dS ready", 12: "grads issued", 13: "products issued"}
    for mode in (0, 1):
        for role in (0, 1):
            n = min(int(bcnt[mode, role]), 4096)
            ev = blog[mode, role, :n]
            print(f"--- backward MODE {mode} role {role}: {n} events")
            lo, hi = n // 5, n - n // 10
            acc = {}
            for i in range(max(lo, 1), hi):
                acc.setdefault((int(ev[i - 1, 0]), int(ev[i, 0])), []).append(int(ev[i, 1] - ev[i - 1, 1]))
            for (a, b), v in sorted(acc.items()):
                print(f"  {bn.get(a, a):>18} -> {bn.get(b, b):<18} mean {np.mean(v):7.0f}  min {np.min(v):6d}  max {np.max(v):6d}  n {len(v)}")
            per = [int(ev[i, 1]) for i in range(lo, hi) if ev[i, 0] == (0 if role == 0 else 10)]
            if len(per) > 2:
                print(f"  period: {np.mean(np.diff(per)):.0f} clk")

# ---- backward, d_head 64 (fused kernel, MODE 2) ----
if dh == 64 and hasattr(L, "b200ft_debug_attn_bwd64_timing"):
    dout = torch.randn(B * T, H * dh, device=dev).to(torch.bfloat16)
    dqkv = torch.empty(B * T, ld, dtype=torch.bfloat16, device=dev)
    ws = ops.attn_bwd_workspace(B, T, H, dh, dev)
    kw = dict(batch=B, seq=T, n_head=H, n_kv_head=KV, d_head=dh, q_off=0, k_off=H * dh, v_off=(H + KV) * dh, scale=1 / math.sqrt(dh))
    for _ in range(3): ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw)
    fn = L.b200ft_debug_attn_bwd64_timing
    fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
    dlog = np.zeros((2, 4096, 2), dtype=np.int64); dcnt = np.zeros(2, dtype=np.int32)
    fn(None, None, 1)
    ops.attn_bwd(qkv, out, dout, lse, ws, dqkv, **kw)
    fn(dlog.ctypes.data_as(ctypes.c_void_p), dcnt.ctypes.data_as(ctypes.c_void_p), 0)
    dn = {0: "wait S", 1: "S ready", 2: "P computed, S freed", 3: "P/dS tiles free", 4: "P stored", 5: "dP ready", 6: "dS stored, handed over", 7: "dQ drained",
          10: "pair start", 11: "S(next) issued", 12: "P,dS ready", 13: "dP(next) issued", 14: "dQ buffer free", 15: "grads issued"}
    for role in (0, 1):
        n = min(int(dcnt[role]), 4096); ev = dlog[role, :n]
        print(f"--- backward d64 fused, role {role}: {n} events")
        lo, hi = n // 5, n - n // 10
        acc = {}
        for i in range(max(lo, 1), hi):
            acc.setdefault((int(ev[i - 1, 0]), int(ev[i, 0])), []).append(int(ev[i, 1] - ev[i - 1, 1]))
        for (a, b), v in sorted(acc.items()):
            print(f"  {dn.get(a, a):>24} -> {dn.get(b, b):<24} mean {np.mean(v):7.0f}  min {np.min(v):6d}  max {np.max(v):6d}  n {len(v)}")
        per = [int(ev[i, 1]) for i in range(lo, hi) if ev[i, 0] == (0 if role == 0 else 10)]
        if len(per) > 2: print(f"  period: {np.mean(np.diff(per)):.0f} clk")
