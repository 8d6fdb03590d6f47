"""Most-sampled SASS lines of `ncu -i X.ncu-rep --page source --csv` (file argument): share of warp-stall samples,
executions, dominant stall reason, plus the opcode mix of the hot loop.  python tools/ncu_top_stalls.py src.csv [n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
h, body = rows[hi], [r for r in rows[hi + 1:] if len(r) > 6]
ci, ei, si = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
stall_cols = [(i, x) for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
tot = sum(float(r[ci] or 0) for r in body)
print(f"total samples {tot:.0f}, warp instructions executed {sum(float(r[ei] or 0) for r in body):.0f}")
agg = {}
for r in body:
    for i, x in stall_cols:
        agg[x] = agg.get(x, 0.0) + float(r[i] or 0)
print("by reason: " + ", ".join(f"{k[6:]} {100 * v / tot:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
for r in sorted(body, key=lambda r: -float(r[ci] or 0))[:n]:
    why = max(stall_cols, key=lambda c: float(r[c[0]] or 0))[1][6:]
    print(f"{100 * float(r[ci]) / tot:5.1f}%  exec {r[ei]:>8}  {why:<14} {r[si].strip()[:100]}")
# opcode mix, weighted by executions
mix = collections.Counter()
for r in body:
    ops = [x for x in r[si].strip().split() if not x.startswith("@")]
    if ops:
        mix[ops[0].split(".")[0]] += float(r[ei] or 0)
allx = sum(mix.values())
print("opcode mix: " + ", ".join(f"{k} {100 * v / allx:.1f}%" for k, v in mix.most_common(16)))
