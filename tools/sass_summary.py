"""Per-kernel counts of the SASS mnemonics that prove a Blackwell-native path (B200_PROFILING.md): tcgen05.mma ->
UTC*MMA, tcgen05.ld/st -> LDTM/STTM, TMA -> UTMALDG/UTMASTG/UTMAREDG/UBLKCP, legacy tensor path -> HMMA (must be 0).
  python tools/sass_summary.py [libb200ft.so] > profiles/r02_sass_summary.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "finetune-controller_b200", "b200ft", "libb200ft.so")
sass = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WANT = ["UTCHMMA.2CTA", "UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAREDG", "UBLKCP", "UBLKRED", "SYNCS",
        "MUFU.EX2", "HMMA", "FFMA2", "STAS"]
counts, order, cur = {}, [], None
for line in sass.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        name = subprocess.run(["c++filt", "-p", m.group(1)], capture_output=True, text=True).stdout.strip()
        cur = name.replace("(anonymous namespace)::", "")
        counts[cur] = collections.Counter()
        order.append(cur)
        continue
    if cur is None:
        continue
    m = re.search(r"^\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\w+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if not m:
        continue
    op = m.group(1)
    counts[cur]["instructions"] += 1
    for w in WANT:
        if op == w or op.startswith(w + "."):
            if w == "UTCHMMA" and ".2CTA" in op:
                continue
            counts[cur][w] += 1
print(f"# cuobjdump -sass {os.path.basename(lib)} (sm_100a): static instruction counts per kernel")
cols = ["instructions"] + WANT
print("kernel".ljust(64) + "".join(c[-12:].rjust(13) for c in cols))
for k in order:
    print(k[:63].ljust(64) + "".join(str(counts[k][c]).rjust(13) for c in cols))
tot = collections.Counter()
for k in order:
    tot.update(counts[k])
print("TOTAL".ljust(64) + "".join(str(tot[c]).rjust(13) for c in cols))
