"""Compact per-kernel table from `ncu -i X.ncu-rep --page raw --csv` (stdin or file): duration, DRAM
bytes and throughput, tensor-pipe activity, occupancy, registers.  Output goes under profiles/."""
import csv
import sys

src = open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin
rows = list(csv.reader(src))
hdr = rows[0]
want = {
    "Kernel Name": "kernel",
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed": "tensor_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occ_pct",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__block_size": "block",
}
idx = {}
for i, h in enumerate(hdr):
    for k, v in want.items():
        if h == k or h.endswith("." + k):
            idx.setdefault(v, i)
units = rows[1]
out = csv.writer(sys.stdout)
cols = [c for c in want.values() if c in idx]
out.writerow([c + (f" [{units[idx[c]]}]" if units[idx[c]] else "") for c in cols])
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    vals = [r[idx[c]] for c in cols]
    vals[0] = vals[0][:90]
    out.writerow(vals)
