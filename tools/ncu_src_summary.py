#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --csv` output: top SASS instructions by stall samples, totals."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = rows[1]
ci = {n: i for i, n in enumerate(h)}
data = rows[2:]
tot_s = sum(int(r[ci["# Samples"]] or 0) for r in data)
tot_i = sum(int(r[ci["Instructions Executed"]] or 0) for r in data)
print("instructions (SASS lines): %d, warp-instructions executed: %d, samples: %d" % (len(data), tot_i, tot_s))
stalls = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
agg = {k: sum(int(r[ci[k]] or 0) for r in data) for k in stalls}
print("stall totals:", sorted(agg.items(), key=lambda kv: -kv[1])[:8])
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# contiguous windows: print top instructions in address order with cumulative
idx = sorted(range(len(data)), key=lambda i: -int(data[i][ci["# Samples"]] or 0))[:top]
for i in sorted(idx):
    r = data[i]
    print("%5d %6s smp %9s exec  %s" % (i, r[ci["# Samples"]], r[ci["Instructions Executed"]], r[ci["Source"]].strip()[:90]))
