#!/usr/bin/env python
"""Summarise `ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`: executed warp instructions and stall samples
per CUDA source line (the rows whose Address column is "-"), grouped per file."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
cur, h, lines = None, None, []
for r in rows:
    if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": h = r; continue
    if not h or len(r) < 10 or r[2] != "-": continue
    ci = {n: i for i, n in enumerate(h)}
    try:
        lines.append((cur, int(r[0]), int(r[ci["Instructions Executed"]] or 0), int(r[ci["# Samples"]] or 0),
                      int(r[ci["L1 Wavefronts Shared Excessive"]] or 0) if "L1 Wavefronts Shared Excessive" in ci else 0, r[1].strip()))
    except ValueError:
        pass
tot = sum(l[2] for l in lines) or 1
ts = sum(l[3] for l in lines) or 1
print("warp instructions %d, samples %d" % (tot, ts))
for f, ln, ie, sm, ex, src in lines:
    if ie * 100.0 / tot >= thr or sm * 100.0 / ts >= thr:
        print("%-18s %4d  inst %5.1f%%  smp %5.1f%%  bankx %9d  %s" % (f, ln, ie * 100.0 / tot, sm * 100.0 / ts, ex, src[:100]))
