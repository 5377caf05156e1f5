#!/usr/bin/env python
"""Summarise an ncu report for profiles/: one row per profiled launch with the counters the roofline is read from.

usage: ncu -i X.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv [images_per_launch] > profiles/NAME.md
"""
import csv
import sys

COLS = [
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram rd"),
    ("dram__bytes_write.sum", "dram wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %"),
    ("l1tex__throughput.avg.pct_of_peak_sustained_active", "l1tex %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2 %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %"),
    ("launch__registers_per_thread", "regs"),
    ("smsp__inst_executed.sum", "warp inst"),
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h, units = rows[0], rows[1]
    ci = {n: i for i, n in enumerate(h)}
    print("| kernel | " + " | ".join(c[1] for c in COLS) + " |")
    print("|---|" + "---|" * len(COLS))
    for r in rows[2:]:
        name = r[ci["Kernel Name"]].split("(")[0]
        cells = []
        for key, _ in COLS:
            if key not in ci:
                cells.append("-")
                continue
            v, u = r[ci[key]], units[ci[key]]
            try:
                f = float(v.replace(",", ""))
                v = ("%.4g" % f)
            except ValueError:
                pass
            cells.append((v + " " + u).strip() if u not in ("%", "inst", "register/thread") else v)
        print("| %s | %s |" % (name, " | ".join(cells)))


if __name__ == "__main__":
    main()
