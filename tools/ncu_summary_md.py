#!/usr/bin/env python
"""Render `ncu -i X.ncu-rep --page raw --csv` as the markdown table of profiles/*_ncu_full_summary.md.

usage: python tools/ncu_summary_md.py raw.csv [raw2.csv ...] > table.md   (one table per file)"""
import csv
import sys

COLS = [("kernel", "Kernel Name"), ("time", "gpu__time_duration.sum"), ("dram rd", "dram__bytes_read.sum"),
        ("dram wr", "dram__bytes_write.sum"), ("sm %", "sm__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("issue %", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
        ("l1tex %", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("l2 %", "lts__throughput.avg.pct_of_peak_sustained_elapsed"),
        ("occ %", "sm__warps_active.avg.pct_of_peak_sustained_active"), ("regs", "launch__registers_per_thread"),
        ("warp inst", "smsp__inst_executed.sum")]


def main():
    for path in sys.argv[1:]:
        rows = list(csv.reader(open(path)))
        h, u = rows[0], rows[1]
        cols = [(t, h.index(n)) for t, n in COLS if n in h]
        print("| " + " | ".join(t for t, _ in cols) + " |")
        print("|" + "---|" * len(cols))
        for r in rows[2:]:
            cells = []
            for t, i in cols:
                v = r[i]
                if t == "kernel":
                    v = v.split("(")[0].split("::")[-1].replace("void ", "")
                elif t in ("time", "dram rd", "dram wr"):
                    v = "%s %s" % (v, u[i])
                elif t == "warp inst":
                    v = "%.3e" % float(v.replace(",", ""))
                elif t != "regs":
                    v = "%.1f" % float(v.replace(",", ""))
                cells.append(v)
            print("| " + " | ".join(cells) + " |")
        print()


if __name__ == "__main__":
    main()
