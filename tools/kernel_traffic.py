#!/usr/bin/env python
"""Rebuild profiles/kernel_traffic.json from `ncu -i X.ncu-rep --page raw --csv` output.

usage: python tools/kernel_traffic.py raw.csv images_per_launch TAG [pipes.csv] > profiles/kernel_traffic.json
Per kernel name: DRAM bytes (read + write) summed over its launches / images, and the time-weighted issue-active
percentage; bench.py reads the first for `roofline.traffic` and the second for `roofline.issue_active_pct_ncu`.
pipes.csv (optional) = the `ncu --metrics sm__inst_executed_pipe_*.sum,smsp__inst_executed.sum --csv` launch list of
the same command: warp instructions per image and per pipe, the numerators of bench.py's `pipe_roofline`.
"""
import csv
import json
import re
import sys


def num(x):
    x = x.replace(",", "")
    return float(x) if x not in ("", "n/a") else 0.0


def kname(x):
    """plain kernel name: no arguments, namespace, `void ` of a template instance or template arguments"""
    x = re.sub(r"\(.*", "", x).split("::")[-1].strip()
    return re.sub(r"<.*", "", re.sub(r"^void\s+", "", x))


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    images = int(sys.argv[2])
    tag = sys.argv[3] if len(sys.argv) > 3 else "?"
    h, units = rows[0], rows[1]
    ci = {n: i for i, n in enumerate(h)}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    t_scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "nsecond": 1e-3, "usecond": 1.0, "msecond": 1e3, "second": 1e6}
    traffic, issue_w, time_sum = {}, {}, {}
    for r in rows[2:]:
        name = kname(r[ci["Kernel Name"]])
        b = 0.0
        for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            b += num(r[ci[m]]) * scale.get(units[ci[m]], 1.0)
        t = num(r[ci["gpu__time_duration.sum"]]) * t_scale.get(units[ci["gpu__time_duration.sum"]], 1.0)
        iss = num(r[ci["smsp__issue_active.avg.pct_of_peak_sustained_active"]])
        traffic[name] = traffic.get(name, 0.0) + b
        issue_w[name] = issue_w.get(name, 0.0) + iss * t
        time_sum[name] = time_sum.get(name, 0.0) + t
    out = {
        "source": "ncu --set full, %s (bench.py --pairs %d: %d images 752x480 nf=1200), dram__bytes_read.sum + "
                  "dram__bytes_write.sum per launch, summed over the launches of a stage" % (tag, images // 2, images),
        "images_per_launch": images,
        "dram_bytes_per_image": {k: int(round(v / images)) for k, v in traffic.items()},
        "issue_active_pct": {k: round(issue_w[k] / time_sum[k], 1) for k in traffic if time_sum[k] > 0},
        "issue_note": "smsp__issue_active.avg.pct_of_peak_sustained_active of the same capture (time-weighted over the "
                      "launches of a stage): the extraction kernels are bound by integer pipes / phase structure, not DRAM",
    }
    if len(sys.argv) > 4:
        pipes = {}
        for r in csv.reader(open(sys.argv[4])):
            if len(r) < 15 or r[0] == "ID":
                continue
            name = kname(r[4])
            m = r[12].replace("sm__inst_executed_pipe_", "").replace("smsp__inst_executed", "all").replace(".sum", "")
            if m.startswith("gpu__"):
                continue
            pipes.setdefault(name, {})
            pipes[name][m] = pipes[name].get(m, 0.0) + num(r[14])
        out["warp_inst_per_image"] = {k: {m: round(v / images, 1) for m, v in d.items()} for k, d in pipes.items()
                                      if k in traffic}
        out["warp_inst_note"] = ("sm__inst_executed_pipe_{alu,fma,fmaheavy,lsu,xu}.sum and smsp__inst_executed.sum of the "
                                 "same command, summed over the launches of a kernel / images")
    json.dump(out, sys.stdout, indent=1)
    sys.stdout.write("\n")


if __name__ == "__main__":
    main()
