#!/usr/bin/env python
"""Compact form of an `ncu --metrics gpu__time_duration.sum --csv` launch list (kernel names shortened, one row per launch
of this library's kernels in launch order + a per-kernel total), for committing under profiles/rNN/.

usage: python profiles/launch_summary.py gpurun_out/launches.csv > profiles/r02/launches.csv
"""
import collections
import csv
import io
import re
import sys


def short(name: str) -> str:
    name = name.replace("void ", "")
    m = re.match(r"(b200grpo::[A-Za-z0-9_]+(<[^>]*>)?)", name)
    return m.group(1) if m else re.sub(r"[<(].*", "", name)[:70]


def main():
    rows = [ln for ln in open(sys.argv[1]) if not ln.startswith("==")]
    launches = list(csv.DictReader(io.StringIO("".join(rows))))
    out = csv.writer(sys.stdout)
    out.writerow(["# ncu --metrics gpu__time_duration.sum --clock-control none (cold-cache, serialised launches: shares, not absolutes)"])
    out.writerow(["id", "kernel", "grid", "block", "time_us"])
    agg = collections.OrderedDict()
    other_n, other_us = 0, 0.0
    for x in launches:
        k, us = short(x["Kernel Name"]), float(x["Metric Value"]) / 1e3
        if k.startswith("b200grpo"):
            out.writerow([x["ID"], k, x["Grid Size"], x["Block Size"], f"{us:.1f}"])
            a = agg.setdefault(k, [0, 0.0])
            a[0] += 1
            a[1] += us
        else:
            other_n += 1
            other_us += us
    out.writerow([])
    out.writerow(["# per kernel: launches, total_us, share of this library's time"])
    total = sum(v[1] for v in agg.values()) or 1.0
    for k, (n, us) in agg.items():
        out.writerow([k, n, f"{us:.1f}", f"{us / total:.4f}"])
    out.writerow(["(torch / other kernels of the same command: workload set-up, not part of a step)", other_n, f"{other_us:.1f}", ""])


if __name__ == "__main__":
    main()
