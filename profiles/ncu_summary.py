#!/usr/bin/env python
"""Summarise an .ncu-rep (read here with `ncu -i`) into the handful of metrics DESIGN.md cites.
usage: python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [more.ncu-rep ...] > profiles/ncu_rNN_summary.md"""
import csv
import io
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "time"),
    ("sm__cycles_elapsed.avg", "sm cycles"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe active %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput %"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("l1tex__m_xbar2l1tex_read_bytes.sum", "L2->SM read"),
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "regs/thread"),
    ("smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct", "stall long_scoreboard %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
]


def main():
    for path in sys.argv[1:]:
        out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(out)))
        hdr, units = rows[0], rows[1]
        print(f"### {path}\n")
        print("| kernel | " + " | ".join(n for _, n in WANT) + " |")
        print("|---|" + "---|" * len(WANT))
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            u = dict(zip(hdr, units))
            name = d["Kernel Name"].split("(")[0].replace("void ", "")
            cells = []
            for k, _ in WANT:
                v = d.get(k, "")
                try:
                    v = f"{float(v):.4g}"
                except ValueError:
                    pass
                cells.append(f"{v} {u.get(k, '')}".strip())
            print(f"| `{name}` | " + " | ".join(cells) + " |")
        print()


if __name__ == "__main__":
    main()
