#!/usr/bin/env python
"""Top stall-sample instructions of one kernel from `ncu -i rep --page source --csv` output.
usage: ncu -i rep --page source --csv --kernel-id ::regex:NAME:N > src.csv ; python profiles/ncu_source_top.py src.csv [N]"""
import csv
import sys

rows = [r for r in csv.reader(open(sys.argv[1]))]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
hi = next(i for i, r in enumerate(rows) if "# Samples" in r)
hdr = rows[hi]
iS, isrc, iex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall_cols = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[hi + 1:]):
    if len(r) <= iS or not r[iS].isdigit():
        continue
    st = sorted(((int(r[i] or 0), h) for i, h in stall_cols), reverse=True)[:2]
    data.append((int(r[iS]), r[isrc].strip(), int(r[iex] or 0), n, st))
tot = sum(d[0] for d in data)
print(f"total samples {tot}, warp instructions executed {sum(d[2] for d in data)}")
for s, src, ex, n, st in sorted(data, reverse=True)[:top]:
    why = " ".join(f"{h[6:]}={c}" for c, h in st if c)
    print(f"{s:7d} {100 * s / max(tot, 1):5.1f}% ex={ex:9d} #{n:5d} {src[:70]:70s} {why}")
