#!/usr/bin/env python3
"""Instructions executed and stall samples per CUDA source line / code section of k_poisson_fused, from
`ncu -i <rep> --page source --csv --print-source cuda,sass` (capture taken with --import-source on, -lineinfo build).
usage: ncu_by_line.py <csv> [top_n]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, data = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr and len(r) > 8 and r[0].isdigit():
        try:
            inst, smp = int(r[hdr.index("Instructions Executed")]), int(r[hdr.index("# Samples")])
        except ValueError:
            continue
        if inst or smp:
            data.append((cur, int(r[0]), inst, smp, r[1].strip()[:100]))
tot, ts = sum(d[2] for d in data), sum(d[3] for d in data)
print(f"warp instructions {tot}   stall samples {ts}")
by_file = collections.Counter()
for f, l, i, s, t in data:
    by_file[f] += i
print("by file:", ", ".join(f"{k} {v / tot * 100:.1f}%" for k, v in by_file.most_common()))
for d in sorted(data, key=lambda d: -d[2])[:top]:
    print(f"{d[0]}:{d[1]:5d} inst {d[2] / tot * 100:5.2f}% smp {d[3] / ts * 100:5.2f}%  {d[4]}")
print("-- by stall samples")
for d in sorted(data, key=lambda d: -d[3])[:15]:
    print(f"{d[0]}:{d[1]:5d} inst {d[2] / tot * 100:5.2f}% smp {d[3] / ts * 100:5.2f}%  {d[4]}")
