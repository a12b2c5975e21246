#!/usr/bin/env python3
"""Attribute an ncu SASS source page to CUDA source lines through nvdisasm's line info.
usage: ncu_lines.py <ncu source csv> <nvdisasm -g -c listing of the kernel> [top_n]
(ncu's own --print-source cuda CSV carries no metrics, so the join is done here by SASS offset.)"""
import collections
import csv
import re
import sys

src_csv, sass, top = sys.argv[1], sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else 40
line_of = {}
cur = None
for ln in open(sass):
    m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', ln)
    if m:
        chain = [(m.group(1).split("/")[-1], int(m.group(2)))]
        for f, l in re.findall(r'inlined at "([^"]+)", line (\d+)', m.group(3)):
            chain.append((f.split("/")[-1], int(l)))
        cur = chain
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*);", ln)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(src_csv)))
H = rows[1]
iA, iS, iI, iT = H.index("Address"), H.index("# Samples"), H.index("Instructions Executed"), H.index("Thread Instructions Executed")
data = []
seen = set()
for r in rows[2:]:
    if len(r) <= iT or r[iA] in seen:
        continue
    seen.add(r[iA])
    try:
        data.append((int(r[iA], 16) if r[iA].startswith("0x") else int(r[iA]), int(r[iS]), int(r[iI]), int(r[iT])))
    except ValueError:
        pass
base = min(d[0] for d in data)
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for a, s, i, t in data:
    chain = line_of.get(a - base, [("?", 0)])
    # attribute to the innermost frame inside our own sources
    key = next(((f, l) for f, l in chain if f.startswith("pic_")), chain[-1])
    g = agg[key]
    g[0] += s; g[1] += i; g[2] += t; g[3] += 1
ts, ti = sum(g[0] for g in agg.values()), sum(g[1] for g in agg.values())
print(f"samples {ts}  warp-inst {ti}")
print("by warp instructions:")
for k, g in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"  {k[0]}:{k[1]:<4d} inst {g[1]/1e6:8.2f}M {100*g[1]/ti:5.1f}%  samples {100*g[0]/ts:5.1f}%  lanes {g[2]/max(1,g[1]):4.1f}  sass {g[3]}")
print("by samples:")
for k, g in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"  {k[0]}:{k[1]:<4d} samples {100*g[0]/ts:5.1f}%  inst {g[1]/1e6:8.2f}M  lanes {g[2]/max(1,g[1]):4.1f}")
