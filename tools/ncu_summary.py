#!/usr/bin/env python3
"""Summarise an ncu export: tools/ncu_summary.py <raw.csv> <source.csv> [top_n]"""
import collections
import csv
import sys

raw, src = sys.argv[1], sys.argv[2]
top_n = int(sys.argv[3]) if len(sys.argv) > 3 else 25
rows = list(csv.reader(open(raw)))
H = rows[0]
for r in rows[2:]:
    for k in ['gpu__time_duration.sum', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
              'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__thread_inst_executed_per_inst_executed.ratio',
              'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
              'lts__t_sectors.sum', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum']:
        if k in H:
            print(k, r[H.index(k)], rows[1][H.index(k)])
    for i, h in enumerate(H):
        if 'issue_stalled' in h and h.endswith('per_issue_active.ratio'):
            v = float(r[i])
            if v > 0.3:
                print('   stall', h.split('stalled_')[1].split('_per')[0], round(v, 2))
rows = list(csv.reader(open(src)))
H = rows[1]
iS, iSrc, iI = H.index('# Samples'), H.index('Source'), H.index('Instructions Executed')
cols = {h: i for i, h in enumerate(H)}
stalls = [h for h in H if h.startswith('stall_') and 'Not Issued' not in h]
data, seen = [], set()
for r in rows[2:]:
    if len(r) <= iS or r[0] in seen:
        continue
    seen.add(r[0])
    try:
        data.append((int(r[iS]), r[iSrc], int(r[iI]), r))
    except ValueError:
        pass
tot = sum(d[0] for d in data)
print('samples', tot, 'sass lines', len(data), 'warp-inst', sum(d[2] for d in data))
agg = collections.Counter()
for s, _, _, r in data:
    for h in stalls:
        agg[h] += int(r[cols[h]] or 0)
print([(k, v) for k, v in agg.most_common(9)])
for s, srcl, inst, r in sorted(data, key=lambda d: -d[0])[:top_n]:
    top = sorted(((int(r[cols[h]] or 0), h) for h in stalls), reverse=True)[:2]
    print(f"{s:6d} {100*s/tot:5.1f}% inst={inst:8d} {srcl[:58]:58s} {top}")
