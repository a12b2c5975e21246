#!/bin/bash
# After tools/gpu_round.sh: copy the round's evidence from gpurun_out/ into profiles/ (tracked).  usage: refresh_profiles.sh r01
R=${1:-r01}; H=$(git rev-parse --short HEAD)
ncu -i gpurun_out/prof_fused.ncu-rep --page raw --csv > /tmp/raw.csv 2>/dev/null
ncu -i gpurun_out/prof_fused.ncu-rep --page source --csv > /tmp/src.csv 2>/dev/null
cp gpurun_out/launches.csv profiles/${R}_launches_bench.csv
{ echo "# ncu --set full capture of k_poisson_fused, round ${R#r} (commit $H): tools/prof_run.py 0 1000000 10 10 4000000 30"
  echo "# = 10 Poisson steps x 10 mobility steps over 1e6 electrons after 30 warm Poisson steps (collisions on), ~1e8 pushes in this launch"
  python tools/ncu_summary.py /tmp/raw.csv /tmp/src.csv 40 | cut -c1-200; } > profiles/${R}_ncu_k_poisson_fused.txt
rm -f profiles/${R}_bench_???????.json profiles/${R}_bench_reference_???????.json
tail -1 gpurun_out/bench.log > profiles/${R}_bench_$H.json
tail -1 gpurun_out/bench_ref.log > profiles/${R}_bench_reference_$H.json
python - "$R" "$H" <<'PY'
import csv, json, re, sys
R, H = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open('/tmp/raw.csv'))); Hd = rows[0]; units = rows[1]; r = rows[2]
def val(k):
    v = float(r[Hd.index(k)]); u = units[Hd.index(k)]
    return v * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}.get(u, 1)
m = re.search(r'pushes_in_profiled_launch (\d+)', open('gpurun_out/plain_prof.log').read())
out = {'commit': H, 'kernel': 'k_poisson_fused', 'command': 'tools/prof_run.py 0 1000000 10 10 4000000 30',
       'dram_bytes_read': val('dram__bytes_read.sum'), 'dram_bytes_write': val('dram__bytes_write.sum'),
       'pushes_in_launch': int(m.group(1)) if m else None, 'warp_instructions': float(r[Hd.index('smsp__inst_executed.sum')]), 'issue_active_pct': float(r[Hd.index('smsp__issue_active.avg.pct_of_peak_sustained_active')]), 'registers_per_thread': float(r[Hd.index('launch__registers_per_thread')]), 'duration_ms': float(r[Hd.index('gpu__time_duration.sum')]) * {'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3}.get(units[Hd.index('gpu__time_duration.sum')], 1.0)}
json.dump(out, open(f'profiles/{R}_ncu_traffic.json', 'w'), indent=1)
print(out)
PY
head -12 profiles/${R}_ncu_k_poisson_fused.txt
