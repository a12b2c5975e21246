#!/bin/bash
# A/B of library builds and environment knobs on ONE box (boxes differ by a few per cent).
#   tools/gpu_ab_libs.sh "<label>|<lib dir or ->|<ENV=val ...>" ...   (lib dir "-" = the in-tree build)
# For every variant: best and median "GPU time of program" over REPS runs of each CLI configuration,
# variants interleaved.  Configurations: CFGS (semicolon-separated pic_run argument lists).
O=gpurun_out; mkdir -p $O
REPS=${REPS:-5}
CFGS=${CFGS:-"30 0 1000000 100 256 100000000 100 10;30 0 10000000 10 256 100000000 100 10;30 0 10000000 10 256 100000000 100 100;30 0 10000000 10 256 100000000 100 1"}
RUN=$PWD/particle-simulation_b200/pic_run
IFS=';' read -ra CFG <<< "$CFGS"
for c in "${CFG[@]}"; do
  echo "== $c"
  for r in $(seq $REPS); do
    for v in "$@"; do
      IFS='|' read -r label lib envs <<< "$v"
      if [ "$lib" = "-" ]; then L=""; else L="LD_LIBRARY_PATH=$PWD/$lib"; fi
      out=$(env $L $envs $RUN $c 2>&1)
      ms=$(echo "$out" | grep -i "GPU time of program" | grep -o "[0-9.]*" | head -1)
      fin=$(echo "$out" | grep -i "Final amount" | grep -o "[0-9]*" | head -1)
      echo "$label $ms $fin"
    done
  done | python3 -c "
import sys, collections, statistics
d = collections.OrderedDict(); fin = {}
for l in sys.stdin:
    p = l.split()
    if len(p) < 2: continue
    try: d.setdefault(p[0], []).append(float(p[1]))
    except ValueError: continue
    fin[p[0]] = p[2] if len(p) > 2 else '?'
for k, v in d.items():
    print('   %-14s best %9.3f  median %9.3f ms  (final n %s, %d runs)' % (k, min(v), statistics.median(v), fin[k], len(v)))
"
done
