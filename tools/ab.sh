#!/bin/bash
# A/B on ONE box (boxes differ by a few per cent): tools/ab.sh VAR valA valB [reps]  -> ms per cfg2 run for each setting, interleaved
VAR=$1; A=$2; B=$3; R=${4:-3}
for i in $(seq $R); do
  for v in $A $B; do
    printf "%s=%s " $VAR $v; env PICB200_BENCH_NO_CPU=1 $VAR=$v python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('ms/step %.3f value %.4g' % (d['ms_per_step'], d['value']))"
  done
done
