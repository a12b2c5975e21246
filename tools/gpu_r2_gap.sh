#!/bin/bash
# Diagnostic session: phase traces of cfg2 with the barrier-approach stamps, default and with the A/B switches
O=gpurun_out; mkdir -p $O
for v in "" "PICB200_WARP_FENCE=0" "PICB200_BALANCE=0"; do
  echo "=== ${v:-default}"
  env $v PICB200_TRACE=$O/trace_gap.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 | tail -2 | cut -c1-200
  python tools/trace_gap.py $O/trace_gap.bin | head -24
  python tools/trace_summary.py $O/trace_gap.bin | head -6
done
bash tools/ab.sh PICB200_WARP_FENCE 1 0 3
bash tools/ab.sh PICB200_BALANCE 1 0 3
