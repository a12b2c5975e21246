#!/bin/bash
O=gpurun_out; mkdir -p $O
for wf in 1 0 1 0; do
  for a in "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 1000000 100 256 100000000 100 10"; do
    echo "WARP_FENCE=$wf $a :: $(PICB200_WARP_FENCE=$wf ./run $a 2>&1 | grep -E 'GPU time' | tr '\n' ' ')"; done; done 2>&1 | tee $O/ab_warp_fence.txt
for wf in 1 0; do PICB200_WARP_FENCE=$wf PICB200_TRACE=$O/trace_wf$wf.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/trace_run.log 2>&1; echo "== WARP_FENCE=$wf"; python tools/trace_summary.py $O/trace_wf$wf.bin 2>/dev/null | head -7; done | tee $O/trace_cfg2_wf.txt
