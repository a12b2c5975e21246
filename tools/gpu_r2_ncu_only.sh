#!/bin/bash
# the two full ncu captures of tools/gpu_r2_round.sh on their own
O=gpurun_out; mkdir -p $O
export PICB200_NO_WARM=1
timeout 300 python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/plain_prof.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -s 1 -c 1 -f -o $O/prof_fused python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/ncu_full.log 2>&1; echo "ncu full cfg2 rc=$?"
timeout 300 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/plain_prof_pt1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -c 1 -f -o $O/prof_pt1 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/ncu_pt1.log 2>&1; echo "ncu full pt1 rc=$?"
