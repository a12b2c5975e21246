#!/bin/bash
# 1 GPU: the memory-side corner (cfg5, poisson_timestep = 1) with the streaming instantiation: phase trace, A/B against the
# plain instantiation, and a full ncu capture.
O=gpurun_out; mkdir -p $O
for ss in 3 0; do for i in 1 2; do echo "PICB200_STREAM_STEPS=$ss :: $(PICB200_STREAM_STEPS=$ss PICB200_PRINT_PUSHES=1 ./run 30 0 10000000 10 256 100000000 100 1 2>&1 | grep -E 'GPU time' | tr '\n' ' ')"; done; done
for ss in 3 0; do echo "pt=2 PICB200_STREAM_STEPS=$ss :: $(PICB200_STREAM_STEPS=$ss ./run 30 0 10000000 10 256 100000000 100 2 2>&1 | grep -E 'GPU time' | tr '\n' ' ')"; done
PICB200_TRACE=$O/trace_pt1.bin python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/trace_pt1_run.log 2>&1; python tools/trace_summary.py $O/trace_pt1.bin | tee $O/trace_pt1.txt
timeout 300 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/plain_prof_pt1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -c 1 -f -o $O/prof_pt1 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/ncu_pt1.log 2>&1
echo "ncu pt1 rc=$?"; tail -3 $O/plain_prof_pt1.log
