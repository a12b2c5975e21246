#!/bin/bash
# 1 GPU: the whole GPU suite with the quota balancer, then A/B (PICB200_BALANCE=0/1) on every corner and the bench.
O=gpurun_out; mkdir -p $O
[ "${SKIP_TESTS:-0}" = "1" ] || { timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log | cut -c1-300; }
for bal in 1 0 1 0; do
  for a in "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 100" "30 0 1000000 100 256 100000000 100 10" "30 0 100000000 10 256 200000000 100 10"; do
    echo "BALANCE=$bal $a :: $(PICB200_BALANCE=$bal ./run $a 2>&1 | grep -E 'GPU time' | tr '\n' ' ')"; done; done 2>&1 | tee $O/ab_balance.txt
PICB200_TRACE=$O/trace_pt1.bin python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/trace_pt1_run.log 2>&1; python tools/trace_summary.py $O/trace_pt1.bin | head -5 | tee $O/trace_pt1.txt
PICB200_TRACE=$O/trace.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/trace_run.log 2>&1; python tools/trace_summary.py $O/trace.bin | tee $O/trace_cfg2.txt
PICB200_BENCH_NO_CPU=1 PICB200_BENCH_NO_CFG4=1 timeout 600 python bench.py --steps 5 > $O/bench_quick.log 2> $O/bench_quick.err; python - <<PY
import json
d=json.loads(open("$O/bench_quick.log").read().strip().splitlines()[-1]); print("value %.4g ms %.3f e2e %.4g" % (d["value"], d["ms_per_step"], d["e2e"]["value"]))
PY
