#!/bin/bash
# Kernel iteration on ONE GPU: the GPU suite, then the timings that matter (cfg2 bench, cfg5 at poisson_timestep 1/10/100).
O=gpurun_out; mkdir -p $O
if [ "${SKIP_TESTS:-0}" != "1" ]; then timeout 1300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log | cut -c1-300; fi
for i in 1 2 3; do PICB200_BENCH_NO_CPU=1 PICB200_BENCH_NO_CFG4=1 PICB200_BENCH_NO_DROPIN=1 python bench.py --steps 5 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('cfg2 ms/step %.3f value %.4g e2e %.4g' % (d['ms_per_step'], d['value'], d['e2e']['value']))"; done
for a in "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 100" "30 0 100000000 10 256 160000000 100 10"; do
  for i in 1 2; do echo "$a :: $(./run $a 2>&1 | grep -E 'GPU time' | tr '\n' ' ')"; done; done
[ -n "${EXTRA:-}" ] && eval "$EXTRA"
true
