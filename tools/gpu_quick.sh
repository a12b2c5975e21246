#!/bin/bash
# Quick iteration cycle on the GPU box: parity tests, phase trace of the fused kernel, short bench.
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
PICB200_TRACE=$O/trace.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/trace_run.log 2>&1; python tools/trace_summary.py $O/trace.bin
python bench.py --steps 5 --warmup 3 2>$O/bench_quick.err | tee $O/bench_quick.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('value %.4g  ms/step %.3f  e2e %.4g  frac %.3f  launches %d' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches'])); print(d.get('cpu_baseline'))"
