#!/bin/bash
# 1 GPU: fixtures from the reference binary, the whole single-GPU suite, quick bench.
set -u
O=gpurun_out; mkdir -p $O
echo "== golden"; timeout 900 python tools/make_golden_from_ref.py dense stats > $O/golden.log 2>&1; echo "golden rc=$?"; tail -8 $O/golden.log
cp $O/golden/*.npz $O/golden/*.json tests/golden/ 2>/dev/null
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q -s > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $O/pytest_gpu.log | cut -c1-400; grep -h "TVD engine" $O/pytest_gpu.log
echo "== bench"; PICB200_BENCH_NO_CPU=1 PICB200_BENCH_NO_CFG4=1 timeout 600 python bench.py --steps 5 > $O/bench_quick.log 2> $O/bench_quick.err; echo "bench rc=$?"; python - <<PY
import json
d=json.loads(open("$O/bench_quick.log").read().strip().splitlines()[-1]); print("value %.4g ms %.3f e2e %.4g dropin %s" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d.get("e2e_dropin")))
PY
