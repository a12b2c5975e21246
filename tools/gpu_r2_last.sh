#!/bin/bash
# last session of round 2: the GPU suite, smoke(), both bench arms and every BASELINE config at the final commit
O=gpurun_out; mkdir -p $O
timeout 1300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 $O/pytest_gpu.log | cut -c1-200
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 900 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench rc=$?"; tail -c 300 $O/bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.log 2> $O/bench_ref.err; echo "bench ref rc=$?"
python - <<PY
import json
d=json.loads(open("$O/bench.log").read().strip().splitlines()[-1]); r=json.loads(open("$O/bench_ref.log").read().strip().splitlines()[-1])
print("ours value %.4g ms %.3f e2e %.4g issue_frac %s dropin %.4g | hbm corner %s | cfg4 %s | reference %.4g (%.1f ms) -> x%.1f device, x%.1f e2e" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"].get("issue_frac"), (d.get("e2e_dropin") or {}).get("value", 0), (d.get("roofline_hbm_corner") or {}).get("frac"), (d.get("cfg4") or {}).get("ms_per_run"), r["value"], r["ms_per_step"], d["value"]/r["value"], d["e2e"]["value"]/r["value"]))
PY
timeout 1500 python tools/measure_configs.py > $O/configs.md 2> $O/configs.err; echo "configs rc=$?"; grep "^|" $O/configs.md | cut -c1-200
