#!/bin/bash
# Round 2 evidence session on ONE GPU: both bench arms, the ncu launch list of the bench command, full ncu captures of the
# dominant kernel at cfg2 and on the memory-side corner (cfg5, poisson_timestep = 1), in-kernel phase traces, and every
# BASELINE config next to the reference's CUDA program.  Everything lands in gpurun_out/ (tools/refresh_profiles_r2.sh
# copies the summaries into profiles/).
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt; lscpu | grep -E 'Model name|^CPU\(s\)' >> $O/gpu.txt
if [ "${SKIP_TESTS:-0}" != "1" ]; then timeout 1300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_gpu.log | cut -c1-200; fi
timeout 900 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench rc=$?"; tail -c 600 $O/bench.err
timeout 600 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.log 2> $O/bench_ref.err; echo "bench ref rc=$?"
python - <<PY
import json
d=json.loads(open("$O/bench.log").read().strip().splitlines()[-1]); r=json.loads(open("$O/bench_ref.log").read().strip().splitlines()[-1])
print("ours value %.4g ms %.3f e2e %.4g issue_frac %s dropin %.4g | hbm corner %s | cfg4 %s | reference %.4g (%.1f ms) -> x%.1f device, x%.1f e2e" % (d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"].get("issue_frac"), (d.get("e2e_dropin") or {}).get("value", 0), (d.get("roofline_hbm_corner") or {}).get("frac"), (d.get("cfg4") or {}).get("ms_per_run"), r["value"], r["ms_per_step"], d["value"]/r["value"], d["e2e"]["value"]/r["value"]))
PY
# ---- ncu: launch list of the bench command (after the plain run above exited 0), then full captures
PICB200_BENCH_NO_CPU=1 PICB200_BENCH_NO_CFG4=1 PICB200_BENCH_NO_DROPIN=1 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_launches.log 2>&1; echo "ncu launches rc=$?"
export PICB200_NO_WARM=1   # (the capture indices below count the launches of prof_run.py, not the zero-step warm-up launch of picb200_create)
timeout 300 python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/plain_prof.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -s 1 -c 1 -f -o $O/prof_fused python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/ncu_full.log 2>&1; echo "ncu full cfg2 rc=$?"
timeout 300 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/plain_prof_pt1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -c 1 -f -o $O/prof_pt1 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/ncu_pt1.log 2>&1; echo "ncu full pt1 rc=$?"
unset PICB200_NO_WARM
# ---- phase traces
PICB200_TRACE=$O/trace.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/trace_run.log 2>&1; python tools/trace_summary.py $O/trace.bin > $O/trace_cfg2.txt 2>&1; head -6 $O/trace_cfg2.txt
PICB200_TRACE=$O/trace_pt1.bin python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/trace_pt1_run.log 2>&1; python tools/trace_summary.py $O/trace_pt1.bin > $O/trace_pt1.txt 2>&1; head -6 $O/trace_pt1.txt
# ---- every BASELINE config, engine and reference
if [ "${SKIP_CONFIGS:-0}" != "1" ]; then timeout 1500 python tools/measure_configs.py > $O/configs.md 2> $O/configs.err; echo "configs rc=$?"; grep "^|" $O/configs.md | cut -c1-230; fi
ls -la $O | head -40
