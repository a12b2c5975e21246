#!/bin/bash
# One GPU-box session: parity tests, both bench arms, the ncu launch list of the bench command and
# one full ncu capture of the dominant kernel.  Everything lands in gpurun_out/.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi -L > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt; lscpu | grep -E 'Model name|^CPU\(s\)' >> $O/gpu.txt
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/pytest_gpu.log
timeout 600 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench rc=$?"; tail -c 3000 $O/bench.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.log 2> $O/bench_ref.err; echo "bench ref rc=$?"; tail -c 1500 $O/bench_ref.log
if [ "${SKIP_NCU:-0}" != "1" ]; then
timeout 300 python bench.py --steps 2 --warmup 3 > $O/plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 > $O/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 300 python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/plain_prof.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -s 1 -c 1 -f -o $O/prof_fused python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/ncu_full.log 2>&1
echo "ncu full rc=$?"
fi
ls -la $O
