#!/bin/bash
# Round 2, first GPU-box session: fixtures from the reference binary, the whole GPU suite, the bench line, a sanitizer
# attempt and the ncu capture of the memory-side corner at HEAD.  Everything lands in gpurun_out/.
set -u
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt
echo "== golden"; timeout 900 python tools/make_golden_from_ref.py dense stats > $O/golden.log 2>&1; echo "golden rc=$?"; tail -8 $O/golden.log
cp $O/golden/*.npz $O/golden/*.json tests/golden/ 2>/dev/null
echo "== pytest"; timeout 1200 python -m pytest tests -m gpu -q -x -s > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 $O/pytest_gpu.log; grep -h "TVD engine" $O/pytest_gpu.log
echo "== bench"; timeout 900 python bench.py > $O/bench.log 2> $O/bench.err; echo "bench rc=$?"; tail -c 5000 $O/bench.log; tail -c 800 $O/bench.err
echo "== sanitizer"
for m in 30 31 32 33; do
  timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 ./run $m 0 20000 3 256 200000 100 10 > $O/sanitizer_memcheck_$m.log 2>&1; echo "memcheck mode $m rc=$?"; tail -4 $O/sanitizer_memcheck_$m.log
done
PICB200_TABLE=$PWD/$O/hot.txt; python -c "open('$O/hot.txt','w').write('2.0 1.5\n'*10000)"
PICB200_TABLE=$PWD/$O/hot.txt timeout 300 compute-sanitizer --tool memcheck --error-exitcode 9 ./run 30 0 20000 3 256 200000 100 10 > $O/sanitizer_memcheck_30_hot.log 2>&1; echo "memcheck hot rc=$?"; tail -4 $O/sanitizer_memcheck_30_hot.log
PICB200_TABLE=$PWD/$O/hot.txt timeout 400 compute-sanitizer --tool racecheck --error-exitcode 9 ./run 30 0 20000 2 256 200000 100 10 > $O/sanitizer_racecheck_30_hot.log 2>&1; echo "racecheck hot rc=$?"; tail -6 $O/sanitizer_racecheck_30_hot.log
PICB200_TABLE=$PWD/$O/hot.txt timeout 400 compute-sanitizer --tool initcheck --error-exitcode 9 ./run 30 0 20000 2 256 200000 100 10 > $O/sanitizer_initcheck_30_hot.log 2>&1; echo "initcheck hot rc=$?"; tail -6 $O/sanitizer_initcheck_30_hot.log
echo "== configs"
for a in "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 100" "31 0 10000000 10 256 100000000 100 10" "32 0 10000000 10 256 100000000 100 10" "30 0 1000000 100 256 100000000 100 10"; do
  for i in 1 2; do echo "$a :: $(PICB200_PRINT_PUSHES=1 ./run $a 2>&1 | grep -E 'GPU time|pushes|Final' | tr '\n' ' ')"; done
done 2>&1 | tee $O/configs_head.txt
echo "== ncu pt=1"
timeout 300 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/plain_prof_pt1.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_poisson_fused -c 1 -f -o $O/prof_pt1 python tools/prof_run.py 0 10000000 10 1 100000000 0 > $O/ncu_pt1.log 2>&1
echo "ncu pt1 rc=$?"; tail -3 $O/plain_prof_pt1.log
echo "== trace cfg2"
PICB200_TRACE=$O/trace.bin python tools/prof_run.py 0 1000000 10 10 4000000 30 > $O/trace_run.log 2>&1; python tools/trace_summary.py $O/trace.bin | tee $O/trace_cfg2.txt
ls -la $O
