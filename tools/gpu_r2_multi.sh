#!/bin/bash
# N GPUs of one box: the whole GPU suite (multi-GPU parity included), bench under torchrun (with its self-check),
# phase trace of the peer exchange, single-GPU bench on the same box.
O=gpurun_out; mkdir -p $O
N=${1:-2}
nvidia-smi -L | head -8
if [ "${FULL_SUITE:-1}" = "1" ]; then timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu_n$N.log 2>&1; echo "pytest rc=$?"; tail -6 $O/pytest_gpu_n$N.log | cut -c1-300
else timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -15; fi
PICB200_BENCH_NO_CFG4=${NO_CFG4:-0} timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 1500 $O/bench_n$N.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n$N.json").read().strip().splitlines()[-1])
    print("N=%d value %.4g ms/step %.3f e2e %.4g" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"]), d.get("parity_check"), d.get("cfg4"))
except Exception as e: print("no bench line", e)
PY
# phase trace: cfg2-shaped slabs, 10 Poisson steps after 30 warm ones
PICB200_TRACE=$O/trace_n$N.bin PICB200_GPUS=$N timeout 300 ./run 30 0 $((1000000*N)) 40 256 $((8000000*N)) 100 10 > $O/trace_run_n$N.log 2>&1; tail -3 $O/trace_run_n$N.log
for r in 0 1; do [ -f $O/trace_n$N.bin.$r ] && python tools/trace_summary.py $O/trace_n$N.bin.$r | tee $O/trace_n${N}_rank$r.txt; done
PICB200_BENCH_NO_CPU=1 PICB200_BENCH_NO_CFG4=1 timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; python - <<PY
import json
d=json.loads(open("$O/bench_n1.json").read().strip().splitlines()[-1]); print("N=1 value %.4g ms/step %.3f" % (d["value"], d["ms_per_step"]))
PY
for a in "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 100"; do
  echo "$a :: $(PICB200_PRINT_PUSHES=1 ./run $a 2>&1 | grep -E 'GPU time|pushes' | tr '\n' ' ')"; done
if [ "${FENCE_AB:-0}" = "1" ]; then
PICB200_PEER_FENCE=sys PICB200_BENCH_NO_CFG4=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 5 --warmup 3 2> $O/bench_sys_n$N.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('PICB200_PEER_FENCE=sys: N=%d value %.4g ms/step %.3f' % (d['n_gpus'], d['value'], d['ms_per_step']), d['parity_check']['ok'])"
fi
