#!/bin/bash
# N GPUs of one box: multi-GPU parity tests with the NVSwitch reduction (default) and with PICB200_NVLS=0, then
# bench.py --gpus N both ways (weak scaling of cfg2 + self-check), and a phase trace.
O=gpurun_out; mkdir -p $O
N=${1:-2}
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -6
PICB200_NVLS=0 timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
for rep in $(seq ${REPS:-2}); do for v in 1 0; do
  PICB200_NVLS=$v PICB200_BENCH_NO_CFG4=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29530+v)) bench.py --gpus $N --steps 10 --warmup 3 2> $O/bench_nvls${v}_n$N.err | tail -1 > $O/bench_nvls${v}_n$N.json
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_nvls${v}_n$N.json").read().strip().splitlines()[-1])
    print("NVLS=$v N=%d value %.4g ms/step %.3f e2e %.4g parity %s" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], (d.get("parity_check") or {}).get("ok")))
except Exception as e:
    print("NVLS=$v: no bench line", e); print(open("$O/bench_nvls${v}_n$N.err").read()[-1500:])
PY
done; done
for v in 1 0; do
PICB200_NVLS=$v PICB200_TRACE=$O/trace_nvls${v}_n$N.bin PICB200_GPUS=$N timeout 300 ./run 30 0 $((1000000*N)) 40 256 $((8000000*N)) 100 10 > $O/trace_run_nvls${v}_n$N.log 2>&1; tail -2 $O/trace_run_nvls${v}_n$N.log
python tools/trace_summary.py $O/trace_nvls${v}_n$N.bin.0 2>/dev/null | head -10 | tee $O/trace_nvls${v}_n${N}_rank0.txt
done
