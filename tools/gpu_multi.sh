#!/bin/bash
# Multi-GPU session (gpurun --gpus N): parity of the slabs, then the bench under torchrun.
O=gpurun_out; mkdir -p $O
N=${1:-2}
nvidia-smi -L | head -8
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -15
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"
tail -c 600 $O/bench_n$N.err; python - <<PY
import json
try:
    d=json.loads(open("$O/bench_n$N.json").read().strip().splitlines()[-1])
    print("N=%d value %.4g ms/step %.3f e2e %.4g" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"]), d["config"]["parallelism"])
except Exception as e: print("no bench line", e)
PY
if [ "${NCCL_TOO:-0}" = "1" ]; then
PICB200_NCCL=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus $N --steps 5 --warmup 3 2>$O/bench_nccl_n$N.err | tee $O/bench_nccl_n$N.json | cut -c1-200
fi
