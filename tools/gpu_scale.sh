#!/bin/bash
# One 8-GPU box: multi-GPU parity tests once, then the weak-scaling bench at N = 8, 4, 2, 1 and cfg4 strong scaling.
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
for N in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --steps 5 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"
done
PICB200_BENCH_NO_CPU=1 timeout 600 python bench.py --steps 5 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench N=1 rc=$?"
python - <<PY
import json
for N in (1, 2, 4, 8):
    try:
        d = json.loads(open("$O/bench_n%d.json" % N).read().strip().splitlines()[-1])
        print("N=%d value %.4g ms/step %.3f e2e %.4g" % (d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"]))
    except Exception as e:
        print("N=%d: no bench line" % N, e)
PY
[ "${CFG4:-1}" = "1" ] && bash tools/gpu_cfg4.sh 8 4 2 1
