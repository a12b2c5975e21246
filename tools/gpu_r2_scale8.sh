#!/bin/bash
# One 8-GPU box: multi-GPU parity tests (k = 2, 4, 8), then bench.py --gpus 8 / 4 / 2 / 1 (weak scaling of cfg2 + the cfg4 sub-record).
O=gpurun_out; mkdir -p $O
nvidia-smi -L | wc -l
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -q 2>&1 | tail -4
for N in 8 4 2; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29520+N)) bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "bench N=$N rc=$?"; tail -c 400 $O/bench_n$N.err | grep -v OMP_NUM | tail -3
done
PICB200_BENCH_NO_CPU=1 timeout 900 python bench.py --steps 10 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench N=1 rc=$?"
python - <<PY
import json
v1=None
for N in (1, 2, 4, 8):
    try:
        d = json.loads(open("$O/bench_n%d.json" % N).read().strip().splitlines()[-1])
        if N == 1: v1 = d["value"]
        c = d.get("cfg4") or {}
        print("N=%d value %.4g ms/step %.3f eff %.3f e2e %.4g parity %s | cfg4 %.4g pushes/s %.1f ms" % (d["n_gpus"], d["value"], d["ms_per_step"], d["value"]/(N*v1) if v1 else 0, d["e2e"]["value"], (d.get("parity_check") or {}).get("ok"), c.get("value", 0), c.get("ms_per_run", 0)))
    except Exception as e:
        print("N=%d: no bench line" % N, e)
PY
PICB200_TRACE=$O/trace_n8.bin PICB200_GPUS=8 timeout 300 ./run 30 0 8000000 40 256 64000000 100 10 > $O/trace_run_n8.log 2>&1; tail -2 $O/trace_run_n8.log
python tools/trace_summary.py $O/trace_n8.bin.0 2>/dev/null | head -12 | tee $O/trace_n8_rank0.txt
