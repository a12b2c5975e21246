#!/bin/bash
# 2 GPUs: bench.py under torchrun with the NVSwitch set-up (a) working, (b) failing on rank 1 -> agreed fall-back to the unicast path
O=gpurun_out; mkdir -p $O
[ "${SKIP_TESTS:-0}" = 1 ] || timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q 2>&1 | tail -3
for tf in none 1; do
  if [ $tf = none ]; then unset PICB200_NVLS_TEST_FAIL; else export PICB200_NVLS_TEST_FAIL=$tf; fi
  PICB200_BENCH_NO_CFG4=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 5 --warmup 3 2> $O/bench_fb.err | tail -1 > $O/bench_fb.json
  python - <<PY
import json
try:
    d=json.loads(open("$O/bench_fb.json").read())
    p=d.get("parity_check") or {}
    print("TEST_FAIL='$tf': value %.4g ms/step %.3f parity %s | %s | %s" % (d["value"], d["ms_per_step"], p.get("ok"), p.get("transport_detail"), p.get("transport_note")))
except Exception as e:
    print("no line", e); print(open("$O/bench_fb.err").read()[-1500:])
PY
done
