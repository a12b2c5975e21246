#!/bin/bash
# 2 GPUs: where does the exchange's time go?  Phase traces with the own / the remote bulk adds switched off (timing only).
O=gpurun_out; mkdir -p $O
N=${1:-2}
for dbg in 0 1 2 3; do
  PICB200_PEER_DEBUG=$dbg PICB200_TRACE=$O/trace_dbg$dbg.bin PICB200_GPUS=$N timeout 300 ./run 30 0 $((1000000*N)) 40 256 $((8000000*N)) 100 10 > $O/trace_run_dbg$dbg.log 2>&1; grep -E "GPU time" $O/trace_run_dbg$dbg.log
  echo "== PICB200_PEER_DEBUG=$dbg"; python tools/trace_summary.py $O/trace_dbg$dbg.bin.0 | grep -E "step time|no compaction" | tee $O/trace_dbg${dbg}_rank0.txt
done
timeout 600 python -m pytest tests/test_multi_gpu.py tests/test_harness_modes.py tests/test_golden.py -m gpu -q 2>&1 | tail -5
