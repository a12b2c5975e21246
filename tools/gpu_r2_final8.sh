#!/bin/bash
# Final 8-GPU session of round 2: NVSwitch probe (push and reduce-scatter forms), multi-GPU parity tests, bench.py --gpus 8/4/2/1.
O=gpurun_out; mkdir -p $O
timeout 120 ab/bin/nvls_probe 8 2>&1 | sort | tee $O/nvls_probe_n8.txt | grep -E "rank 0|multicast"
bash tools/gpu_r2_scale8.sh
