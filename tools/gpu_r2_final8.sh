#!/bin/bash
# Final 8-GPU session of round 2: NVSwitch probe (push and reduce-scatter forms), multi-GPU parity tests, bench.py --gpus 8/4/2/1.
O=gpurun_out; mkdir -p $O
nvcc -O2 -gencode arch=compute_100a,code=sm_100a -o ab/bin/nvls_probe tools/nvls_probe.cu -lcuda 2>/dev/null; timeout 120 ab/bin/nvls_probe 8 2>&1 | sort | tee $O/nvls_probe_n8.txt | grep -E "rank 0|multicast"
bash tools/gpu_r2_scale8.sh
