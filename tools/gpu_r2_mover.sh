#!/bin/bash
# windowed mover launches and grid-phase kernels (modes 31/32): the GPU suite, then CLI A/B of the in-tree build against ab/base
O=gpurun_out; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
REPS=${REPS:-4} CFGS="32 0 1000000 100 256 100000000 100 10;31 0 1000000 100 256 100000000 100 10;32 0 10000000 10 256 100000000 100 10;31 0 10000000 10 256 100000000 100 10;32 0 10000000 10 256 100000000 100 1;31 0 10000000 10 256 100000000 100 1;32 0 10000 10 256 1000000 100 10;31 0 10000 10 256 1000000 100 10;30 0 1000000 100 256 100000000 100 10;30 0 10000 10 256 1000000 100 10" \
  bash tools/gpu_ab_libs.sh "base|ab/base|" "new|-|"
