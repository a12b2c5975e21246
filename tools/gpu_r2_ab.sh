#!/bin/bash
# One box: GPU suite on the in-tree build, then in-tree build against ab/base (the previous commit's library), interleaved.
O=gpurun_out; mkdir -p $O
if [ "${SKIP_TESTS:-0}" != "1" ]; then timeout 1300 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_gpu.log | cut -c1-400; fi
REPS=${REPS:-4} bash tools/gpu_ab_libs.sh "base|ab/base|" "new|-|" ${AB_EXTRA:-}
