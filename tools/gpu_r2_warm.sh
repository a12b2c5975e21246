#!/bin/bash
# CLI `GPU time of program` with and without the kernel warm-up of picb200_create (PICB200_NO_WARM=1), after the GPU suite
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for w in 0 1; do for a in "30 0 10000 10 256 1000000 100 10" "33 0 10000 10 256 1000000 100 10" "30 0 1000000 100 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 1" "30 0 10000000 10 256 100000000 100 10" "30 0 10000000 10 256 100000000 100 100" "32 0 10000 10 256 1000000 100 10"; do
  if [ $w = 0 ]; then e="PICB200_NO_WARM=1"; else e="X=1"; fi
  echo "warm=$w :: $a :: $(env $e ./run $a 2>&1 | grep -E "GPU time" | tr "\n" " ") $(env $e ./run $a 2>&1 | grep -E "GPU time" | tr "\n" " ")"
done; done
