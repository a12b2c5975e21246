#!/bin/bash
# modes 32/31 at cfg1 / cfg2 / cfg5 sizes: per-cell field from 4 (default) / 2 / never, CLI `GPU time of program`
for fr in default 2 -1; do for a in "32 0 10000 10 256 1000000 100 10" "32 0 1000000 100 256 100000000 100 10" "31 0 1000000 100 256 100000000 100 10" "32 0 10000000 10 256 100000000 100 1"; do
  if [ $fr = default ]; then e=""; else e="PICB200_FIELD_RATIO=$fr"; fi
  echo "field ratio $fr :: $a :: $(env $e ./run $a 2>&1 | grep -E "GPU time" | tr "\n" " ") $(env $e ./run $a 2>&1 | grep -E "GPU time" | tr "\n" " ")"
done; done
