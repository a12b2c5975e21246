#!/bin/bash
# the collision-heavy corner of BASELINE cfg5 (pt = 100: 2.8 children per electron and Poisson step) + cfg2, this engine only
for a in "30 0 10000000 10 256 100000000 100 100" "31 0 10000000 10 256 100000000 100 100" "30 0 10000000 10 256 100000000 100 10" "30 0 1000000 100 256 100000000 100 10"; do
  for i in 1 2; do echo "$a :: $(PICB200_PRINT_PUSHES=1 ./run $a 2>&1 | grep -E 'GPU time|pushes|Final' | tr '\n' ' ')"; done
done
