#!/bin/bash
# phase breakdown (CUDA-event spans) of the host-synchronised modes at cfg5 sizes
for m in 2 1; do for pt in 1 10; do
  echo "== mode $m pt $pt"; python tools/prof_run.py $m 10000000 10 $pt 100000000 0 2>&1 | tail -2 | cut -c1-400
done; done
echo "== field off"; PICB200_FIELD_RATIO=-1 python tools/prof_run.py 2 10000000 10 1 100000000 0 2>&1 | tail -1 | cut -c1-300
