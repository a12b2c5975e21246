#!/bin/bash
# BASELINE cfg4: 1e9 electrons, 10 Poisson x 10 mobility steps, sharded over N GPUs of one box (strong scaling; N=1 fits one B200)
O=gpurun_out; mkdir -p $O
for N in ${@:-8 4 2 1}; do
  PICB200_PRINT_PUSHES=1 PICB200_GPUS=$N timeout 600 ./run 30 0 1000000000 10 256 1600000000 100 10 > $O/cfg4_n$N.log 2> $O/cfg4_n$N.err
  echo "N=$N rc=$? $(grep -E 'Final|added|removed|GPU time' $O/cfg4_n$N.log | tr '\n' ' ') $(cat $O/cfg4_n$N.err | tail -1)"
done
