#!/bin/bash
# Determinism soak: the same run repeated must give byte-identical id-sorted snapshots (slot order is a race, the
# electrons are not).  usage: gpu_soak.sh [reps]
R=${1:-10}; W=/tmp/soak; mkdir -p $W/src; cp src/cross_section.txt $W/src/
cd $W
for cfg in "30 0 1000000 100 256 100000000 100 10" "30 0 2000000 10 256 100000000 100 100" "33 0 300000 40 256 3000000 100 7" \
           "30 0 4000000 10 256 100000000 100 1" "31 0 1000000 20 256 100000000 100 10" "32 0 1000000 20 256 100000000 100 10"; do
  rm -f sums.txt
  for i in $(seq $R); do
    PICB200_DUMP_FINAL=$W/f.bin $GRAFT_REPO_ROOT/particle-simulation_b200/pic_run $cfg > out.txt 2>&1 || { echo "run failed"; tail -3 out.txt; }
    sha256sum f.bin | cut -c1-16 >> sums.txt
  done
  echo "$cfg :: $(sort sums.txt | uniq -c | tr '\n' ' ') $(grep -E 'Final|added' out.txt | tr '\n' ' ')"
done
