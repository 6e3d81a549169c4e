#!/bin/bash
set -u
OUT=gpurun_out
mkdir -p $OUT
for f in 0 1 2 4 3 7 8 15; do
  echo "== NBODY_PERSIST_EXP=$f"
  NBODY_PERSIST_EXP=$f timeout 60 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,1000,2000,4000,8000 | tail -5
done > $OUT/r02c_persist_exp.txt 2>&1
echo "== no persist" >> $OUT/r02c_persist_exp.txt
NBODY_NO_PERSIST=1 timeout 60 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,1000,2000,4000,8000 | tail -5 >> $OUT/r02c_persist_exp.txt
cat $OUT/r02c_persist_exp.txt
