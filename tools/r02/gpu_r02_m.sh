#!/bin/bash
# Round-2 run M: which part of balanced kernel v2 costs what (MODE switches), three sizes.
set -u
OUT=gpurun_out; mkdir -p $OUT
B=nbody_b200/bin/nbody-bench
{
for v in shipped b16x2 b16x2m1 b16x2m2 b16x2m4 b16x2m7 b16x4 b16x4m7 b32x1; do
  echo "== $v"
  for n in 4736 10000 16384 20000; do
    if [ $v = shipped ]; then timeout 60 $B --gpu --fresh-seed --n $n | tail -1; else NBODY_VARIANT=$v timeout 60 $B --gpu --fresh-seed --n $n 2>&1 | tail -1; fi
  done
done
} > $OUT/r02m_balanced_modes.txt 2>&1; cat $OUT/r02m_balanced_modes.txt
