#!/bin/bash
# Round-2 run F: the small-World kernel -- parity tests, then nbody-bench against the (tile, slab)-per-CTA kernel.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_world or ragged or fixture or closed_form or massless or chained" > $OUT/r02f_small_tests.log 2>&1
echo "pytest rc=$?" >> $OUT/r02f_small_tests.log; tail -4 $OUT/r02f_small_tests.log
SIZES=250,500,800,1000,1200,1600,2000,3000,4000,5000,6000,8000,10000,12000,16384
{
echo "== step_kernel (NBODY_SMALL_MAX_N=0)"; NBODY_SMALL_MAX_N=0 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $SIZES
for v in w16x1 w16x2 w8x1; do
  echo "== small_world_kernel $v (NBODY_VARIANT)"; NBODY_VARIANT=$v nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $SIZES
done
echo "== planned"; nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $SIZES
} > $OUT/r02f_small_world_kernel.txt 2>&1
cat $OUT/r02f_small_world_kernel.txt
