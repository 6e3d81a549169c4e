#!/bin/bash
# Round-2 run B: the persistent multi-step kernel (tests first, bounded), nbody-bench at the reference's sizes with
# and without it, and the schedule lottery of the GROUPED kernel.
set -u
OUT=gpurun_out
mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "persistent or chained or large_worlds" > $OUT/r02b_persist_tests.log 2>&1
echo "pytest rc=$?" >> $OUT/r02b_persist_tests.log
tail -3 $OUT/r02b_persist_tests.log
if grep -q "rc=0" $OUT/r02b_persist_tests.log; then
  timeout 600 python -m pytest tests -m gpu -q -x > $OUT/r02b_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02b_gpu_tests.log
  tail -3 $OUT/r02b_gpu_tests.log
  timeout 120 nbody_b200/bin/nbody-bench --gpu > $OUT/r02b_nbody_bench_gpu_persist.txt 2>&1
  NBODY_NO_PERSIST=1 timeout 120 nbody_b200/bin/nbody-bench --gpu > $OUT/r02b_nbody_bench_gpu_nopersist.txt 2>&1
  timeout 120 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,500,1000,2000,3000,4000,6000,8000,10000,12000 > $OUT/r02b_nbody_bench_small_persist.txt 2>&1
  NBODY_NO_PERSIST=1 timeout 120 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,500,1000,2000,3000,4000,6000,8000,10000,12000 > $OUT/r02b_nbody_bench_small_nopersist.txt 2>&1
  paste $OUT/r02b_nbody_bench_small_persist.txt $OUT/r02b_nbody_bench_small_nopersist.txt
fi
TUNE_SET=ship timeout 60 tools/lottery/g_000000 sweep 1048576 523884 37 > $OUT/r02b_tune_ship.txt 2>&1
bash tools/lottery/run_grouped.sh > $OUT/r02b_lottery_grouped.log 2>&1
sort -k2 -n $OUT/lottery_grouped.txt | head -8
grep packed $OUT/r02b_tune_ship.txt
