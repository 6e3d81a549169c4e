#!/bin/bash
# Round-2 run J: reference-exact verification mode -- bit-for-bit tests, then its speed against the fast kernels.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_exact.py -m gpu -q -x > $OUT/r02j_exact_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02j_exact_tests.log; tail -5 $OUT/r02j_exact_tests.log
{
echo "== fast kernels (planned)"; nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 1000,6000,16384,50000,100000,262144 --steps 10 --warmup 2
for b in avx sse scalar; do echo "== NBODY_EXACT=$b"; NBODY_EXACT=$b nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 1000,6000,16384,50000,100000,262144 --steps 10 --warmup 2; done
echo "== NBODY_EXACT=avx N=1M"; NBODY_EXACT=avx nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 1048576 --steps 2 --warmup 1
} > $OUT/r02j_exact_speed.txt 2>&1; cat $OUT/r02j_exact_speed.txt
