#!/bin/bash
# Round-2 run O: the cluster-resident multi-step kernel of tiny Worlds -- bit-identity, then nbody-bench rows.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "resident or graph or persistent" > $OUT/r02o_resident_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02o_resident_tests.log; tail -5 $OUT/r02o_resident_tests.log
B=nbody_b200/bin/nbody-bench
SIZES=250,500,800,1000,1200,1536,2000,3000,4000
{
echo "== resident cluster kernel (planned up to N = 640; here forced up to 4000)"; NBODY_RESIDENT_MAX_N=4000 timeout 120 $B --gpu --fresh-seed --n $SIZES
echo "== CUDA-graph chunks (NBODY_NO_RESIDENT=1)"; NBODY_NO_RESIDENT=1 timeout 120 $B --gpu --fresh-seed --n $SIZES
echo "== launch by launch (NBODY_NO_RESIDENT=1 NBODY_NO_GRAPH=1)"; NBODY_NO_RESIDENT=1 NBODY_NO_GRAPH=1 timeout 120 $B --gpu --fresh-seed --n $SIZES
echo "== default plan, the reference's sizes"; timeout 120 $B --gpu
} > $OUT/r02o_resident_bench.txt 2>&1; cat $OUT/r02o_resident_bench.txt
