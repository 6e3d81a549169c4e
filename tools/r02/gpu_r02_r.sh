#!/bin/bash
# Round-2 run R: the cluster-split small-World kernel -- bit-identity, then nbody-bench with and without the split.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "cluster_split or graph or fixture" > $OUT/r02r_split_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02r_split_tests.log; tail -4 $OUT/r02r_split_tests.log
B=nbody_b200/bin/nbody-bench
SIZES=250,500,700,800,1000,1200,1600,2000,2368,2400,3000,4000
{
echo "== slabs of a tile over a cluster of 2..4 CTAs (default); chained steps as CUDA-graph chunks (NBODY_NO_RESIDENT=1)"; NBODY_NO_RESIDENT=1 timeout 120 $B --gpu --fresh-seed --n $SIZES
echo "== one CTA per tile (NBODY_NO_SPLIT=1 NBODY_NO_RESIDENT=1)"; NBODY_NO_SPLIT=1 NBODY_NO_RESIDENT=1 timeout 120 $B --gpu --fresh-seed --n $SIZES
echo "== default plan, the reference's sizes"; timeout 120 $B --gpu
echo "== single synchronous steps (--steps 1 per call x 200 calls is not a CLI mode: 100 chained steps above)"
} > $OUT/r02r_split_bench.txt 2>&1; cat $OUT/r02r_split_bench.txt
