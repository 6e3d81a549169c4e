#!/bin/bash
# Round-2 run A: GPU suite with the parity columns, default bench, the GROUPED kernel against the ungrouped one
# at N = 1M (forced with NBODY_PART_BUDGET), and N = 2M / 4M with the planner's own grouped plan.
set -u
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $OUT/r02a_gpu.txt
python -m pytest tests -m gpu -q -s -x > $OUT/r02a_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02a_gpu_tests.log
python bench.py --steps 10 --warmup 3 > $OUT/r02a_bench_n1m.json 2> $OUT/r02a_bench_n1m.err
B="python bench.py --no-cpu-baseline --no-e2e --warmup 3"
$B --steps 6 > $OUT/r02a_n1m_ungrouped.json 2>> $OUT/r02a.err
for g in 1 2 4 8 16; do
  NBODY_PART_BUDGET=$((g * 1048576 * 8 + 8)) $B --steps 6 > $OUT/r02a_n1m_groups$g.json 2>> $OUT/r02a.err
done
$B --particles 2097152 --steps 3 --warmup 1 > $OUT/r02a_n2m.json 2>> $OUT/r02a.err
NBODY_PART_BUDGET=100000000000 $B --particles 2097152 --steps 3 --warmup 1 > $OUT/r02a_n2m_ungrouped.json 2>> $OUT/r02a.err
$B --particles 4194304 --steps 3 --warmup 1 > $OUT/r02a_n4m.json 2>> $OUT/r02a.err
NBODY_PART_BUDGET=100000000000 $B --particles 4194304 --steps 3 --warmup 1 > $OUT/r02a_n4m_ungrouped.json 2>> $OUT/r02a.err
grep -h -o '"ms_per_step": [0-9.]*' $OUT/r02a_n*.json | paste - - - - - - - - - - - 
ls -la $OUT
