#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -q -s > $OUT/r02h_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02h_gpu_tests.log
tail -5 $OUT/r02h_gpu_tests.log
nbody_b200/bin/nbody-bench --gpu > $OUT/r02h_nbody_bench_gpu.txt 2>&1; cat $OUT/r02h_nbody_bench_gpu.txt
NBODY_SMALL_MAX_N=0 nbody_b200/bin/nbody-bench --gpu > $OUT/r02h_nbody_bench_gpu_stepkernel_only.txt 2>&1
python tools/bench_configs.py c2 > $OUT/r02h_config2.jsonl 2>&1; cat $OUT/r02h_config2.jsonl
NBODY_SMALL_MAX_N=0 python tools/bench_configs.py c2 >> $OUT/r02h_config2.jsonl 2>&1; tail -1 $OUT/r02h_config2.jsonl
python tools/bench_configs.py ladder > $OUT/r02h_ladder.jsonl 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > $OUT/r02h_smoke.log 2>&1; tail -2 $OUT/r02h_smoke.log
