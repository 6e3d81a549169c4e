#!/bin/bash
# Round-2 run K: the final tree on one B200 -- full GPU suite, smoke() (fast path and exact mode against the oracle),
# the default bench line and the reference arm.
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -q > $OUT/r02k_gpu_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02k_gpu_tests.log; tail -4 $OUT/r02k_gpu_tests.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/r02k_smoke.log 2>&1; echo "smoke rc=$?" >> $OUT/r02k_smoke.log; tail -3 $OUT/r02k_smoke.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $OUT/r02k_bench_ref.json 2> $OUT/r02k_bench_ref.err; echo "ref rc=$?"; tail -1 $OUT/r02k_bench_ref.json
timeout 600 python bench.py > $OUT/r02k_bench.json 2> $OUT/r02k_bench.err; echo "bench rc=$?"; tail -1 $OUT/r02k_bench.json
