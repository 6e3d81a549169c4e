#!/bin/bash
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "graph or chained or small_world" > $OUT/r02i_graph_tests.log 2>&1; echo "pytest rc=$?" >> $OUT/r02i_graph_tests.log; tail -4 $OUT/r02i_graph_tests.log
{
echo "== graph chunks (default)"; nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,500,800,1000,1200,1600,2000,3000,4000,4736
echo "== NBODY_NO_GRAPH=1"; NBODY_NO_GRAPH=1 nbody_b200/bin/nbody-bench --gpu --fresh-seed --n 250,500,800,1000,1200,1600,2000,3000,4000,4736
} > $OUT/r02i_graph_bench.txt 2>&1; cat $OUT/r02i_graph_bench.txt
python tools/launch_bound_probe.py 100 2>&1 | head -4
