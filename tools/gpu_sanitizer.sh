#!/bin/bash
# compute-sanitizer over the small-N GPU tests (memcheck, then racecheck and synccheck on a subset).
set -u
OUT=gpurun_out; mkdir -p $OUT
CS=/usr/local/cuda/bin/compute-sanitizer
SEL="ragged or two_body or massless or no_massive or chained or slab_groups or persistent or zero_steps or get_between"
timeout 900 $CS --tool memcheck --error-exitcode 9 --leak-check no python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$SEL" > $OUT/r02_sanitizer_memcheck.log 2>&1
echo "memcheck rc=$?" | tee -a $OUT/r02_sanitizer_memcheck.log
SEL2="two_body or chained or (slab_groups and 4003) or (persistent and 1000)"
timeout 900 $CS --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$SEL2" > $OUT/r02_sanitizer_racecheck.log 2>&1
echo "racecheck rc=$?" | tee -a $OUT/r02_sanitizer_racecheck.log
timeout 900 $CS --tool synccheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$SEL2" > $OUT/r02_sanitizer_synccheck.log 2>&1
echo "synccheck rc=$?" | tee -a $OUT/r02_sanitizer_synccheck.log
timeout 600 $CS --tool initcheck --error-exitcode 9 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "$SEL2" > $OUT/r02_sanitizer_initcheck.log 2>&1
echo "initcheck rc=$?" | tee -a $OUT/r02_sanitizer_initcheck.log
tail -5 $OUT/r02_sanitizer_*.log
