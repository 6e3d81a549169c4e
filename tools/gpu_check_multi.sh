#!/bin/bash
# Round-2 multi-GPU check (run with gpurun --gpus 2|4|8): GPU suite incl. the multi-GPU tests, then bench.py at 1..G GPUs.
set -u
OUT=gpurun_out; mkdir -p $OUT
G=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests -m gpu -q -x > $OUT/r02e_gpu_tests_g$G.log 2>&1; echo "pytest rc=$?" >> $OUT/r02e_gpu_tests_g$G.log
tail -4 $OUT/r02e_gpu_tests_g$G.log
bash tools/gpu_scaling.sh $G
for g in 1 2 4 8; do [ ! -f $OUT/scale_g$g.json ] || cp $OUT/scale_g$g.json $OUT/r02e_scale_box${G}_g$g.json; done
