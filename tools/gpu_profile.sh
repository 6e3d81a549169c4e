#!/bin/bash
# tools/gpu_profile.sh -- the profiling recipe of /opt/skills/guides/B200_PROFILING.md for this repo.
# Run on the GPU box:  gpurun --timeout 1500 -- 'bash tools/gpu_profile.sh <tag>'
# Writes gpurun_out/<tag>_*.  Plain run first; ncu only if it exited 0.
set -u
TAG=${1:-prof}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
    --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
$CMD > $OUT/${TAG}_plain2.json 2> $OUT/${TAG}_plain2.err &&
ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 3 -c 1 \
    -f -o $OUT/${TAG}_step $CMD > $OUT/${TAG}_ncu_full.log 2>&1
ls -la $OUT
