#!/bin/bash
# Round-2 run N: per-warp timelines of one step, small-World kernel against the balanced kernel.
set -u
OUT=gpurun_out; mkdir -p $OUT
{
for n in 16384 10000 20000; do
  for v in ${TLV:-w16x4t b16x2t b16x4t}; do timeout 120 python tools/step_timeline.py $n $v; done
done
} > $OUT/r02n_timelines.txt 2>&1; cat $OUT/r02n_timelines.txt
