#!/bin/bash
# Round-2 run N: per-warp timelines of one step, small-World kernel against the balanced kernel.
set -u
OUT=gpurun_out; mkdir -p $OUT
{
for n in 16384 10000; do
  for v in w16x4t b16x2t b16x2m7t b16x2m2t; do timeout 120 python tools/step_timeline.py $n $v; done
done
timeout 120 python tools/step_timeline.py 14208 w16x3t
timeout 120 python tools/step_timeline.py 4736 w32x1t
timeout 120 python tools/step_timeline.py 4736 b16x2t
} > $OUT/r02n_timelines.txt 2>&1; cat $OUT/r02n_timelines.txt
