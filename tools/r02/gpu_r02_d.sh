#!/bin/bash
# Does a single wave of CTAs that each walk 2-3 slabs beat two waves of one-slab CTAs at mid N?  (forced with NBODY_PART_BUDGET)
set -u
OUT=gpurun_out
mkdir -p $OUT
{
for n in 10000 16384 20000 30000; do
  echo "== N=$n ungrouped"
  nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $n --json
  for g in 4 5 6 7 8 9 10 11 12 13 14; do
    tiles=$(( (n + 255) / 256 ))
    b=$(( g * tiles * 256 * 8 + 8 ))
    echo "== N=$n max groups $g (budget $b)"
    NBODY_PART_BUDGET=$b nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $n --json
  done
done
} > $OUT/r02d_midN_groups.txt 2>&1
grep -v "^==" $OUT/r02d_midN_groups.txt | python3 -c "
import sys,json
for l in sys.stdin:
    try: d=json.loads(l); print(d['n'], d['gpu_us_per_step'])
    except Exception as e: print(l.strip())
"
