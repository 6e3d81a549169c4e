#!/bin/bash
# tools/sweep_plan.sh "N N ..." -- chained-step time (nbody-bench --gpu, 200 steps) for forced (variant, slab length)
# against the planner's own choice; checks the planner's model on the box.  Output: one line per run.
for n in ${1:-4000 10000 16384 20000 50000}; do
  m=$((n / 2))
  echo "N=$n planner: $(nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $n --steps 200 --warmup 20 --json)"
  for v in ${VARIANTS:-t128x2 t128x4 t64x4 t256x4}; do
    for ns in ${SLABS:-6 8 10 13 16 20 23 27 32 41}; do
      len=$(( (m + ns - 1) / ns )); len=$(( (len + 15) / 16 * 16 ))
      r=$(NBODY_VARIANT=$v NBODY_SLAB=$len nbody_b200/bin/nbody-bench --gpu --fresh-seed --n $n --steps 200 --warmup 20 --json 2>&1 | tail -1)
      us=$(echo "$r" | sed -n 's/.*"gpu_us_per_step": \([0-9.]*\).*/\1/p')
      echo "N=$n $v nslab~$ns slab_len=$len us/step=$us"
    done
  done
done
