#!/bin/bash
# tools/gpu_scaling.sh [max_gpus] -- bench.py at 1,2,4,8 GPUs back to back (one rank per GPU), like the driver does.
MAXG=${1:-8}
OUT=gpurun_out
mkdir -p $OUT
PORT=29600
for g in 1 2 4 8; do
  [ $g -gt $MAXG ] && break
  PORT=$((PORT+1))
  if [ $g -eq 1 ]; then
    timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 ${BENCH_ARGS:-} > $OUT/scale_g$g.json 2> $OUT/scale_g$g.err
  else
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $g --master-addr 127.0.0.1 --master-port $PORT \
        bench.py --gpus $g --steps 10 --warmup 3 ${BENCH_ARGS:-} > $OUT/scale_g$g.json 2> $OUT/scale_g$g.err
  fi
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/scale_g$g.json").read().strip().splitlines()[-1])
    print("gpus", d["n_gpus"], "ms/step %.3f" % d["ms_per_step"], "Gpairs/s %.1f" % d["value"], "frac %.4f" % d["roofline"]["frac"], "e2e %.1f" % d["e2e"]["value"])
except Exception as e:
    print("gpus $g failed", e)
PY
done
