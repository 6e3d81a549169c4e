#!/bin/bash
# BASELINE.json config 5 on an 8-GPU box: N=4M strong scaling (1 and 8 GPUs), weak scaling at constant N*M/g
# (1.48M/2.10M/2.97M/4.19M on 1/2/4/8 GPUs), and the CUDA-only 1,000-step drift at N=1M on 8 GPUs.
OUT=gpurun_out; mkdir -p $OUT; PORT=29700
run() { # gpus n tag
  PORT=$((PORT+1))
  if [ $1 -eq 1 ]; then
    timeout 400 python bench.py --gpus 1 --particles $2 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/c5_$3.json 2> $OUT/c5_$3.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port $PORT \
      bench.py --gpus $1 --particles $2 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/c5_$3.json 2> $OUT/c5_$3.err
  fi
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/c5_$3.json").read().strip().splitlines()[-1])
    print("$3 gpus", d["n_gpus"], "N", d["config"]["total_len"], "M", d["config"]["mass_len"], "ms/step %.2f" % d["ms_per_step"], "Gpairs/s %.1f" % d["value"], "frac %.4f" % d["roofline"]["frac"], "e2e ms %.2f" % d["e2e"]["ms_per_step"])
except Exception as e:
    print("$3 failed", e)
PY
}
if [ "${ONLY_MULTI:-0}" = "1" ]; then
  run 8 4194304 strong4m_g8; run 2 2097152 weak_g2; run 4 2965821 weak_g4; exit 0
fi
run 8 4194304 strong4m_g8
run 1 4194304 strong4m_g1
run 1 1482910 weak_g1
run 2 2097152 weak_g2
run 4 2965821 weak_g4
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29799 \
  tools/drift_large.py 1048576 1000 2 > $OUT/c5_drift_1m_g8.jsonl 2> $OUT/c5_drift_1m_g8.err
cat $OUT/c5_drift_1m_g8.jsonl | cut -c1-330
