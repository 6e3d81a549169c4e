#!/usr/bin/env python
"""CUDA-only 1,000-step drift at sizes the CPU reference cannot finish (BASELINE.json config 5), one rank per GPU.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 \
      tools/drift_large.py [N] [steps] [checkpoints]

Diagnostics (P = sum m v, K, U over massive pairs with target softening, E = K + U) are computed in FP64 ON THE
DEVICE (NbSimDiagnostics: the same definitions as oracle_diagnostics_f64, cross-checked in
tests/test_gpu_drift.py::test_device_diagnostics_match_oracle).  Prints one JSON line per checkpoint on rank 0.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import nbody_b200 as nb

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    checks = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    dt = 0.01  # ref src/main.c:14
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
    torch.cuda.set_device(local)
    uid = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        box = [nb.get_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        uid = box[0]
    ps = nb.make_galaxies(n, 2)
    with nb.World(ps) as w0:
        arr, m = w0.particles(), w0.mass_len
    with nb.SimPipeline(len(arr), m, device=local, local_gpus=1, rank=rank, world=world, unique_id=uid) as sim:
        sim.set_data(arr)
        d0 = sim.diagnostics()
        rows = [(0, d0, 0.0)]
        per = steps // checks
        for c in range(checks):
            t0 = time.perf_counter()
            sim.update(per, dt)
            rows.append(((c + 1) * per, sim.diagnostics(), time.perf_counter() - t0))
        if rank == 0:
            e0, p0 = abs(d0["E"]), float(np.hypot(*d0["P"]))
            for step, d, secs in rows:
                print(json.dumps({
                    "n": n, "mass_len": m, "gpus": world, "step": step, "dt": dt,
                    "E": d["E"], "K": d["K"], "U": d["U"], "Px": d["P"][0], "Py": d["P"][1],
                    "dE_over_E0": (d["E"] - d0["E"]) / e0,
                    "dP_over_P0": float(np.hypot(d["P"][0] - d0["P"][0], d["P"][1] - d0["P"][1])) / p0,
                    "ms_per_step": secs / per * 1e3 if step else None}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
