#!/usr/bin/env python
"""One rank per GPU (torchrun): the sharded pipeline must reproduce the single-GPU result bit for bit.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tools/check_multirank.py [N] [steps]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    import nbody_b200 as nb

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 50000
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ps = nb.make_galaxies(n, 2)
    with nb.World(ps) as w0:
        arr, m = w0.particles(), w0.mass_len
    box = [nb.get_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    with nb.SimPipeline(len(arr), m, device=local, local_gpus=1, rank=rank, world=world, unique_id=box[0]) as sim:
        # sharded upload: a rank reads only the particles it owns out of the array it is given -- poison the rest
        mine = sim.local_ranges()
        poisoned = arr.copy()
        own = np.zeros(len(arr), dtype=bool)
        for lo, hi in mine:
            own[lo:hi] = True
        for k in ("px", "py", "vx", "vy", "mass", "radius"):
            poisoned[k][~own] = np.nan
        sim.set_data(poisoned)
        sim.update(steps, 0.01)
        sharded = sim.get_data()                      # gathers the other ranks' particles
        mine_only = np.zeros(len(arr), dtype=arr.dtype)
        sim.get_data_local(mine_only)                     # sharded read-back: own ranges only, rest untouched
        ms = sim.last_step_ms()
    local_ok = all(mine_only[lo:hi].tobytes() == sharded[lo:hi].tobytes() for lo, hi in mine) and \
        not mine_only[~own].view(np.uint8).any() and sum(hi - lo for lo, hi in mine) > 0
    flags = torch.tensor([int(local_ok)], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    local_ok = bool(flags.item())
    ok = True
    if rank == 0:
        with nb.SimPipeline(len(arr), m, device=local, local_gpus=1) as one:
            one.set_data(arr)
            one.update(steps, 0.01)
            ref = one.get_data()
        ok = sharded.tobytes() == ref.tobytes() and local_ok
        print(f"check_multirank: N={n} M={m} world={world} steps={steps}: "
              f"{'bit-identical to 1 GPU, sharded upload/read-back ok' if ok else 'MISMATCH'}; last call {ms:.3f} ms", flush=True)
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
