#!/usr/bin/env python
"""Generate tests/golden/*.npz by RUNNING THE REFERENCE (oracle/_ref, compiled from /root/reference).

The reference ships no golden vectors for the force kernel or the generator, so these fixtures
are outputs of the reference's own code executed in the build container:

  galaxy_<n>_<g>.npz : srand(11037); MakeGalaxies(n, g)                     (ref src/lib/galaxy.c)
                       + CreateWorld order (ref src/lib/world.c:32-46)
                       + UpdateWorld_CPU(w, dt, 1) and (w, dt, 10), AVX build (ref world.c:99-110)
                       + the same with the SSE and scalar builds (1 step)

They travel with the repository so the GPU box (which has no /root/reference) can check against
them.  Re-run:  python tools/make_golden.py   (needs oracle/_ref; `make oracle` builds it).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _oracle import RefLib  # noqa: E402

CASES = [(1000, 2, 0.01), (4003, 2, 0.01), (6000, 3, 0.01)]


def main():
    out_dir = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out_dir, exist_ok=True)
    avx, sse, scalar = RefLib("avx"), RefLib("sse"), RefLib("scalar")
    for n, g, dt in CASES:
        ps = avx.make_galaxies(n, g, seed=11037)
        world0 = avx.world_state(ps)
        step1 = avx.update_cpu(ps, dt, 1)
        step10 = avx.update_cpu(ps, dt, 10)
        step1_sse = sse.update_cpu(ps, dt, 1)
        step1_scalar = scalar.update_cpu(ps, dt, 1)
        path = os.path.join(out_dir, f"galaxy_{n}_{g}.npz")
        np.savez_compressed(
            path, input=ps, world0=world0, step1=step1, step10=step10, step1_sse=step1_sse,
            step1_scalar=step1_scalar, dt=np.float32(dt), seed=np.int64(11037),
        )
        print(path, os.path.getsize(path), "bytes; mass_len =", int(np.count_nonzero(ps["mass"] > 0)))


if __name__ == "__main__":
    main()
