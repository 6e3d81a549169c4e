#!/usr/bin/env python3
"""Where the clocks of a resident-cluster step go (resident_kernel.cuh built with its TL switch): per CTA, thread 0's
totals over the steps of one call for its pair loop, the __syncthreads, slab sums + integrator + remote stores, and
the cluster barrier.  Usage (GPU box):  python tools/resident_breakdown.py N [steps]"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
os.environ["NBODY_RESIDENT_TL"] = "1"
os.environ.setdefault("NBODY_RESIDENT_MAX_N", "4000")


def main():
    n = int(sys.argv[1])
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
    import nbody_b200 as nb
    from nbody_b200 import api

    with nb.World(nb.make_galaxies(n, 2)) as w:
        m = w.mass_len
        arr = w.particles().copy()
    with nb.SimPipeline(n, m) as sim:
        sim.set_data(arr)
        info = sim.info()
        sim.update(100, 0.01)
        sim.update(steps, 0.01)
        us = sim.last_step_ms() * 1e3 / steps
        buf = np.zeros((160 * 32, 4), dtype=np.uint64)
        api.lib().NbSimDebugTimeline(buf.ctypes.data_as(C.c_void_p), 160 * 32)
    c = info["resident_ctas"]
    t = buf.reshape(-1)[: c * 4].reshape(c, 4).astype(np.float64) / steps
    clk = info["sm_clock_khz"] * 1e-3  # cycles per us
    print(f"N={n} M={m} slabs={info['slabs']} slab_len={info['slab_len']} cluster of {c} CTAs: {us:.2f} us/step = {us * clk:.0f} clk")
    print("  clocks per step, thread 0 of each CTA:   pair loop   __syncthreads   sums+integrate+stores   cluster barrier   total")
    for r in range(c):
        print(f"    CTA {r:2d}                               {t[r,0]:9.0f} {t[r,1]:15.0f} {t[r,2]:23.0f} {t[r,3]:17.0f} {t[r].sum():7.0f}")
    print(f"    mean                                 {t[:,0].mean():9.0f} {t[:,1].mean():15.0f} {t[:,2].mean():23.0f} {t[:,3].mean():17.0f} {t.sum(axis=1).mean():7.0f}")


if __name__ == "__main__":
    main()
