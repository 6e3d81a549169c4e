#!/usr/bin/env python3
"""Where the time of ONE step goes, per warp: the kernel variants built with the timeline switch (b16x2t, w16x4t, ...)
record %globaltimer after the dependency wait, when the sources are in shared memory, after the warp's last unit and
at its end.  Usage (GPU box):  python tools/step_timeline.py N variant [slab_len]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))


def main():
    n = int(sys.argv[1])
    variant = sys.argv[2]
    slab = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    os.environ["NBODY_VARIANT"] = variant
    import ctypes as C

    import nbody_b200 as nb
    from nbody_b200 import api

    ps = nb.make_galaxies(n, 2)
    with nb.World(ps) as w:           # CreateWorld's partition: massive particles first
        m = w.mass_len
        arr = w.particles().copy()
    with nb.SimPipeline(n, m, slab_len=slab) as sim:
        sim.set_data(arr)
        info = sim.info()
        sim.update(20, 0.01)              # warm
        sim.update(50, 0.01)
        chained = sim.last_step_ms() * 1e3 / 50
        sim.update(1, 0.01)
        single = sim.last_step_ms() * 1e3
        buf = np.zeros((160 * 32, 4), dtype=np.uint64)
        api.lib().NbSimDebugTimeline(buf.ctypes.data_as(C.c_void_p), 160 * 32)
    warps = info["threads"] // 32
    ctas_max = 160 * 32 // warps
    t = buf[: ctas_max * warps].reshape(ctas_max, warps, 4).astype(np.int64)
    live = t[:, 0, 0] > 0
    t = t[live]
    t0 = t[:, :, 0].min()
    rel = (t - t0) / 1e3               # us
    start, loaded, pairs, end = rel[:, :, 0], rel[:, :, 1], rel[:, :, 2], rel[:, :, 3]
    busy = pairs > loaded + 1e-9       # warps that had at least one unit
    print(f"N={n} variant={variant} slab_len={info['slab_len']} slabs={info['slabs']} CTAs={t.shape[0]} warps/CTA={warps}"
          f"  chained {chained:.1f} us/step, single step {single:.1f} us (events)")
    def row(name, a, mask=None):
        a = a[mask] if mask is not None else a.ravel()
        print(f"  {name:34s} min {a.min():7.2f}  mean {a.mean():7.2f}  max {a.max():7.2f}")
    row("start after dependency wait", start)
    row("sources in shared memory", loaded)
    row("  load time", loaded - start)
    row("last unit done (busy warps)", pairs, busy)
    row("  pair time (busy warps)", (pairs - loaded), busy)
    cta_pairs = np.where(busy, pairs, 0).max(axis=1)
    row("CTA: last warp out of pair loop", cta_pairs)
    row("CTA: first busy warp out", np.where(busy, pairs, 1e9).min(axis=1))
    row("end of warp", end)
    row("CTA: end", end.max(axis=1))
    print(f"  busy warps per CTA: {busy.sum(axis=1).min()} .. {busy.sum(axis=1).max()};  span {end.max():.2f} us")


if __name__ == "__main__":
    main()
