#!/usr/bin/env python
"""Is a chain of n small steps bound by the host's launch rate or by the GPU?  Times NbSimStepAsync's return (host
enqueue of n launches) against the device time of the same call (CUDA events) at small N.
Usage: python tools/launch_bound_probe.py [n_steps]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbody_b200 as nb  # noqa: E402
from _oracle import Oracle  # noqa: E402  (partition only: test infrastructure used by a tool, not by the product)

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
oracle = Oracle()
for n in (250, 1000, 2000, 4000, 10000, 16384):
    arr, m = oracle.partition(nb.make_galaxies(n, 2))
    with nb.SimPipeline(n, m) as sim:
        sim.set_data(arr)
        sim.update(10, 1.0)
        best = None
        for _ in range(5):
            t0 = time.perf_counter()
            sim.step_async(steps, 1.0)
            t1 = time.perf_counter()
            sim.synchronize()
            t2 = time.perf_counter()
            row = dict(n=n, steps=steps, host_enqueue_us_per_step=(t1 - t0) / steps * 1e6,
                       wall_us_per_step=(t2 - t0) / steps * 1e6, device_us_per_step=sim.last_step_ms() / steps * 1e3)
            if best is None or row["wall_us_per_step"] < best["wall_us_per_step"]:
                best = row
        print(json.dumps(best), flush=True)
