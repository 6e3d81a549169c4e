#!/usr/bin/env python
"""BASELINE.json configs 2 and 3 (plus a size ladder) on one B200, through the C ABI.

  config 2: N=16384 galaxy init, single-step latency: median of 1000 synchronous UpdateWorld_GPU(w, dt, 1) calls
  config 3: N=262144 galaxy init, chained multi-step: UpdateWorld_GPU(w, dt, 100) per call, 5 calls
  ladder  : the reference bench's own sizes (ref src/bench.c:38) + 16K/64K/256K, 10 warm-up + 100 steps in one call

Prints one JSON line per row; device time from CUDA events (NbSimLastStepMs), wall time around the call.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_b200 as nb  # noqa: E402
import ctypes as C  # noqa: E402


def last_ms(world):
    sim = nb.lib().NbWorldPipeline(world._w)
    return float(nb.lib().NbSimLastStepMs(sim))


def row(tag, n, steps_per_call, calls, warm_calls, dt=1.0, galaxies=2):
    ps = nb.make_galaxies(n, galaxies)
    with nb.World(ps) as w:
        m = w.mass_len
        for _ in range(warm_calls):
            w.update_gpu(dt, steps_per_call)
        wall, dev = [], []
        for _ in range(calls):
            t0 = time.perf_counter()
            w.update_gpu(dt, steps_per_call)          # synchronous, like the reference
            wall.append(time.perf_counter() - t0)
            dev.append(last_ms(w))
        info = w.pipeline_info()
    pairs = float(n) * m * steps_per_call
    wall_med, dev_med = float(np.median(wall)), float(np.median(dev)) * 1e-3
    peak = info["sm_count"] * 128 * 2 * info["sm_clock_khz"] * 1e3
    out = {
        "tag": tag, "n": n, "mass_len": m, "steps_per_call": steps_per_call, "calls": calls,
        "us_per_step_wall": wall_med / steps_per_call * 1e6, "us_per_step_device": dev_med / steps_per_call * 1e6,
        "gpairs_per_s_wall": pairs / wall_med * 1e-9, "gpairs_per_s_device": pairs / dev_med * 1e-9,
        "frac_fp32_peak_device": pairs / dev_med * 14 / peak,
        "variant": info["variant"], "slabs": info["slabs"],
        "slab_len": info["slab_len"],
    }
    print(json.dumps(out), flush=True)


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "all"
    if which in ("all", "c2"):
        row("config2 N=16K single-step latency", 16384, 1, 1000, 10)
    if which in ("all", "c3"):
        row("config3 N=256K 100 steps per call", 262144, 100, 5, 1)
    if which in ("all", "ladder"):
        for n in (250, 500, 800, 1200, 2000, 4000, 10000, 16384, 20000, 50000, 65536, 100000, 262144):
            row("ladder 10+100 steps (ref bench.c protocol)", n, 100, 1, 1 if n > 100000 else 1)


if __name__ == "__main__":
    main()
