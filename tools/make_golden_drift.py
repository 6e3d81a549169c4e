#!/usr/bin/env python
"""Generate tests/golden/drift_*.npz by RUNNING THE REFERENCE (oracle/_ref): 1,000 steps of UpdateWorld_CPU
(AVX build) at dt = 0.01 (ref src/main.c:14), FP64 diagnostics (oracle_diagnostics_f64) every 100 steps.

north_star: "total-energy and momentum drift matching the reference within 1% over 1,000 steps".
The fixture holds the reference's drift so the GPU box (no /root/reference, few CPU minutes) can compare.
Re-run:  python tools/make_golden_drift.py   (needs oracle/_ref; `make oracle` builds it)
"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _oracle import Oracle, RefLib, _ptr, mass_len_of  # noqa: E402

CASES = [(6000, 3, 0.01, 1000, 100), (16384, 2, 0.01, 1000, 100)]


def run_flavour(ref, oracle, ps, dt, steps, every):
    w = ref.lib.CreateWorld(_ptr(ps), len(ps))
    m = mass_len_of(ref._get(w))
    rows = []

    def snap(step):
        d = oracle.diagnostics(ref._get(w), m)
        rows.append([step, d["P"][0], d["P"][1], d["K"], d["U"], d["E"], d["M"], d["C"][0], d["C"][1]])

    snap(0)
    for s in range(every, steps + 1, every):
        ref.lib.UpdateWorld_CPU(w, dt, every)
        snap(s)
    final = ref._get(w)
    ref.lib.DestroyWorld(w)
    return rows, final, m


def main():
    ref, oracle = RefLib("avx"), Oracle()
    for n, g, dt, steps, every in CASES:
        ps = ref.make_galaxies(n, g, seed=11037)
        rows, final, m = run_flavour(ref, oracle, ps, dt, steps, every)
        # the reference's OWN sensitivity to summation order: its SSE and scalar builds on the same input
        rows_sse, final_sse, _ = run_flavour(RefLib("sse"), oracle, ps, dt, steps, every)
        rows_scalar, _, _ = run_flavour(RefLib("scalar"), oracle, ps, dt, steps, every)
        path = os.path.join(ROOT, "tests", "golden", f"drift_{n}_{g}.npz")
        np.savez_compressed(path, diag=np.array(rows, dtype=np.float64),
                            diag_sse=np.array(rows_sse, dtype=np.float64),
                            diag_scalar=np.array(rows_scalar, dtype=np.float64),
                            final_pos_sse=np.stack([final_sse["px"], final_sse["py"]], axis=1), dt=np.float32(dt), seed=np.int64(11037),
                            n=np.int64(n), galaxies=np.int64(g), mass_len=np.int64(m),
                            columns=np.array(["step", "Px", "Py", "K", "U", "E", "M", "Cx", "Cy"]),
                            final_pos=np.stack([final["px"], final["py"]], axis=1))
        r = np.array(rows)
        for tag, rr in (("sse", np.array(rows_sse)), ("scalar", np.array(rows_scalar))):
            print("   %s build: dE %.6e (avx %.6e, ratio %.4f)" % (tag, rr[-1, 5] - rr[0, 5], r[-1, 5] - r[0, 5],
                                                               (rr[-1, 5] - rr[0, 5]) / (r[-1, 5] - r[0, 5])))
        print(path, "E0 %.6e E1000 %.6e dE/E %.3e |P0| %.3e |P1000| %.3e" % (
            r[0, 5], r[-1, 5], (r[-1, 5] - r[0, 5]) / abs(r[0, 5]), np.hypot(r[0, 1], r[0, 2]), np.hypot(r[-1, 1], r[-1, 2])))


if __name__ == "__main__":
    main()
