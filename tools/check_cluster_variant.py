#!/usr/bin/env python
"""Cluster/DSMEM slab reduction (NBODY_VARIANT=c128x2 / c128x4) against the ticket variants with the same
tiles and slabs: results must be bit-identical (same summation order), and the chained step time is printed
for both.  Each case runs in its own process (the variant is read from the environment at pipeline creation,
and a CUDA error aborts the process).

    python tools/check_cluster_variant.py [N ...]
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = r"""
import sys, json, time, hashlib
sys.path.insert(0, %r)
import nbody_b200 as nb
n, steps = int(sys.argv[1]), int(sys.argv[2])
ps = nb.make_galaxies(n, 2)
with nb.World(ps) as w:
    w.update_gpu(0.01, 3)
    digest = hashlib.sha1(w.particles().tobytes()).hexdigest()
    info = w.pipeline_info()
    w.update_gpu(0.01, 20)
    t0 = time.perf_counter(); w.update_gpu(0.01, steps); dt = time.perf_counter() - t0
print(json.dumps(dict(digest=digest, us_per_step=dt / steps * 1e6, slabs=info["slabs"], slab_len=info["slab_len"],
                      tile=info["tile_targets"])))
""" % ROOT


def run(n, variant, slab_len, steps=200):
    env = dict(os.environ, NBODY_VARIANT=variant, NBODY_SLAB=str(slab_len))
    out = subprocess.run([sys.executable, "-c", CHILD, str(n), str(steps)], env=env, capture_output=True, text=True,
                         timeout=120)
    if out.returncode != 0:
        return dict(error=(out.stderr or out.stdout)[-400:])
    return json.loads(out.stdout.strip().splitlines()[-1])


def main():
    sizes = [int(a) for a in sys.argv[1:]] or [4000, 10000, 16384]
    ok = True
    for n in sizes:
        m = n // 2
        for pair, slabs in ((("t128x2", "c128x2"), (8, 16)), (("t128x4", "c128x4"), (16,))):
            for ns in slabs:
                slab_len = -(-(-(-m // ns)) // 16) * 16
                a, b = run(n, pair[0], slab_len), run(n, pair[1], slab_len)
                same = "error" not in a and "error" not in b and a["digest"] == b["digest"]
                ok &= same
                print(json.dumps(dict(n=n, slabs=a.get("slabs"), slab_len=slab_len, ticket=pair[0], cluster=pair[1],
                                      bit_identical=same, ticket_us=a.get("us_per_step"), cluster_us=b.get("us_per_step"),
                                      error=b.get("error") or a.get("error"))), flush=True)
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
