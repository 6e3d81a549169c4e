#!/usr/bin/env python3
"""Build the large-N step kernel under every setting of the source-order knobs (force_kernel.cuh NB_ORD_*),
group the builds by the opcode sequence ptxas gave the pair loop, and write one benchmark binary per distinct
schedule plus the script that times them on a B200.

    python tools/schedule_lottery.py            # here (no GPU): build into tools/lottery/, print the table
    gpurun -- 'bash tools/lottery/run.sh'       # on the box: gpurun_out/lottery.txt, one "name ms" line each
    rm -rf tools/lottery                        # the binaries are scratch (git-ignored)

Why: the loop's instruction ORDER, which ptxas derives from the order of the source, moves N = 1M between 156.3
and 160.7 ms with identical instruction counts (profiles/r01_schedule_lottery.txt).  All settings compute
bit-identical results except NB_ORD_R2=1 (the two squares are added to the softening in the other order), which
is therefore left out unless --with-r2 is given.
"""
import argparse
import concurrent.futures as cf
import hashlib
import itertools
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sass_fingerprint as fp  # noqa: E402

AXES = ("ACC", "R2", "GI", "LOOP", "F", "SUB")
RANGES = {"ACC": (0, 1, 2), "R2": (0, 1), "GI": (0, 1), "LOOP": (0, 1), "F": (0, 1), "SUB": (0, 1, 2)}
OUT = os.path.join(ROOT, "tools", "lottery")


GROUPED = False  # --grouped: the knobs of the GROUPED build (NB_ORDG_*), judged on its own loop


def build(setting):
    name = ("g_" if GROUPED else "s_") + "".join(str(v) for v in setting)
    exe = os.path.join(OUT, name)
    cmd = ["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-DTUNE_MINIMAL",
           "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "nbody_b200", "csrc")]
    prefix = "NB_ORDG_" if GROUPED else "NB_ORD_"
    cmd += [f"-D{prefix}{a}={v}" for a, v in zip(AXES, setting) if not (GROUPED and a == "R2")]
    cmd += [os.path.join(ROOT, "tools", "tune.cu"), "-o", exe]
    subprocess.run(cmd, capture_output=True, check=True)
    loop = fp.pair_loop(fp.opcodes(exe, fp.KERNELS["g512x4" if GROUPED else "t512x4"]))
    return name, hashlib.sha1("\n".join(loop).encode()).hexdigest()[:10]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--with-r2", action="store_true", help="include NB_ORD_R2=1 (changes rounding)")
    ap.add_argument("--jobs", type=int, default=8)
    ap.add_argument("--grouped", action="store_true", help="the GROUPED build's knobs (NB_ORDG_*) and loop")
    args = ap.parse_args()
    global GROUPED
    GROUPED = args.grouped
    os.makedirs(OUT, exist_ok=True)
    ranges = [RANGES[a] if (a != "R2" or (args.with_r2 and not GROUPED)) else (0,) for a in AXES]
    settings = list(itertools.product(*ranges))
    with cf.ThreadPoolExecutor(args.jobs) as ex:
        built = list(ex.map(build, settings))
    groups = {}
    for name, h in built:
        groups.setdefault(h, []).append(name)
    keep = []
    print(f"{len(built)} builds, {len(groups)} distinct schedules of the pair loop   (digits: {' '.join(AXES)})")
    for h, names in groups.items():
        keep.append(names[0])
        print(f"  {h}  {names[0]}  (+{len(names) - 1} identical)")
        for n in names[1:]:
            os.remove(os.path.join(OUT, n))
    tune_set, pat, out = ("ship-grouped", "grouped", "lottery_grouped") if GROUPED else ("ship", "packed", "lottery")
    with open(os.path.join(OUT, "run_grouped.sh" if GROUPED else "run.sh"), "w") as f:
        f.write("#!/bin/bash\n# times every distinct schedule at N = 1M (37 slabs); one 'name ms' line each\n"
                "mkdir -p gpurun_out\nfor b in " + " ".join(keep) + "; do\n"
                f"  echo \"$b $(TUNE_SET={tune_set} timeout 60 tools/lottery/$b sweep 1048576 523884 37 | grep {pat} | "
                "sed -E 's/.*best +([0-9.]+) ms.*/\\1/')\"\ndone | tee gpurun_out/" + out + ".txt\n"
                "sort -k2 -n gpurun_out/" + out + ".txt | head -5\n")


if __name__ == "__main__":
    main()
