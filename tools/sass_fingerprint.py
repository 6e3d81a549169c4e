#!/usr/bin/env python3
"""Fingerprint of the machine code of the step kernel's inner loop.

ptxas chooses the order of the ~200 instructions of the pair loop, and that order is worth a few per
cent: an extra 4-byte kernel parameter used only in the epilogue changed the register allocation
(122 -> 121) and the schedule, and N = 1M went from 157.6 to 161.4 ms per step with the same instruction
counts.  This tool reduces the loop to its opcode sequence (register names and
addresses dropped) and hashes it, so that a change of schedule is noticed on the build machine, before
anything is measured.  `tests/test_host.py` compares the hash of the built library with
`profiles/kernel_fingerprint.json`; after an intended kernel change, re-measure on a B200 and run

    python tools/sass_fingerprint.py --update

Usage: sass_fingerprint.py [--lib PATH] [--update] [--dump]
"""
import argparse
import hashlib
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DEFAULT_LIB = os.path.join(ROOT, "nbody_b200", "lib", "libnbody.so")
FINGERPRINT = os.path.join(ROOT, "profiles", "kernel_fingerprint.json")

# kernels of the plan's candidate table (sim_cuda.cu NB_VARIANT): THREADS, T, TJP, NS, MINB, UNR, FORM
# ...Lb0: one CTA per slab; ...Lb1 ("g" names): the GROUPED build, a CTA walks a group of slabs (large Worlds)
KERNELS = {
    "t512x4": "_ZN2nb11step_kernelILi512ELi4ELi128ELi4ELi1ELi2ELi0ELb0EEEvNS_10StepParamsE",
    "t256x4": "_ZN2nb11step_kernelILi256ELi4ELi128ELi4ELi2ELi2ELi0ELb0EEEvNS_10StepParamsE",
    "t128x4": "_ZN2nb11step_kernelILi128ELi4ELi128ELi4ELi4ELi2ELi0ELb0EEEvNS_10StepParamsE",
    "t128x2": "_ZN2nb11step_kernelILi128ELi2ELi64ELi4ELi6ELi2ELi0ELb0EEEvNS_10StepParamsE",
    "g512x4": "_ZN2nb11step_kernelILi512ELi4ELi128ELi4ELi1ELi2ELi0ELb1EEEvNS_10StepParamsE",
}


def cuobjdump():
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    return exe if os.path.exists(exe) else None


def opcodes(lib, symbol):
    """Opcode sequence of one kernel (predicates and operands dropped)."""
    out = subprocess.run([cuobjdump(), "-sass", "-fun", symbol, lib], capture_output=True, text=True, check=True).stdout
    ops = []
    for line in out.splitlines():
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(.*?);", line)
        if not m:
            continue
        tok = m.group(1).split()
        if tok and tok[0].startswith("@"):
            tok = tok[1:]
        if tok:
            ops.append(tok[0])
    return ops


def pair_loop(ops):
    """The pair loop: from the wait on a stage's 'full' barrier to the arrive on its 'empty' barrier."""
    waits = [i for i, o in enumerate(ops) if o.startswith("SYNCS.PHASECHK")]
    arrives = [i for i, o in enumerate(ops) if o.startswith("SYNCS.ARRIVE.TRANS64.A1T0")]
    if not arrives:
        return []
    end = arrives[0]
    start = max((w for w in waits if w < end), default=0)
    return ops[start:end + 1]


def toolkit_version():
    """`release 12.9, V12.9.86` of the nvcc that would build the library (ptxas ships with it)."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    try:
        out = subprocess.run([nvcc, "--version"], capture_output=True, text=True).stdout
    except OSError:
        return "unknown"
    m = re.search(r"release [0-9.]+, V[0-9.]+", out)
    return m.group(0) if m else "unknown"


def fingerprint(lib):
    fp = {}
    for name, sym in KERNELS.items():
        ops = opcodes(lib, sym)
        loop = pair_loop(ops)
        fp[name] = {
            "loop_sha1": hashlib.sha1("\n".join(loop).encode()).hexdigest(),
            "loop_instructions": len(loop),
            "loop_f32x2": sum(o in ("FFMA2", "FMUL2", "FADD2") for o in loop),
            "loop_mufu": sum(o.startswith("MUFU") for o in loop),
            "loop_lds": sum(o.startswith("LDS") for o in loop),
            "kernel_instructions": len(ops),
        }
    return fp


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default=DEFAULT_LIB)
    ap.add_argument("--update", action="store_true", help="write profiles/kernel_fingerprint.json")
    ap.add_argument("--dump", action="store_true", help="print the opcode sequence of the t512x4 pair loop")
    args = ap.parse_args()
    if cuobjdump() is None:
        sys.exit("cuobjdump not found")
    if args.dump:
        print("\n".join(pair_loop(opcodes(args.lib, KERNELS["t512x4"]))))
        return
    fp = fingerprint(args.lib)
    print(json.dumps(fp, indent=1))
    if args.update:
        fp["_meta"] = {"toolkit": toolkit_version(),
                       "procedure": "after an intended kernel or toolkit change: make tune && gpurun "
                                    "'TUNE_SET=ship tools/bin/tune sweep 1048576 523884 37' (and python bench.py), "
                                    "compare with profiles/*_tune_sweep*.txt / DESIGN.md section 7, then "
                                    "python tools/sass_fingerprint.py --update and commit the new numbers"}
        with open(FINGERPRINT, "w") as f:
            json.dump(fp, f, indent=1)
            f.write("\n")
    elif os.path.exists(FINGERPRINT):
        want = json.load(open(FINGERPRINT))
        bad = [k for k in fp if not k.startswith("_") and want.get(k, {}).get("loop_sha1") != fp[k]["loop_sha1"]]
        if bad:
            sys.exit("schedule of the pair loop changed for: " + ", ".join(bad))


if __name__ == "__main__":
    main()
