"""INTEGRATION.md level B, executed: the reference's OWN, unmodified world.c (compiled from /root/reference into
oracle/_ref/libnbody_ref_cuda.so by oracle/Makefile) drives the product's CUDA pipeline through the five
functions of its GPU seam (ref src/lib/sim_gpu.h:21-36).  If the seam is a faithful drop-in, the reference's
World and the product's World give bit-identical particles."""
import ctypes as C
import os

import numpy as np
import pytest

from _oracle import ORACLE_DIR, PARTICLE, _ptr

pytestmark = pytest.mark.gpu
LIB = os.path.join(ORACLE_DIR, "_ref", "libnbody_ref_cuda.so")


class RefWorldOnCuda:
    def __init__(self):
        self.lib = L = C.CDLL(LIB, mode=C.RTLD_LOCAL)
        L.CreateWorld.restype = C.c_void_p
        L.CreateWorld.argtypes = [C.c_void_p, C.c_uint32]
        L.DestroyWorld.argtypes = [C.c_void_p]
        L.GetWorldParticles.restype = C.c_void_p
        L.GetWorldParticles.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
        L.UpdateWorld_CPU.argtypes = [C.c_void_p, C.c_float, C.c_uint32]
        L.UpdateWorld_GPU.argtypes = [C.c_void_p, C.c_float, C.c_uint32]

    def get(self, w):
        n = C.c_uint32(0)
        p = self.lib.GetWorldParticles(w, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value * 32,)).view(PARTICLE).copy()


@pytest.fixture(scope="module")
def refcuda():
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libnbody_ref_cuda.so not built (needs /root/reference at build time)")
    return RefWorldOnCuda()


@pytest.mark.parametrize("n,gal", [(1000, 2), (6000, 3), (20000, 2)])
def test_reference_world_c_on_cuda_seam_matches_product_world(nb, refcuda, n, gal):
    ps = nb.make_galaxies(n, gal)
    w = refcuda.lib.CreateWorld(_ptr(ps), len(ps))              # reference world.c: partition + CreateSimPipeline
    assert refcuda.get(w).tobytes() == nb.World(ps).particles().tobytes()
    refcuda.lib.UpdateWorld_GPU(w, 0.01, 1)                      # SetSimulationData + PerformSimUpdate
    refcuda.lib.UpdateWorld_GPU(w, 0.01, 4)
    ref_gpu = refcuda.get(w)                                     # GetSimulationData
    refcuda.lib.UpdateWorld_CPU(w, 0.01, 2)                      # reference AVX path on the downloaded state
    refcuda.lib.UpdateWorld_GPU(w, 0.02, 3)                      # re-upload, new dt
    ref_mixed = refcuda.get(w)
    refcuda.lib.DestroyWorld(w)                                  # DestroySimPipeline

    with nb.World(ps) as mine:
        mine.update_gpu(0.01, 1)
        mine.update_gpu(0.01, 4)
        assert mine.particles().tobytes() == ref_gpu.tobytes()
        mine.update_cpu(0.01, 2)                                 # product UpdateWorld_CPU == reference AVX, bit for bit
        mine.update_gpu(0.02, 3)
        assert mine.particles().tobytes() == ref_mixed.tobytes()


def test_reference_bench_binary_runs_on_cuda_seam():
    import subprocess
    exe = os.path.join(ORACLE_DIR, "_ref", "nbody-bench-ref-cuda")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/nbody-bench-ref-cuda not built")
    out = subprocess.run([exe, "--gpu"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-500:]
    rows = [l.split() for l in out.stdout.strip().splitlines()[1:]]
    assert [int(r[0]) for r in rows] == [250, 500, 800, 1200, 2000, 4000, 10000, 20000, 50000, 100000]
    print("\n[level B] unmodified ref src/bench.c --gpu on the CUDA seam (us/step):", {int(r[0]): int(r[1]) for r in rows})
