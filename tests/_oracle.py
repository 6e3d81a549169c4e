"""ctypes loaders for the CHECKERS under oracle/ (test infrastructure only).

* ``Oracle``  -> oracle/_build/liboracle.so, our C restatement (oracle/nbody_oracle.c)
* ``RefLib``  -> oracle/_ref/libnbody_ref*.so, the reference itself compiled from
                /root/reference by oracle/Makefile (prebuilt files travel to the GPU box)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")

# include/nbody.h Particle: pos, vel, acc (V2 each), mass, radius -- 32 bytes
PARTICLE = np.dtype(
    [("px", "<f4"), ("py", "<f4"), ("vx", "<f4"), ("vy", "<f4"),
     ("ax", "<f4"), ("ay", "<f4"), ("mass", "<f4"), ("radius", "<f4")]
)
assert PARTICLE.itemsize == 32


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def build_oracle():
    """(Re)build oracle/_build/liboracle.so; also oracle/_ref when the reference is present."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True)
    if os.path.isdir("/root/reference/src/lib"):
        subprocess.run(["make", "-s", "-C", ORACLE_DIR, "ref"], check=True)


class Oracle:
    def __init__(self):
        path = os.path.join(ORACLE_DIR, "_build", "liboracle.so")
        if not os.path.exists(path):
            build_oracle()
        self.lib = L = C.CDLL(path)
        L.oracle_partition.restype = C.c_uint32
        L.oracle_partition.argtypes = [C.c_void_p, C.c_uint32]
        L.oracle_partition_int.restype = C.c_uint32
        L.oracle_partition_int.argtypes = [C.c_void_p, C.c_uint32]
        L.oracle_step_lanes_range.restype = None
        L.oracle_step_lanes_range.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32,
                                              C.c_float, C.c_int]
        L.oracle_update.restype = None
        L.oracle_update.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_float, C.c_uint32, C.c_int]
        L.oracle_acc_f64.restype = None
        L.oracle_acc_f64.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.oracle_diagnostics_f64.restype = None
        L.oracle_diagnostics_f64.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]
        L.oracle_max_threads.restype = C.c_int

    def threads(self):
        return int(self.lib.oracle_max_threads())

    def partition(self, ps):
        """Returns (partitioned copy, mass_len)."""
        arr = np.ascontiguousarray(ps, dtype=PARTICLE).copy()
        m = self.lib.oracle_partition(_ptr(arr), len(arr))
        return arr, int(m)

    def partition_int(self, xs):
        arr = np.ascontiguousarray(xs, dtype=np.int32).copy()
        m = self.lib.oracle_partition_int(_ptr(arr), len(arr))
        return arr, int(m)

    def update(self, arr, mass_len, dt, n=1, lanes=8):
        """n Jacobi steps on a partitioned array (copy returned)."""
        out = np.ascontiguousarray(arr, dtype=PARTICLE).copy()
        self.lib.oracle_update(_ptr(out), len(out), mass_len, dt, n, lanes)
        return out

    def step_range(self, arr, mass_len, first, last, dt, lanes=8):
        out = np.ascontiguousarray(arr, dtype=PARTICLE).copy()
        self.lib.oracle_step_lanes_range(_ptr(out), mass_len, first, last, dt, lanes)
        return out

    def acc_f64(self, arr, mass_len, first=0, last=None):
        arr = np.ascontiguousarray(arr, dtype=PARTICLE)
        last = len(arr) if last is None else last
        acc = np.empty((last - first, 2), dtype=np.float64)
        nrm = np.empty(last - first, dtype=np.float64)
        self.lib.oracle_acc_f64(_ptr(arr), mass_len, first, last, _ptr(acc), _ptr(nrm))
        return acc, nrm

    def diagnostics(self, arr, mass_len):
        arr = np.ascontiguousarray(arr, dtype=PARTICLE)
        out = np.empty(7, dtype=np.float64)
        self.lib.oracle_diagnostics_f64(_ptr(arr), len(arr), mass_len, _ptr(out))
        return dict(P=out[0:2].copy(), K=out[2], U=out[3], E=out[2] + out[3], M=out[4], C=out[5:7].copy())


class RefLib:
    """The reference's own code (include/nbody.h API + MakeGalaxies), CPU half only."""

    FLAVOURS = {"avx": "libnbody_ref.so", "sse": "libnbody_ref_sse.so", "scalar": "libnbody_ref_scalar.so"}

    @classmethod
    def available(cls, flavour="avx"):
        return os.path.exists(os.path.join(ORACLE_DIR, "_ref", cls.FLAVOURS[flavour]))

    def __init__(self, flavour="avx"):
        path = os.path.join(ORACLE_DIR, "_ref", self.FLAVOURS[flavour])
        if not os.path.exists(path):
            build_oracle()
        self.lib = L = C.CDLL(path)
        self.libc = C.CDLL(None)
        L.CreateWorld.restype = C.c_void_p
        L.CreateWorld.argtypes = [C.c_void_p, C.c_uint32]
        L.DestroyWorld.restype = None
        L.DestroyWorld.argtypes = [C.c_void_p]
        L.GetWorldParticles.restype = C.c_void_p
        L.GetWorldParticles.argtypes = [C.c_void_p, C.POINTER(C.c_uint32)]
        L.UpdateWorld_CPU.restype = None
        L.UpdateWorld_CPU.argtypes = [C.c_void_p, C.c_float, C.c_uint32]
        L.MakeGalaxies.restype = C.c_void_p
        L.MakeGalaxies.argtypes = [C.c_uint32, C.c_uint32]
        L.refglue_step_range.restype = None
        L.refglue_step_range.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_float]
        L.refglue_simd_width.restype = C.c_int
        L.refglue_max_threads.restype = C.c_int
        self.libc.free.argtypes = [C.c_void_p]
        self.libc.srand.argtypes = [C.c_uint]

    def threads(self):
        return int(self.lib.refglue_max_threads())

    def make_galaxies(self, n, galaxies, seed=11037):
        """srand(seed); MakeGalaxies(n, galaxies) -- the reference generator (galaxy.c:31-221)."""
        self.libc.srand(seed)
        p = self.lib.MakeGalaxies(n, galaxies)
        out = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n * 32,)).view(PARTICLE).copy()
        self.libc.free(p)
        return out

    def world_state(self, ps):
        """CreateWorld(ps) then GetWorldParticles: the partitioned order the reference exposes."""
        ps = np.ascontiguousarray(ps, dtype=PARTICLE)
        w = self.lib.CreateWorld(_ptr(ps), len(ps))
        out = self._get(w)
        self.lib.DestroyWorld(w)
        return out

    def _get(self, w):
        n = C.c_uint32(0)
        p = self.lib.GetWorldParticles(w, C.byref(n))
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value * 32,)).view(PARTICLE).copy()

    def update_cpu(self, ps, dt, n=1, calls=1):
        """CreateWorld(ps); UpdateWorld_CPU(w, dt, n) x calls; GetWorldParticles."""
        ps = np.ascontiguousarray(ps, dtype=PARTICLE)
        w = self.lib.CreateWorld(_ptr(ps), len(ps))
        for _ in range(calls):
            self.lib.UpdateWorld_CPU(w, dt, n)
        out = self._get(w)
        self.lib.DestroyWorld(w)
        return out

    def step_range(self, arr, mass_len, first, last, dt):
        """One reference step restricted to targets [first, last) (oracle/ref_glue.c)."""
        out = np.ascontiguousarray(arr, dtype=PARTICLE).copy()
        self.lib.refglue_step_range(_ptr(out), mass_len, first, last, dt)
        return out


def mass_len_of(arr):
    """Number of leading particles with mass > 0 in a partitioned array."""
    return int(np.count_nonzero(arr["mass"] > 0))


def rel_acc_error(a, b):
    """Per-particle ||a-b|| / ||b|| on (n,2) arrays; 0 where both are exactly equal."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    d = np.hypot(a[:, 0] - b[:, 0], a[:, 1] - b[:, 1])
    nb = np.hypot(b[:, 0], b[:, 1])
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.where(d == 0, 0.0, d / nb)
    return r


def acc_of(arr):
    return np.stack([arr["ax"], arr["ay"]], axis=1)
