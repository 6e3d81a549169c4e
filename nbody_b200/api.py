"""ctypes bindings of libnbody.so (see package docstring)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libnbody.so")

# include/nbody.h Particle: V2 pos, vel, acc; float mass, radius -- 32 bytes, ABI
PARTICLE = np.dtype(
    [("px", "<f4"), ("py", "<f4"), ("vx", "<f4"), ("vy", "<f4"),
     ("ax", "<f4"), ("ay", "<f4"), ("mass", "<f4"), ("radius", "<f4")]
)
assert PARTICLE.itemsize == 32

UNIQUE_ID_BYTES = 128
FLAG_NO_PEER_STORE = 1
FLAG_NO_OVERLAP = 2
FLAG_PERSIST = 4
FLAG_NO_GRAPH = 8
FLAG_EXACT = 16
FLAG_NO_RESIDENT = 32


class WorldData(C.Structure):  # include/nbody_sim.h (ref sim_gpu.h:8-12)
    _fields_ = [("total_len", C.c_uint32), ("mass_len", C.c_uint32), ("dt", C.c_float)]


class SimConfig(C.Structure):  # include/nbody_sim.h NbSimConfig
    _fields_ = [
        ("struct_size", C.c_uint32),
        ("device", C.c_int32),
        ("local_gpus", C.c_int32),
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("unique_id", C.c_void_p),
        ("stream", C.c_void_p),
        ("slab_len", C.c_uint32),
        ("flags", C.c_uint32),
    ]


class SimPlanInfo(C.Structure):  # include/nbody_sim.h NbSimPlanInfo
    _fields_ = [
        ("variant", C.c_char * 16),
        ("threads", C.c_uint32),
        ("tile_targets", C.c_uint32),
        ("tiles", C.c_uint32),
        ("massive_tiles", C.c_uint32),
        ("slab_len", C.c_uint32),
        ("slabs", C.c_uint32),
        ("ctas_per_sm", C.c_uint32),
        ("group_slabs", C.c_uint32),
        ("groups", C.c_uint32),
        ("resident_ctas", C.c_uint32),
        ("resident_rows", C.c_uint32),
        ("resident_warps", C.c_uint32),
        ("cluster_split", C.c_uint32),
    ]


class SimInfo(C.Structure):  # include/nbody_sim.h NbSimInfo
    _fields_ = [
        ("device_name", C.c_char * 64),
        ("device", C.c_int32),
        ("local_gpus", C.c_int32),
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("sm_count", C.c_int32),
        ("sm_clock_khz", C.c_int32),
        ("cc_major", C.c_int32),
        ("cc_minor", C.c_int32),
        ("total_len", C.c_uint32),
        ("mass_len", C.c_uint32),
        ("tile_targets", C.c_uint32),
        ("threads", C.c_uint32),
        ("slab_len", C.c_uint32),
        ("slabs", C.c_uint32),
        ("tiles_local", C.c_uint32),
        ("kernels_launched", C.c_uint64),
        ("pairs_per_step", C.c_uint64),
        ("group_slabs", C.c_uint32),
        ("persistent_ctas", C.c_uint32),
        ("persistent_launches", C.c_uint64),
        ("variant", C.c_char * 16),
        ("graph_launches", C.c_uint64),
        ("graph_steps", C.c_uint32),
        ("resident_ctas", C.c_uint32),
        ("resident_launches", C.c_uint64),
        ("cluster_split", C.c_uint32),
    ]

    def as_dict(self):
        d = {}
        for name, _ in self._fields_:
            v = getattr(self, name)
            d[name] = v.decode() if isinstance(v, bytes) else int(v)
        return d


_lib = None


def lib():
    """Load libnbody.so once.  Raises if the library was not built: there is no fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make` (or __graft_entry__.build()); "
            "nbody_b200 has no Python/CPU fallback for the CUDA path"
        )
    L = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    vp, u32, f32 = C.c_void_p, C.c_uint32, C.c_float
    sigs = {
        # include/nbody.h
        "CreateWorld": (vp, [vp, u32]),
        "DestroyWorld": (None, [vp]),
        "GetWorldParticles": (vp, [vp, C.POINTER(u32)]),
        "UpdateWorld_CPU": (None, [vp, f32, u32]),
        "UpdateWorld_GPU": (None, [vp, f32, u32]),
        # include/galaxy.h
        "MakeGalaxies": (vp, [u32, u32]),
        # include/nbody_sim.h section 1
        "CreateSimPipeline": (vp, [WorldData]),
        "DestroySimPipeline": (None, [vp]),
        "GetSimulationData": (None, [vp, vp]),
        "SetSimulationData": (None, [vp, vp]),
        "PerformSimUpdate": (None, [vp, u32, f32]),
        # include/nbody_sim.h section 2
        "NbSimCreate": (vp, [WorldData, C.POINTER(SimConfig)]),
        "NbSimGetUniqueId": (None, [vp]),
        "NbSimStepAsync": (None, [vp, u32, f32]),
        "NbSimSynchronize": (None, [vp]),
        "NbSimLastStepMs": (f32, [vp]),
        "NbSimHostAlloc": (vp, [C.c_size_t]),
        "NbSimHostFree": (None, [vp]),
        "NbSimGetDataLocal": (None, [vp, vp]),
        "NbSimLocalRanges": (C.c_uint32, [vp, vp, C.c_uint32]),
        "NbSimGetInfo": (None, [vp, C.POINTER(SimInfo)]),
        "NbSimDeviceCount": (C.c_int32, []),
        "NbSimPlan": (None, [u32, u32, C.c_int32, C.c_int32, C.POINTER(SimPlanInfo), vp]),
        "NbSimDiagnostics": (None, [vp, vp]),
        "NbSimAccF64": (None, [vp, u32, u32, vp]),
        "NbSimDebugTimeline": (u32, [vp, u32]),
        "NbStepNoAlloc": (C.c_int32, [vp, vp, u32, u32, f32, vp]),
        "NbSimFitCamera": (None, [vp, C.c_int32, C.c_int32, C.POINTER(f32), C.POINTER(f32), C.POINTER(f32)]),
        "NbSimRenderRGBA": (None, [vp, C.c_int32, C.c_int32, f32, f32, f32, f32, vp]),
        # world.c extras
        "NbWorldPipeline": (vp, [vp]),
        "NbWorldMassLen": (u32, [vp]),
        "NbWorldPrepareGPU": (None, [vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


_libc = C.CDLL(None)
_libc.free.argtypes = [C.c_void_p]
_libc.srand.argtypes = [C.c_uint]


def _as_particles(ps):
    a = np.ascontiguousarray(ps, dtype=PARTICLE)
    return a


def _view(ptr, n):
    if n == 0:
        return np.empty(0, dtype=PARTICLE)
    buf = (C.c_uint8 * (n * 32)).from_address(ptr)
    return np.frombuffer(buf, dtype=PARTICLE)


def plan(total_len, mass_len, shards=1, sm_count=0):
    """Host-only work decomposition (NbSimPlan): returns (info dict, [(a0, a1, b0, b1)] per shard)."""
    info = SimPlanInfo()
    ranges = np.zeros((shards, 4), dtype=np.uint32)
    lib().NbSimPlan(total_len, mass_len, shards, sm_count, C.byref(info), ranges.ctypes.data_as(C.c_void_p))
    d = {name: (getattr(info, name).decode() if name == "variant" else int(getattr(info, name)))
         for name, _ in info._fields_}
    return d, [tuple(int(x) for x in r) for r in ranges]


def device_count():
    return int(lib().NbSimDeviceCount())


# ---- the five World functions, C names (ref include/nbody.h:61-73) ---------------------------

def CreateWorld(ps, size=None):
    a = _as_particles(ps)
    size = len(a) if size is None else size
    return lib().CreateWorld(a.ctypes.data_as(C.c_void_p), size)


def DestroyWorld(w):
    lib().DestroyWorld(w)


def GetWorldParticles(w):
    """Returns a NumPy view of the World-owned array (borrowed; valid until DestroyWorld)."""
    n = C.c_uint32(0)
    p = lib().GetWorldParticles(w, C.byref(n))
    return _view(p, n.value)


def UpdateWorld_CPU(w, dt, n):
    lib().UpdateWorld_CPU(w, dt, n)


def UpdateWorld_GPU(w, dt, n):
    lib().UpdateWorld_GPU(w, dt, n)


def MakeGalaxies(particle_count, galaxy_count):
    """include/galaxy.h MakeGalaxies; returns an owned NumPy copy (the C array is freed)."""
    p = lib().MakeGalaxies(particle_count, galaxy_count)
    out = _view(p, particle_count).copy()
    _libc.free(p)
    return out


def make_galaxies(particle_count, galaxy_count, seed=11037):
    """srand(seed); MakeGalaxies(...) -- the benchmark input definition (ref src/bench.c:42,53)."""
    _libc.srand(seed)
    return MakeGalaxies(particle_count, galaxy_count)


class PinnedParticles:
    """A Particle array in page-locked host memory (NbSimHostAlloc), so SetSimulationData /
    GetSimulationData run at link speed.  `.array` is a NumPy view; freed on close()/GC."""

    def __init__(self, n):
        self._ptr = lib().NbSimHostAlloc(max(n, 1) * 32)
        self.array = _view(self._ptr, n)

    def close(self):
        if self._ptr:
            self.array = None
            lib().NbSimHostFree(self._ptr)
            self._ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_ppm(path, rgba):
    """Binary PPM (P6) of an RGBA image from SimPipeline.render -- the 'window' of a headless box."""
    h, w = rgba.shape[:2]
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(rgba[:, :, :3]).tobytes())


def get_unique_id():
    buf = (C.c_uint8 * UNIQUE_ID_BYTES)()
    lib().NbSimGetUniqueId(buf)
    return bytes(buf)


class World:
    """Object wrapper over the five functions; context manager destroys the World."""

    def __init__(self, ps):
        self._w = CreateWorld(ps)

    def update_gpu(self, dt, n=1):
        UpdateWorld_GPU(self._w, dt, n)

    def update_cpu(self, dt, n=1):
        UpdateWorld_CPU(self._w, dt, n)

    def particles(self, copy=True):
        v = GetWorldParticles(self._w)
        return v.copy() if copy else v

    @property
    def mass_len(self):
        return int(lib().NbWorldMassLen(self._w))

    def pipeline_info(self):
        sim = lib().NbWorldPipeline(self._w)
        if not sim:
            return None
        info = SimInfo()
        lib().NbSimGetInfo(sim, C.byref(info))
        return info.as_dict()

    def close(self):
        if self._w:
            DestroyWorld(self._w)
            self._w = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class SimPipeline:
    """The sim_gpu.h seam + NbSim* additions, for callers that manage the partition themselves."""

    def __init__(self, total_len, mass_len, device=-1, local_gpus=0, rank=0, world=1, unique_id=None,
                 stream=None, slab_len=0, flags=0):
        L = lib()
        self._uid = None
        cfg = SimConfig()
        cfg.struct_size = C.sizeof(SimConfig)
        cfg.device = device
        cfg.local_gpus = local_gpus
        cfg.rank = rank
        cfg.world = world
        if unique_id is not None:
            self._uid = C.create_string_buffer(bytes(unique_id), UNIQUE_ID_BYTES)
            cfg.unique_id = C.cast(self._uid, C.c_void_p)
        cfg.stream = stream
        cfg.slab_len = slab_len
        cfg.flags = flags
        self.total_len = total_len
        self.mass_len = mass_len
        self._sim = L.NbSimCreate(WorldData(total_len, mass_len, 0.0), C.byref(cfg))

    def set_data(self, ps):
        a = _as_particles(ps)
        assert len(a) == self.total_len
        lib().SetSimulationData(self._sim, a.ctypes.data_as(C.c_void_p))

    def get_data(self, out=None):
        if out is None:
            out = np.empty(self.total_len, dtype=PARTICLE)
        lib().GetSimulationData(self._sim, out.ctypes.data_as(C.c_void_p))
        return out

    def get_data_local(self, out):
        """Sharded read-back: only the particles owned by this process' GPUs are written to OUT (full-size array)."""
        assert len(out) == self.total_len and out.dtype == PARTICLE
        lib().NbSimGetDataLocal(self._sim, out.ctypes.data_as(C.c_void_p))
        return out

    def local_ranges(self):
        """[(first, last), ...] particle index ranges owned by this process' GPUs."""
        buf = (C.c_uint32 * 64)()
        n = lib().NbSimLocalRanges(self._sim, buf, 32)
        return [(int(buf[2 * k]), int(buf[2 * k + 1])) for k in range(min(n, 32))]

    def update(self, n, dt):
        lib().PerformSimUpdate(self._sim, n, dt)

    def step_async(self, n, dt):
        lib().NbSimStepAsync(self._sim, n, dt)

    def synchronize(self):
        lib().NbSimSynchronize(self._sim)

    def last_step_ms(self):
        return float(lib().NbSimLastStepMs(self._sim))

    def info(self):
        info = SimInfo()
        lib().NbSimGetInfo(self._sim, C.byref(info))
        return info.as_dict()

    def diagnostics(self):
        out = np.zeros(7, dtype=np.float64)
        lib().NbSimDiagnostics(self._sim, out.ctypes.data_as(C.c_void_p))
        return dict(P=out[0:2].copy(), K=out[2], U=out[3], E=out[2] + out[3], M=out[4], C=out[5:7].copy())

    def fit_camera(self, width, height):
        """(target_x, target_y, zoom) fitting all particles into 90 % of the image (ref main.c:198-225)."""
        tx, ty, zoom = C.c_float(), C.c_float(), C.c_float()
        lib().NbSimFitCamera(self._sim, width, height, C.byref(tx), C.byref(ty), C.byref(zoom))
        return tx.value, ty.value, zoom.value

    def render(self, width, height, camera=None, min_radius=0.0):
        """Headless DrawParticles (ref main.c:237-250): (height, width, 4) uint8 RGBA image."""
        tx, ty, zoom = camera if camera is not None else self.fit_camera(width, height)
        img = np.empty((height, width, 4), dtype=np.uint8)
        lib().NbSimRenderRGBA(self._sim, width, height, tx, ty, zoom, min_radius, img.ctypes.data_as(C.c_void_p))
        return img

    def acc_f64(self, first=0, last=None):
        last = self.total_len if last is None else last
        out = np.empty((last - first, 2), dtype=np.float64)
        lib().NbSimAccF64(self._sim, first, last, out.ctypes.data_as(C.c_void_p))
        return out

    def close(self):
        if self._sim:
            lib().DestroySimPipeline(self._sim)
            self._sim = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
