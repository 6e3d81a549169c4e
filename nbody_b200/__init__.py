"""nbody_b200 -- Python mirror of the World API (include/nbody.h) over libnbody.so.

The product is the C library: host code in C (csrc/world.c, csrc/galaxy.c) calling hand-written
sm_100a CUDA through the C-ABI seam of include/nbody_sim.h.  This module only binds it with ctypes
so tests and bench.py can call it exactly as a C program would: same function names, argument
meaning and abort-on-error behaviour as the reference header (ref include/nbody.h:61-73,
include/galaxy.h:64, src/lib/sim_gpu.h:21-36).  There is no Python or CPU implementation of the GPU
path here; if libnbody.so is missing, importing the bindings raises.
"""
from .api import (  # noqa: F401
    PARTICLE,
    PinnedParticles,
    LIB_PATH,
    CreateWorld,
    DestroyWorld,
    GetWorldParticles,
    MakeGalaxies,
    SimConfig,
    SimPipeline,
    UpdateWorld_CPU,
    UpdateWorld_GPU,
    World,
    device_count,
    get_unique_id,
    lib,
    make_galaxies,
    plan,
    write_ppm,
)

__all__ = [
    "PARTICLE", "PinnedParticles", "LIB_PATH", "CreateWorld", "DestroyWorld", "GetWorldParticles", "MakeGalaxies",
    "SimConfig", "SimPipeline", "UpdateWorld_CPU", "UpdateWorld_GPU", "World", "device_count",
    "get_unique_id", "lib", "make_galaxies", "plan", "write_ppm",
]
