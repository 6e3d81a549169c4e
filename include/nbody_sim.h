/*
 * nbody_sim.h -- C-ABI of the CUDA simulation pipeline (libnbody_sim / libnbody.so).
 *
 * Section 1 is, name for name and argument for argument, the seam the reference's
 * world.c binds to reach its GPU simulator (ref src/lib/sim_gpu.h:8-36).  The
 * reference implements it with Vulkan (src/lib/sim_gpu.c + vulkan_ctx.c + the GLSL
 * shader); here it is a CUDA context with device-resident SoA buffers and the sm_100a
 * force+integrator kernel.  Because the signatures are Vulkan-free plain C, the
 * reference's own world.c links against this library unchanged.
 *
 * Section 2 are additions the reference has no counterpart for (device choice, target
 * sharding across GPUs, asynchronous stepping, pinned host memory, introspection).
 * Plain pointers and sizes only; no CUDA, torch or NCCL types appear in a signature.
 *
 * Error convention (ref src/lib/util.h:17-29): no return codes; any CUDA/NCCL failure
 * prints "file:line [func] ..." to stderr and abort()s.  There is no CPU fallback:
 * creating a pipeline on a machine without a usable B200-class device aborts.
 */
#ifndef NB_SIM_H
#define NB_SIM_H

#include <stddef.h>
#include <stdint.h>

#include "nbody.h"

#ifdef __cplusplus
extern "C" {
#endif

/* ---------------- Section 1: the reference seam (sim_gpu.h) ------------------------ */

/* Kernel parameters; only dt may change after creation.  (ref sim_gpu.h:8-12) */
typedef struct WorldData {
    uint32_t total_len; /* all particles = targets                         */
    uint32_t mass_len;  /* leading particles with mass > 0 = sources        */
    float dt;           /* ignored at creation; passed per PerformSimUpdate */
} WorldData;

typedef struct SimPipeline SimPipeline; /* ref sim_gpu.h:15 */

/* Allocate device state for DATA.total_len particles (ref sim_gpu.h:21).  The number of
 * GPUs driven by this process comes from env NBODY_NGPUS (default 1), the first device
 * from NBODY_DEVICE (default 0): the public API has no parameter for either. */
SimPipeline *CreateSimPipeline(WorldData data);

/* Free everything; NULL is ignored.  (ref sim_gpu.h:24) */
void DestroySimPipeline(SimPipeline *sim);

/* Device -> PS: total_len Particles in World order.  (ref sim_gpu.h:27) */
void GetSimulationData(const SimPipeline *sim, Particle *ps);

/* PS -> device: total_len Particles in World order (massive first).  (ref sim_gpu.h:30)
 * With several GPUs (NBODY_NGPUS, or one process per GPU) every GPU copies only the particles it
 * owns out of PS and the source data is replicated GPU to GPU; every rank passes the same array. */
void SetSimulationData(SimPipeline *sim, const Particle *ps);

/* N > 0 chained steps of size DT; returns after the device finished.  Data MUST have
 * been set before.  (ref sim_gpu.h:32-36, sim_gpu.c:258-361) */
void PerformSimUpdate(SimPipeline *sim, uint32_t n, float dt);

/* ---------------- Section 2: additions --------------------------------------------- */

#define NB_SIM_UNIQUE_ID_BYTES 128

/* Optional creation parameters; zero-initialise, set struct_size, fill what you need. */
typedef struct NbSimConfig {
    uint32_t struct_size;  /* sizeof(NbSimConfig), for forward compatibility            */
    int32_t device;        /* first CUDA device ordinal; < 0: env NBODY_DEVICE or 0      */
    int32_t local_gpus;    /* GPUs driven by THIS process; <= 0: env NBODY_NGPUS or 1    */
    int32_t rank;          /* this process' index among `world` processes (default 0)   */
    int32_t world;         /* processes sharing the particle set (default 1)             */
    const void *unique_id; /* NB_SIM_UNIQUE_ID_BYTES from NbSimGetUniqueId on rank 0;    */
                           /* required iff world > 1                                     */
    void *stream;          /* cudaStream_t to run on (single-GPU only); NULL: own stream */
    uint32_t slab_len;     /* sources per partial-sum slab; 0: automatic                 */
    uint32_t flags;        /* NB_SIM_FLAG_*                                              */
} NbSimConfig;

#define NB_SIM_FLAG_NO_PEER_STORE 1u /* single process, several GPUs: exchange positions with   */
                                     /* NCCL instead of peer stores from the kernel epilogue   */
#define NB_SIM_FLAG_NO_OVERLAP    2u /* do not overlap the position exchange with the massless  */
                                     /* targets' kernel                                         */
#define NB_SIM_FLAG_NO_GRAPH      8u /* never replay chained steps of a small World as CUDA-graph    */
                                     /* chunks (env NBODY_NO_GRAPH=1 does the same)                   */
#define NB_SIM_FLAG_EXACT        16u /* verification mode: reproduce the reference's AVX CPU update    */
                                     /* BIT FOR BIT (IEEE sqrt and division, its 8-lane summation      */
                                     /* order; exact_kernel.cuh) at about a quarter of the fast        */
                                     /* kernels' rate.  env NBODY_EXACT=avx|sse|scalar does the same   */
                                     /* for the reference's three SIMD builds                          */
#define NB_SIM_FLAG_NO_RESIDENT  32u /* never run chained steps of a tiny World as one launch of one     */
                                     /* thread-block cluster with the state resident on chip             */
                                     /* (resident_kernel.cuh; env NBODY_NO_RESIDENT=1 does the same)    */
#define NB_SIM_FLAG_PERSIST       4u /* run n >= 2 chained steps of a small World as ONE cooperative */
                                     /* launch with a grid barrier between steps (env NBODY_PERSIST=1  */
                                     /* does the same).  Bit-identical to n launches and measured       */
                                     /* slower on B200 (persist_kernel.cuh): off by default             */

SimPipeline *NbSimCreate(WorldData data, const NbSimConfig *cfg);

/* Fill OUT (NB_SIM_UNIQUE_ID_BYTES) with a fresh NCCL unique id; rank 0 calls this and
 * ships the bytes to the other ranks by its own means (torch.distributed, MPI, a file). */
void NbSimGetUniqueId(void *out);

/* Enqueue N steps and return immediately / wait for everything enqueued so far. */
void NbSimStepAsync(SimPipeline *sim, uint32_t n, float dt);
void NbSimSynchronize(SimPipeline *sim);

/* Device-side time of the most recent NbSimStepAsync/PerformSimUpdate call in ms (CUDA
 * events on the launching stream, max over this process' GPUs); synchronises. */
float NbSimLastStepMs(SimPipeline *sim);

/* One process per GPU: like GetSimulationData, but only the particles owned by THIS process' GPUs are
 * written to PS (same indices as in the full array; the rest of PS is left alone) and nothing is gathered
 * from the other ranks -- the sharded read-back of a sharded World.  NbSimLocalRanges returns those index
 * ranges: up to 2 per GPU of this process, OUT[2k], OUT[2k+1] = [first, last); returns the number of ranges. */
void NbSimGetDataLocal(SimPipeline *sim, Particle *ps);
uint32_t NbSimLocalRanges(const SimPipeline *sim, uint32_t *out, uint32_t max_ranges);

/* Page-locked host memory for Particle arrays, so uploads/downloads run at link speed. */
void *NbSimHostAlloc(size_t bytes);
void NbSimHostFree(void *p);

typedef struct NbSimInfo {
    char device_name[64];
    int32_t device;         /* first device ordinal                       */
    int32_t local_gpus;     /* GPUs driven by this process                */
    int32_t rank, world;    /* process grid                               */
    int32_t sm_count;       /* multiprocessors on the first device        */
    int32_t sm_clock_khz;   /* max SM clock reported by the driver        */
    int32_t cc_major, cc_minor;
    uint32_t total_len, mass_len;
    uint32_t tile_targets;  /* targets per CTA (threads x targets/thread) */
    uint32_t threads;       /* threads per CTA                            */
    uint32_t slab_len;      /* sources per slab                           */
    uint32_t slabs;         /* slabs per target tile                      */
    uint32_t tiles_local;   /* target tiles computed by this process      */
    uint64_t kernels_launched; /* kernels this pipeline has launched so far */
    uint64_t pairs_per_step;   /* total_len * mass_len                      */
    uint32_t group_slabs;   /* consecutive slabs one CTA walks (1: a CTA per slab); partials per target = */
                            /* ceil(slabs / group_slabs)                                                  */
    uint32_t persistent_ctas;     /* NB_SIM_FLAG_PERSIST only: largest grid that runs n >= 2 chained steps as  */
                                  /* ONE cooperative launch; 0: every step is its own launch (the default)     */
    uint64_t persistent_launches; /* such launches so far (they are counted in kernels_launched too)            */
    char variant[16];             /* kernel variant of the plan (NbSimPlanInfo.variant)                         */
    uint64_t graph_launches;      /* cudaGraphLaunch calls so far: chained steps of small Worlds are replayed   */
    uint32_t graph_steps;         /* as CUDA-graph chunks of this many steps (0: this World never uses graphs)  */
    uint32_t resident_ctas;       /* CTAs of the one cluster that runs n >= 2 chained steps of this (tiny) World */
                                  /* in one launch, state resident on chip (resident_kernel.cuh); 0: not used   */
    uint64_t resident_launches;   /* such launches so far (counted in kernels_launched too)                     */
    uint32_t cluster_split;       /* NbSimPlanInfo.cluster_split of this pipeline                               */
} NbSimInfo;

void NbSimGetInfo(const SimPipeline *sim, NbSimInfo *out);

/* The work decomposition the pipeline would use for these sizes, computed on the host alone (no
 * device needed).  TILE_RANGES, when not NULL, receives 4 values per shard q in [0, SHARDS):
 * massive target tiles [a0, a1) and massless target tiles [b0, b1); a tile is TILE_TARGETS
 * consecutive particles in World order.  SM_COUNT <= 0 means a B200 (148). */
typedef struct NbSimPlanInfo {
    char variant[16];        /* kernel variant name                               */
    uint32_t threads;        /* threads per CTA                                   */
    uint32_t tile_targets;   /* targets per tile (= per CTA)                      */
    uint32_t tiles;          /* ceil(total_len / tile_targets)                    */
    uint32_t massive_tiles;  /* tiles containing at least one massive particle    */
    uint32_t slab_len;       /* sources per partial-sum slab                      */
    uint32_t slabs;          /* slabs per tile; fixes the summation order         */
    uint32_t ctas_per_sm;    /* resident CTAs per SM the plan assumes             */
    uint32_t group_slabs;    /* consecutive slabs one CTA walks and sums in slab order itself (1: one CTA per  */
                             /* slab).  > 1 only where one partial per (slab, target) would exceed the budget */
    uint32_t groups;         /* CTAs per tile = partial sums per target = ceil(slabs / group_slabs); they are */
                             /* added in group order                                                           */
    uint32_t resident_ctas;  /* one GPU, tiny Worlds: n >= 2 chained steps run as ONE launch of one cluster of */
                             /* this many CTAs with the state resident on chip (resident_kernel.cuh); 0: no   */
    uint32_t resident_rows;  /* target rows (32 targets) per CTA of that cluster                               */
    uint32_t resident_warps; /* warps per CTA of that cluster                                                  */
    uint32_t cluster_split;  /* small Worlds with few tiles: CTAs (one thread-block cluster) that share the    */
                             /* slabs of a tile (split_kernel.cuh); 1: one CTA per tile                        */
} NbSimPlanInfo;

void NbSimPlan(uint32_t total_len, uint32_t mass_len, int32_t shards, int32_t sm_count, NbSimPlanInfo *out,
               uint32_t *tile_ranges);

/* Number of CUDA devices visible to the process (0 when there is no driver/GPU). */
int32_t NbSimDeviceCount(void);

/* FP64 diagnostics computed on the device from the current state (drift harness):
 * out[0..1] = sum m v, out[2] = kinetic, out[3] = potential (massive pairs, target
 * softening), out[4] = sum m, out[5..6] = sum m x.  Synchronises. */
void NbSimDiagnostics(SimPipeline *sim, double *out7);

/* Headless stand-in for the reference GUI's DrawParticles / CreateCamera (ref src/main.c:198-250):
 * a CUDA point splat of the current device state into a WIDTH x HEIGHT RGBA8 image (row-major, top row
 * first, 4 bytes per pixel) copied to RGBA.  Every particle is a filled disc centred on its position
 * truncated to an integer, radius max(radius, MIN_RADIUS) world units, colour by mass class, painted in
 * array order (higher index wins); screen = (world - target) * zoom + (WIDTH/2, HEIGHT/2).
 * NbSimFitCamera returns the camera that fits all particles into 90 % of the image (zoom <= 1), like the
 * reference's CreateCamera.  Both synchronise; with several processes every rank must call them. */
void NbSimFitCamera(SimPipeline *sim, int32_t width, int32_t height, float *target_x, float *target_y, float *zoom);
void NbSimRenderRGBA(SimPipeline *sim, int32_t width, int32_t height, float target_x, float target_y, float zoom,
                     float min_radius, uint8_t *rgba);

/* FP64 accelerations of targets [first, last) from the current device state, written
 * as (ax, ay) doubles to OUT (2 * (last - first) values).  Does not advance the state. */
void NbSimAccF64(SimPipeline *sim, uint32_t first, uint32_t last, double *out);

/* Diagnostics for tools/step_timeline.py: the kernel variants built with the timeline switch (NBODY_VARIANT=b16x2t ...)
 * record %globaltimer (ns) per warp at four points of a step -- after the dependency wait, sources in shared memory,
 * last unit done, end.  Copies [warps][4] values of the last such launch on the current device; returns the warps. */
uint32_t NbSimDebugTimeline(uint64_t *out, uint32_t max_warps);

/*
 * One step of the hot path WITHOUT any device-memory allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  No pipeline object: D_IN and D_OUT are the caller's own device-accessible arrays of
 * TOTAL_LEN 32-byte Particles in World order (the MASS_LEN massive particles first); D_OUT receives the state after
 * one step of length DT (pos, vel, acc; mass and radius copied) and must not be D_IN (a step reads OLD positions).
 * The work is enqueued on STREAM (a cudaStream_t; NULL: the default stream) of the CURRENT device and not waited for.
 *
 * The scratch the kernel needs (slab partials and arrival tickets, nbody_b200/csrc/noalloc.cu) is a static __device__
 * array of 4 MiB + 256 B that every GPU which loads the library holds until exit; it limits the call to
 * NB_NOALLOC_MAX_PARTICLES particles.  The function calls no cudaMalloc, cudaFree or allocating library routine.
 * All calls share the array: calls that overlap on the GPU (two streams, two host threads) race on it.  The first call
 * on each device looks the array's device address up (cudaGetSymbolAddress) and should be made outside graph capture.
 *
 * Returns 0 on success; 1 for arguments it rejects (more than NB_NOALLOC_MAX_PARTICLES particles, mass_len >
 * total_len, null or identical arrays); 2 for a CUDA error.  A message goes to stderr; the process is not aborted.
 * Same arithmetic as the other kernels (force_kernel.cuh) with slabs of 512 sources: agrees with UpdateWorld_GPU to
 * FP32 rounding (tests/test_zz_no_alloc.py), not bit for bit.
 */
#define NB_NOALLOC_MAX_PARTICLES 16384u
int32_t NbStepNoAlloc(const Particle *d_in, Particle *d_out, uint32_t total_len, uint32_t mass_len, float dt,
                      void *stream);

#ifdef __cplusplus
}
#endif

#endif /* NB_SIM_H */
