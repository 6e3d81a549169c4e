/*
 * noalloc.cu -- NbStepNoAlloc: one step of the hot path for callers that cannot allocate device memory (code under CUDA
 * graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Nothing in the other entry points goes through this file.
 *
 * The kernel is a variant of step_kernel.cuh's (target tile, slab)-per-CTA scheme -- the slab partials meet in scratch
 * memory and the last CTA of a tile to arrive (atomic ticket) adds them in slab order and integrates -- with that
 * scratch in a file-scope static __device__ array of a fixed, modest size instead of a cudaMalloc'd buffer:
 *
 *     NB_NOALLOC_MAX_PARTICLES = 16 384 particles, slabs of 512 sources  =>  at most 32 slabs x 16 384 targets x 8 B
 *     = 4 MiB of partials + 64 tickets; every GPU that loads the library holds these 4 MiB + 256 B until exit.
 *
 * The caller passes the input and output particle arrays (device-accessible, 32-byte AoS, World order: the mass_len
 * massive particles first).  The pair arithmetic is force_kernel.cuh's (interact_pack: packed f32x2, rsqrt.approx,
 * even/odd chains, slabs added in slab order), the integrator the reference's (ref src/lib/sim_cpu.c:191-193).
 *
 * In host code the array's NAME is a host-side stand-in, not a device address: the address the kernel gets comes from
 * cudaGetSymbolAddress, looked up once per device with that device current.  Every call shares the array and it keeps
 * the last call's contents: the tickets are cleared at the start of every call (on the caller's stream), the partials
 * need no clearing (every slot that is read was written by this call first).  Calls that overlap on the GPU -- two
 * streams, or two host threads -- race on it.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <mutex>

#include "force_kernel.cuh"
#include "nbody.h"
#include "nbody_sim.h"

namespace nb {

constexpr uint32_t kNoAllocMax = NB_NOALLOC_MAX_PARTICLES; /* particles */
constexpr uint32_t kNoAllocTile = 256;                     /* targets per CTA, one per thread */
constexpr uint32_t kNoAllocSlab = 512;                     /* sources per slab: chains of 256 terms */
constexpr uint32_t kNoAllocSlabPacks = kNoAllocSlab / 4;
constexpr uint32_t kNoAllocMaxSlabs = kNoAllocMax / kNoAllocSlab;
constexpr uint32_t kNoAllocMaxTiles = kNoAllocMax / kNoAllocTile;
constexpr size_t kNoAllocPartBytes = (size_t)kNoAllocMaxSlabs * kNoAllocMax * sizeof(float2);
constexpr size_t kNoAllocTicketBytes = 256; /* kNoAllocMaxTiles x 4 B */
static_assert(kNoAllocMaxTiles * sizeof(unsigned) <= kNoAllocTicketBytes, "tickets do not fit");

/* [kNoAllocMaxSlabs][kNoAllocMax] float2 partials, then kNoAllocMaxTiles tickets */
static __device__ __align__(256) unsigned char nb_noalloc_scratch[kNoAllocPartBytes + kNoAllocTicketBytes];

struct NoAllocParams {
    const Particle *in;
    Particle *out;
    float2 *part;      /* scratch: part[slab * kNoAllocMax + i] */
    unsigned *tickets; /* scratch: one per target tile, zero at kernel start */
    uint32_t total_len, mass_len, nslab;
    float dt;
};

__global__ void __launch_bounds__(kNoAllocTile) noalloc_step_kernel(const NoAllocParams p) {
    __shared__ float4 s_xy[kNoAllocSlabPacks * 2]; /* per pack: xs, ys */
    __shared__ float4 s_gm[kNoAllocSlabPacks];
    __shared__ int s_last;

    const uint32_t tid = threadIdx.x;
    const uint32_t slab = blockIdx.x % p.nslab, tile = blockIdx.x / p.nslab;
    const uint32_t i = tile * kNoAllocTile + tid;
    const bool live = i < p.total_len;

    /* this thread's target */
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(0.f, 0.f, 0.f, 1.f); /* pos vel | acc mass radius */
    if (live) {
        a = reinterpret_cast<const float4 *>(p.in)[2 * (size_t)i];
        b = reinterpret_cast<const float4 *>(p.in)[2 * (size_t)i + 1];
    }

    /* this CTA's slab of sources: AoS -> packs of 4 in shared memory (x[4], y[4], G*m[4]); lanes past mass_len are
     * padding sources far away with G*m = 0: their contribution is exactly +-0 (force_kernel.cuh kPadPos) */
    const uint32_t j0 = slab * kNoAllocSlab;
    const uint32_t cnt = min(kNoAllocSlab, p.mass_len > j0 ? p.mass_len - j0 : 0u);
    const uint32_t npk = (cnt + 3u) / 4u;
    for (uint32_t k = tid; k < npk * 4u; k += kNoAllocTile) {
        float sx = kPadPos, sy = kPadPos, sg = 0.f;
        if (k < cnt) {
            const float4 sa = reinterpret_cast<const float4 *>(p.in)[2 * (size_t)(j0 + k)];
            const float4 sb = reinterpret_cast<const float4 *>(p.in)[2 * (size_t)(j0 + k) + 1];
            sx = sa.x;
            sy = sa.y;
            sg = __fmul_rn(sb.z, NB_G); /* one rounding, like the reference's per-pair m * G (ref sim_cpu.c:179) */
        }
        const uint32_t q = k >> 2, l = k & 3u;
        reinterpret_cast<float *>(&s_xy[2 * q])[l] = sx;
        reinterpret_cast<float *>(&s_xy[2 * q + 1])[l] = sy;
        reinterpret_cast<float *>(&s_gm[q])[l] = sg;
    }
    __syncthreads();

    float2 ntx[1] = {make_float2(-a.x, -a.x)}, nty[1] = {make_float2(-a.y, -a.y)}, rho[1] = {make_float2(b.w, b.w)};
    float2 ax[1] = {make_float2(0.f, 0.f)}, ay[1] = {make_float2(0.f, 0.f)};
#pragma unroll 2
    for (uint32_t q = 0; q < npk; q++) interact_pack<1, kFormPacked>(s_xy[2 * q], s_xy[2 * q + 1], s_gm[q], ntx, nty, rho, ax, ay);
    float2 acc = make_float2(__fadd_rn(ax[0].x, ax[0].y), __fadd_rn(ay[0].x, ay[0].y));

    if (p.nslab > 1) {
        /* slab partial -> scratch; the last CTA of the tile adds all of them in slab order */
        __stcg(&p.part[(size_t)slab * kNoAllocMax + i], acc);
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned ticket = atomicAdd(&p.tickets[tile], 1u);
            s_last = (ticket == p.nslab - 1);
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        acc = __ldcg(&p.part[i]);
        for (uint32_t s = 1; s < p.nslab; s++) {
            const float2 v = __ldcg(&p.part[(size_t)s * kNoAllocMax + i]);
            acc.x = __fadd_rn(acc.x, v.x);
            acc.y = __fadd_rn(acc.y, v.y);
        }
    }
    if (!live) return;
    /* vel += acc * dt; pos += vel * dt with the new vel; every product and sum rounded on its own */
    const float vx = __fadd_rn(a.z, __fmul_rn(acc.x, p.dt)), vy = __fadd_rn(a.w, __fmul_rn(acc.y, p.dt));
    const float px = __fadd_rn(a.x, __fmul_rn(vx, p.dt)), py = __fadd_rn(a.y, __fmul_rn(vy, p.dt));
    reinterpret_cast<float4 *>(p.out)[2 * (size_t)i] = make_float4(px, py, vx, vy);
    reinterpret_cast<float4 *>(p.out)[2 * (size_t)i + 1] = make_float4(acc.x, acc.y, b.z, b.w);
}

/* device address of the scratch array, per device ordinal; looked up once, with that device current */
constexpr int kNoAllocMaxDevices = 64;
static unsigned char *g_noalloc_ptr[kNoAllocMaxDevices];
static std::mutex g_noalloc_mutex;

}  // namespace nb

#define NB_NOALLOC_FAIL(CODE, ...)                                        \
    do {                                                                  \
        fprintf(stderr, "%s:%d [%s] ", __FILE__, __LINE__, __func__);     \
        fprintf(stderr, __VA_ARGS__);                                     \
        fprintf(stderr, "\n");                                            \
        return (CODE);                                                    \
    } while (0)

extern "C" int32_t NbStepNoAlloc(const Particle *d_in, Particle *d_out, uint32_t total_len, uint32_t mass_len, float dt,
                                 void *stream) {
    using namespace nb;
    if (total_len > kNoAllocMax)
        NB_NOALLOC_FAIL(1, "NbStepNoAlloc: %u particles, the static scratch array is sized for at most %u "
                        "(NB_NOALLOC_MAX_PARTICLES); use UpdateWorld_GPU / PerformSimUpdate for larger Worlds",
                        total_len, kNoAllocMax);
    if (mass_len > total_len) NB_NOALLOC_FAIL(1, "NbStepNoAlloc: mass_len %u > total_len %u", mass_len, total_len);
    if (total_len == 0) return 0;
    if (!d_in || !d_out || d_in == d_out)
        NB_NOALLOC_FAIL(1, "NbStepNoAlloc: input and output must be two distinct device arrays (a step reads OLD positions)");

    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) NB_NOALLOC_FAIL(2, "NbStepNoAlloc: cudaGetDevice: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= kNoAllocMaxDevices) NB_NOALLOC_FAIL(2, "NbStepNoAlloc: device ordinal %d not supported", dev);
    unsigned char *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_noalloc_mutex);
        if (!g_noalloc_ptr[dev]) { /* set-up, once per device, this device current: NOT the array's name */
            void *ptr = nullptr;
            e = cudaGetSymbolAddress(&ptr, nb_noalloc_scratch);
            if (e != cudaSuccess || !ptr)
                NB_NOALLOC_FAIL(2, "NbStepNoAlloc: cudaGetSymbolAddress on device %d: %s", dev, cudaGetErrorString(e));
            g_noalloc_ptr[dev] = static_cast<unsigned char *>(ptr);
        }
        scratch = g_noalloc_ptr[dev];
    }

    NoAllocParams p;
    p.in = d_in;
    p.out = d_out;
    p.part = reinterpret_cast<float2 *>(scratch);
    p.tickets = reinterpret_cast<unsigned *>(scratch + kNoAllocPartBytes);
    p.total_len = total_len;
    p.mass_len = mass_len;
    p.nslab = mass_len ? (mass_len + kNoAllocSlab - 1) / kNoAllocSlab : 1u;
    p.dt = dt;
    const uint32_t tiles = (total_len + kNoAllocTile - 1) / kNoAllocTile;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    /* the array keeps the last call's contents: the tickets must start at zero (a call that died half-way, or a
     * replayed graph, must not leave a count behind) */
    e = cudaMemsetAsync(p.tickets, 0, kNoAllocTicketBytes, st);
    if (e != cudaSuccess) NB_NOALLOC_FAIL(2, "NbStepNoAlloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
    noalloc_step_kernel<<<tiles * p.nslab, kNoAllocTile, 0, st>>>(p);
    e = cudaGetLastError();
    if (e != cudaSuccess) NB_NOALLOC_FAIL(2, "NbStepNoAlloc: kernel launch: %s", cudaGetErrorString(e));
    return 0;
}
