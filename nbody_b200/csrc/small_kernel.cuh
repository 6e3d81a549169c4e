/*
 * small_kernel.cuh -- the step kernel of SMALL Worlds: the slabs of a target tile are split over the WARPS of one
 * CTA instead of over CTAs.
 *
 * step_kernel.cuh gives every (target tile, slab) its own CTA and adds the slab partials through global memory
 * (partial store, fence, ticket, partial loads by the last CTA): three dependent L2 round trips, ~2.5 us, on top
 * of the first source tile's round trip.  Below N ~ 6 K that chain IS the step (N = 1 000: 6.5 us per step, of
 * which < 0.3 us is arithmetic; the reference's own benchmark sizes 500 ... 4 000, ref src/bench.c:38).  Here
 *
 *   - one CTA = one tile of 32 * T targets; every warp holds the SAME targets (lane l: targets k * 32 + l, k < T);
 *   - ALL sources are staged in shared memory once per CTA (one TMA bulk copy per array, 12 B per source:
 *     24 KB at M = 2 000), no ring, no refills;
 *   - warp w walks slabs w, w + WARPS, ... with the same pair arithmetic and the same even/odd chains as the big
 *     kernel (interact_pack), leaves each slab's even + odd sum in shared memory;
 *   - after one __syncthreads the first 32 * T threads add the slab sums of their target IN SLAB ORDER and
 *     integrate, with pos and vel fetched at kernel entry (off the critical path).
 *
 * The summation order is the plan's (slab boundaries, even/odd chains, slabs in order), exactly as in
 * step_kernel.cuh: with the same slab length the two kernels produce the same bits
 * (tests/test_gpu_parity.py::test_small_world_kernel_equals_step_kernel_bitwise).  The step's dependent chain is
 * now: bulk copy -> pair loop -> shared-memory sums -> stores.
 */
#pragma once

#include "force_kernel.cuh"

namespace nb {

/* dynamic shared memory: [src_packs_pad][2] float4 xy | [src_packs_pad] float4 gm | [nslab][32 T] float2 | mbarrier */
__host__ __device__ inline size_t small_smem_bytes(uint32_t src_packs_pad, uint32_t nslab, int T) {
    return (size_t)src_packs_pad * 48 + (size_t)nslab * 32 * T * sizeof(float2) + 16;
}

/* Diagnostics (kernels built with the timeline switch; tools/step_timeline.py): %globaltimer of every warp after the dependency wait, after the
 * sources landed, after its last unit, and at its end. */
constexpr uint32_t kTimelineWarps = 160 * 32;
__device__ unsigned long long nb_timeline[kTimelineWarps * 4];
__device__ __forceinline__ void timeline_mark(bool on, uint32_t slot, int what) {
    if (on && (threadIdx.x & 31u) == 0 && slot < kTimelineWarps) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        nb_timeline[slot * 4 + what] = t;
    }
}

/* integrate_store (force_kernel.cuh) with pos and vel already in registers */
__device__ __forceinline__ void integrate_store_vx(const StepParams &p, uint32_t i, float2 a, float2 v, float2 x) {
    v.x = __fadd_rn(v.x, __fmul_rn(a.x, p.dt));
    v.y = __fadd_rn(v.y, __fmul_rn(a.y, p.dt));
    x.x = __fadd_rn(x.x, __fmul_rn(v.x, p.dt));
    x.y = __fadd_rn(x.y, __fmul_rn(v.y, p.dt));
    p.acc[i] = a;
    p.vel[i] = v;
    p.pos[i] = x;
    if (i < p.mass_len) {
        const uint32_t o = (i >> 2) * 8u + (i & 3u);
#pragma unroll 1
        for (int d = 0; d < p.ndst; d++) {
            p.dst_xy[d][o] = x.x;
            p.dst_xy[d][o + 4] = x.y;
        }
    }
}

template <int WARPS, int T, int UNR, bool TL = false>
__global__ void __launch_bounds__(WARPS * 32, 1) small_world_kernel(const __grid_constant__ StepParams p) {
    constexpr int TILE = 32 * T;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_xy = reinterpret_cast<float4 *>(smem_raw);                       /* [src_packs_pad][2] */
    float4 *s_gm = s_xy + (size_t)p.src_packs_pad * 2;                         /* [src_packs_pad]    */
    float2 *s_part = reinterpret_cast<float2 *>(s_gm + p.src_packs_pad);       /* [nslab][TILE]      */
    uint64_t *bar = reinterpret_cast<uint64_t *>(s_part + (size_t)p.nslab * TILE);

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t tile = p.tile_first + blockIdx.x;
    const uint32_t t0 = tile * TILE;

    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    /* programmatic dependent launch: nothing global before the previous step's writes are visible */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t tl_slot = blockIdx.x * WARPS + warp;
    timeline_mark(TL, tl_slot, 0);

    if (tid == 0) { /* all sources -> shared memory; bulk copies of at most 32 KB each */
        const uint32_t xy_bytes = p.src_packs_pad * 32u, gm_bytes = p.src_packs_pad * 16u;
        mbar_expect_tx(bar, xy_bytes + gm_bytes);
        for (uint32_t o = 0; o < xy_bytes; o += 32768u)
            bulk_g2s(reinterpret_cast<unsigned char *>(s_xy) + o, reinterpret_cast<const unsigned char *>(p.src_xy) + o,
                     min(32768u, xy_bytes - o), bar);
        for (uint32_t o = 0; o < gm_bytes; o += 32768u)
            bulk_g2s(reinterpret_cast<unsigned char *>(s_gm) + o, reinterpret_cast<const unsigned char *>(p.src_gm) + o,
                     min(32768u, gm_bytes - o), bar);
    }

    /* this lane's targets (the same in every warp), and the state of the target this thread will integrate */
    float2 ntx[T], nty[T], rho[T];
#pragma unroll
    for (int k = 0; k < T; k++) {
        const uint32_t i = t0 + k * 32 + lane;
        float2 x = make_float2(0.f, 0.f);
        float r = 1.f;
        if (i < p.total_len) {
            x = p.pos[i];
            r = p.radius[i];
        }
        ntx[k] = make_float2(-x.x, -x.x);
        nty[k] = make_float2(-x.y, -x.y);
        rho[k] = make_float2(r, r);
    }
    float2 my_v = make_float2(0.f, 0.f), my_x = make_float2(0.f, 0.f);
    const uint32_t my_i = t0 + tid;
    if (tid < (uint32_t)TILE && my_i < p.total_len) {
        my_v = p.vel[my_i];
        my_x = p.pos[my_i];
    }

    mbar_wait(bar, 0);
    timeline_mark(TL, tl_slot, 1);

    for (uint32_t s = warp; s < p.nslab; s += WARPS) {
        const uint32_t pk0 = s * p.slab_packs;
        const uint32_t pk1 = min(pk0 + p.slab_packs, p.src_packs_pad);
        float2 ax[T], ay[T];
#pragma unroll
        for (int k = 0; k < T; k++) {
            ax[k] = make_float2(0.f, 0.f);
            ay[k] = make_float2(0.f, 0.f);
        }
#pragma unroll 1
        for (uint32_t q = pk0; q < pk1; q += UNR) {
#pragma unroll
            for (int u = 0; u < UNR; u++) {
                const float4 xs = s_xy[2 * (q + u)];
                const float4 ys = s_xy[2 * (q + u) + 1];
                const float4 gs = s_gm[q + u];
                interact_pack<T, kFormPacked>(xs, ys, gs, ntx, nty, rho, ax, ay);
            }
        }
#pragma unroll
        for (int k = 0; k < T; k++)
            s_part[(size_t)s * TILE + k * 32 + lane] =
                make_float2(__fadd_rn(ax[k].x, ax[k].y), __fadd_rn(ay[k].x, ay[k].y));
    }

    /* pair loop done: the next step's grid may be scheduled into the slots that free up from here on */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    timeline_mark(TL, tl_slot, 2);
    __syncthreads();

    if (tid < (uint32_t)TILE && my_i < p.total_len) {
        float2 sum = s_part[tid];
        for (uint32_t s = 1; s < p.nslab; s++) {
            const float2 v = s_part[(size_t)s * TILE + tid];
            sum.x = __fadd_rn(sum.x, v.x);
            sum.y = __fadd_rn(sum.y, v.y);
        }
        integrate_store_vx(p, my_i, sum, my_v, my_x);
    }
    timeline_mark(TL, tl_slot, 3);
}

}  // namespace nb
