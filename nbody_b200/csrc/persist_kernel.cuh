/*
 * persist_kernel.cuh -- n chained steps of a SMALL World in ONE launch.
 *
 * UpdateWorld_GPU(w, dt, n) is n kernel launches on one stream (sim_cuda.cu step_async).  Below N ~ 10K a step is
 * shorter than what separates two dependent launches (3.7 us per step at N = 250 with programmatic dependent
 * launch, of which < 0.1 us is arithmetic), and that is exactly the reference's own benchmark protocol: 100 steps
 * per call at N = 250 ... 4 000 (ref src/bench.c:21-38).  Here the whole grid of a step -- one CTA per
 * (target tile, slab), the SAME work plan and therefore the same summation order, bit for bit, as n single
 * launches -- stays resident (cooperative launch) and separates the steps with a grid-wide barrier instead of a
 * kernel boundary:
 *
 *   per step:  TMA ring over this CTA's slab -> slab partial -> ticket -> last CTA of the tile adds the partials
 *              in slab order, integrates, writes pos/vel/acc and the NEXT source buffer -> grid barrier
 *
 * What a kernel boundary gave for free and is explicit here:
 *   - other SMs' L1 may hold last step's pos/vel lines: state that changes between steps is read with ld.global.cg
 *     (L2 is the point of coherence); radius never changes and is loaded once;
 *   - the source tiles are fetched by the async proxy (cp.async.bulk) but were written by ordinary stores of
 *     other CTAs one step earlier: writers execute fence.proxy.async before they arrive on the grid barrier and
 *     the producer thread once more before its first bulk copy of a step;
 *   - the mbarrier ring keeps running across steps (stage and phase follow a tile counter that is never reset).
 *
 * The grid barrier is a monotonic arrival counter in global memory (one red.release per CTA, ld.acquire polling);
 * the host passes the counter's value at launch, so nothing is reset between launches.  All CTAs must be
 * co-resident: the launch is cooperative and sim_cuda.cu only takes this path when the grid fits.
 *
 * MEASURED ON B200 (round 2, profiles/r02_persistent_kernel.txt; nbody-bench protocol, us per step, 100 steps per
 * call): bit-identical to n launches at every size tried -- and SLOWER than them:
 *       N          250    1000    2000    4000    8000
 *   persistent     5.8     8.4     8.5    11.4    19.4
 *   n launches     3.7     6.5     6.3     8.1    18.3      (programmatic dependent launch, step_kernel.cuh)
 * With every fence of the barrier and of the ticket removed (unsafe, timing only) it is still 5.7 / 7.7 / 7.5 /
 * 10.5 / 17.7: on this GPU a kernel boundary with programmatic dependent launch costs LESS than one software grid
 * barrier (a red, then ld.acquire polling across 148 SMs, then every CTA's first L2 round trips all over again --
 * which the dependent launch overlaps with the previous step's tail).  So this kernel is NOT the default: it runs
 * only with NBODY_PERSIST=1 / NB_SIM_FLAG_PERSIST, and is kept as the record of that measurement plus its parity
 * test (tests/test_gpu_parity.py).
 */
#pragma once

#include "force_kernel.cuh"
#include "step_kernel.cuh"

namespace nb {

struct PersistParams {
    StepParams s;       /* s.src_xy: positions read by the FIRST step; s.dst_xy[0]: the other buffer (ping-pong) */
    unsigned *bar;      /* grid barrier: arrivals so far, monotonic (wraps modulo 2^32)                          */
    unsigned bar_base;  /* its value when this launch starts                                                     */
    uint32_t nsteps;    /* >= 1                                                                                   */
};

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned *p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned *p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() {
    asm volatile("fence.proxy.async.global;" ::: "memory");
}

/* integrate_store (force_kernel.cuh) for state that another SM wrote one step ago: L2 loads */
__device__ __forceinline__ void integrate_store_cg(const StepParams &p, float *dst_xy, uint32_t i, float2 a) {
    float2 v = __ldcg(&p.vel[i]);
    float2 x = __ldcg(&p.pos[i]);
    v.x = __fadd_rn(v.x, __fmul_rn(a.x, p.dt));
    v.y = __fadd_rn(v.y, __fmul_rn(a.y, p.dt));
    x.x = __fadd_rn(x.x, __fmul_rn(v.x, p.dt));
    x.y = __fadd_rn(x.y, __fmul_rn(v.y, p.dt));
    p.acc[i] = a;
    p.vel[i] = v;
    p.pos[i] = x;
    if (i < p.mass_len) {
        const uint32_t o = (i >> 2) * 8u + (i & 3u);
        dst_xy[o] = x.x;
        dst_xy[o + 4] = x.y;
    }
}

template <int THREADS, int T, int TJP, int NS, int MINB, int UNR>
__global__ void __launch_bounds__(THREADS, MINB) step_kernel_persistent(const __grid_constant__ PersistParams pp) {
    static_assert(TJP % UNR == 0, "tile must be a multiple of the unroll");
    constexpr int PF = NS - 2;
    constexpr int NWARP = THREADS / 32;
    static_assert(PF >= 1, "need at least 3 stages");
    const StepParams &p = pp.s;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *st_xy = reinterpret_cast<float4 *>(smem_raw);
    float4 *st_gm = st_xy + (size_t)NS * TJP * 2;
    uint64_t *full = reinterpret_cast<uint64_t *>(st_gm + (size_t)NS * TJP);
    uint64_t *empty = full + NS;
    __shared__ int s_last;

    const uint32_t tid = threadIdx.x;
    const uint32_t slab = blockIdx.x % p.nslab;
    const uint32_t tile = p.tile_first + blockIdx.x / p.nslab;
    const uint32_t i0 = tile * (THREADS * T) + tid;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NS; s++) {
            mbar_init(&full[s], 1);
            mbar_init(&empty[s], NWARP);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    asm volatile("griddepcontrol.wait;" ::: "memory"); /* the previous launch on the stream may have been a PDL primary */

    const uint32_t pk0 = slab * p.slab_packs;
    const uint32_t pk1 = min(pk0 + p.slab_packs, p.src_packs_pad);
    const uint32_t npk = pk1 > pk0 ? pk1 - pk0 : 0;
    const uint32_t ntile = (npk + TJP - 1) / TJP;

    float2 rho[T];
#pragma unroll
    for (int k = 0; k < T; k++) {
        const uint32_t i = i0 + k * THREADS;
        const float r = i < p.total_len ? p.radius[i] : 1.f;
        rho[k] = make_float2(r, r);
    }
    __syncthreads();

    uint32_t kbase = 0; /* source tiles this CTA has consumed in earlier steps: ring position */
    for (uint32_t step = 0; step < pp.nsteps; step++) {
        const float4 *src_xy = (step & 1u) ? reinterpret_cast<const float4 *>(p.dst_xy[0]) : p.src_xy;
        float *dst_xy = (step & 1u) ? reinterpret_cast<float *>(const_cast<float4 *>(p.src_xy)) : p.dst_xy[0];

        auto issue = [&](uint32_t k) { /* thread 0: fetch source tile k of this step into the ring */
            const uint32_t kk = kbase + k;
            const uint32_t s = kk % NS;
            const uint32_t f = kk / NS;
            if (f > 0) mbar_wait(&empty[s], (f - 1) & 1);
            const uint32_t first = pk0 + k * TJP;
            const uint32_t cnt = min((uint32_t)TJP, pk1 - first);
            mbar_expect_tx(&full[s], cnt * 48u);
            bulk_g2s(st_xy + (size_t)s * TJP * 2, src_xy + (size_t)first * 2, cnt * 32u, &full[s]);
            bulk_g2s(st_gm + (size_t)s * TJP, p.src_gm + first, cnt * 16u, &full[s]);
        };
        if (tid == 0) {
            fence_proxy_async_global(); /* the positions were written by ordinary stores of other CTAs */
            for (uint32_t k = 0; k < (uint32_t)PF && k < ntile; k++) issue(k);
        }

        float2 ntx[T], nty[T], ax[T], ay[T];
#pragma unroll
        for (int k = 0; k < T; k++) {
            const uint32_t i = i0 + k * THREADS;
            float2 x = make_float2(0.f, 0.f);
            if (i < p.total_len) x = __ldcg(&p.pos[i]);
            ntx[k] = make_float2(-x.x, -x.x);
            nty[k] = make_float2(-x.y, -x.y);
            ax[k] = make_float2(0.f, 0.f);
            ay[k] = make_float2(0.f, 0.f);
        }

        for (uint32_t k = 0; k < ntile; k++) {
            if (tid == 0 && k + PF < ntile) issue(k + PF);
            const uint32_t kk = kbase + k;
            const uint32_t s = kk % NS;
            mbar_wait(&full[s], (kk / NS) & 1);
            const float4 *xy = st_xy + (size_t)s * TJP * 2;
            const float4 *gm = st_gm + (size_t)s * TJP;
            const uint32_t cnt = min((uint32_t)TJP, npk - k * TJP);
#pragma unroll 1
            for (uint32_t q = 0; q < cnt; q += UNR) {
#pragma unroll
                for (int u = 0; u < UNR; u++) {
                    const float4 xs = xy[2 * (q + u)];
                    const float4 ys = xy[2 * (q + u) + 1];
                    const float4 gs = gm[q + u];
                    interact_pack<T, kFormPacked>(xs, ys, gs, ntx, nty, rho, ax, ay);
                }
            }
            __syncwarp();
            if ((tid & 31) == 0) mbar_arrive(&empty[s]);
        }
        kbase += ntile;

        float2 a[T];
#pragma unroll
        for (int k = 0; k < T; k++) a[k] = make_float2(__fadd_rn(ax[k].x, ax[k].y), __fadd_rn(ay[k].x, ay[k].y));

        if (p.nslab == 1) {
#pragma unroll
            for (int k = 0; k < T; k++) {
                const uint32_t i = i0 + k * THREADS;
                if (i < p.total_len) integrate_store_cg(p, dst_xy, i, a[k]);
            }
        } else {
#pragma unroll
            for (int k = 0; k < T; k++) {
                const uint32_t i = i0 + k * THREADS;
                __stcg(&p.part[(size_t)slab * p.part_stride + i], a[k]);
            }
            __threadfence();
            __syncthreads();
            if (tid == 0) {
                const unsigned ticket = atomicAdd(&p.tile_cnt[tile], 1u);
                const int last = (ticket == p.nslab - 1);
                if (last) p.tile_cnt[tile] = 0;
                s_last = last;
            }
            __syncthreads();
            if (s_last) {
                __threadfence();
                float2 sum[T];
                for (uint32_t s0 = 0; s0 < p.nslab; s0 += 8) {
                    float2 v[T][8];
#pragma unroll
                    for (int k = 0; k < T; k++) {
                        const uint32_t i = i0 + k * THREADS;
#pragma unroll
                        for (int u = 0; u < 8; u++)
                            if (s0 + u < p.nslab) v[k][u] = __ldcg(&p.part[(size_t)(s0 + u) * p.part_stride + i]);
                    }
#pragma unroll
                    for (int k = 0; k < T; k++) {
#pragma unroll
                        for (int u = 0; u < 8; u++) {
                            if (s0 + u < p.nslab) {
                                if (s0 + u == 0) {
                                    sum[k] = v[k][0];
                                } else {
                                    sum[k].x = __fadd_rn(sum[k].x, v[k][u].x);
                                    sum[k].y = __fadd_rn(sum[k].y, v[k][u].y);
                                }
                            }
                        }
                    }
                }
#pragma unroll
                for (int k = 0; k < T; k++) {
                    const uint32_t i = i0 + k * THREADS;
                    if (i < p.total_len) integrate_store_cg(p, dst_xy, i, sum[k]);
                }
            }
        }

        if (step + 1 == pp.nsteps) break; /* the kernel boundary separates the last step from what follows */

        /* grid barrier: every CTA's stores of this step (state, next source buffer, zeroed ticket) are visible
         * to every CTA -- to its ordinary loads and to its bulk copies -- before any CTA starts the next step */
        fence_proxy_async_global();
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            red_release_gpu_add(pp.bar, 1u);
            const unsigned target = pp.bar_base + (step + 1u) * gridDim.x;
            while ((int)(ld_acquire_gpu_u32(pp.bar) - target) < 0) {
            }
        }
        __syncthreads();
    }
}

}  // namespace nb
