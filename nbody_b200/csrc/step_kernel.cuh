/*
 * step_kernel.cuh -- the __global__ entry of the hot path (see force_kernel.cuh for the design).
 *
 * grid  = (target tiles of this launch) * nslab, slab index fastest so that the CTAs that
 *         share a target tile are co-resident and its last ticket is drawn early;
 * block = THREADS threads, T targets per thread, target i = tile*THREADS*T + k*THREADS + tid;
 * smem  = NS stages of TJP packs (48 B per pack) + 2*NS mbarriers.
 */
#pragma once

#include "cluster_reduce.cuh"
#include "force_kernel.cuh"

namespace nb {

template <int TJP, int NS>
constexpr size_t step_smem_bytes() {
    return (size_t)NS * TJP * 48 + 2 * NS * sizeof(uint64_t) + 16;
}
/* GROUPED kernels keep the running sum over the slabs of their group in shared memory (one float2 per target:
 * registers are spoken for -- 122 of 128 at 512 threads) */
template <int THREADS, int T, int TJP, int NS>
constexpr size_t step_smem_bytes_grouped() {
    return step_smem_bytes<TJP, NS>() + (size_t)THREADS * T * sizeof(float2);
}

/*
 * GROUPED = false: one CTA = one (target tile, slab).
 * GROUPED = true:  one CTA = one (target tile, group of consecutive slabs).  The summation order keeps its slab
 *   structure -- even/odd chains inside a slab, slab sum = even + odd -- but the CTA adds the slab sums of its
 *   group in slab order itself (p.fold_tiles source tiles per slab) and only the GROUP sum goes through the
 *   partial buffer: p.nslab and p.slab_packs then count groups.  Used for Worlds whose per-slab partials would
 *   not fit a fixed budget (sim_cuda.cu plan_host): the buffer is nslab x N x 8 B, and nslab grows with M.
 */
template <int THREADS, int T, int TJP, int NS, int MINB, int UNR, int FORM = kFormPacked, bool GROUPED = false>
__global__ void __launch_bounds__(THREADS, MINB) step_kernel(const __grid_constant__ StepParams p) {
    constexpr bool TP = FORM == kFormTargetPacked;
    static_assert(!GROUPED || FORM == kFormPacked, "slab groups exist for the packed form only");
    static_assert(!TP || T % 2 == 0, "the target-packed form needs an even number of targets per thread");
    constexpr int TR = TP ? T / 2 : T; /* f32x2 register sets per thread */
    static_assert(TJP % UNR == 0, "tile must be a multiple of the unroll");
    constexpr int PF = NS - 2;          /* tiles in flight ahead of the consumer */
    constexpr int NWARP = THREADS / 32;
    static_assert(PF >= 1, "need at least 3 stages");

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *st_xy = reinterpret_cast<float4 *>(smem_raw);                 /* [NS][TJP][2] */
    float4 *st_gm = st_xy + (size_t)NS * TJP * 2;                         /* [NS][TJP]    */
    uint64_t *full = reinterpret_cast<uint64_t *>(st_gm + (size_t)NS * TJP);
    uint64_t *empty = full + NS;
    float2 *run_s = reinterpret_cast<float2 *>(empty + NS); /* GROUPED: [T][THREADS] running sum over slabs */
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

    /* Programmatic dependent launch (no-op when the launch did not ask for it): this CTA may have been
     * scheduled while the previous step was still draining; do not touch global memory before that grid
     * has completed and its writes are visible. */
    asm volatile("griddepcontrol.wait;" ::: "memory");

    /* this CTA's slab, in packs */
    const uint32_t pk0 = slab * p.slab_packs;
    const uint32_t pk1 = min(pk0 + p.slab_packs, p.src_packs_pad);
    const uint32_t npk = pk1 > pk0 ? pk1 - pk0 : 0;
    const uint32_t ntile = (npk + TJP - 1) / TJP;

    auto issue = [&](uint32_t k) { /* thread 0 only: fetch source tile k into stage k % NS */
        const uint32_t s = k % NS;
        const uint32_t f = k / NS;
        if (f > 0) mbar_wait(&empty[s], (f - 1) & 1);
        const uint32_t first = pk0 + k * TJP;
        const uint32_t cnt = min((uint32_t)TJP, pk1 - first);
        mbar_expect_tx(&full[s], cnt * 48u);
        bulk_g2s(st_xy + (size_t)s * TJP * 2, p.src_xy + (size_t)first * 2, cnt * 32u, &full[s]);
        bulk_g2s(st_gm + (size_t)s * TJP, p.src_gm + first, cnt * 16u, &full[s]);
    };

    if (tid == 0) {
        for (uint32_t k = 0; k < (uint32_t)PF && k < ntile; k++) issue(k);
    }

    /* targets -> registers, negated.  Source-packed: one target duplicated into both halves of an
     * f32x2; target-packed: targets 2k and 2k+1 of this thread side by side. */
    float2 ntx[TR], nty[TR], rho[TR], ax[TR], ay[TR];
#pragma unroll
    for (int k = 0; k < TR; k++) {
        float2 x[2];
        float r[2];
#pragma unroll
        for (int h = 0; h < (TP ? 2 : 1); h++) {
            const uint32_t i = i0 + (TP ? 2 * k + h : k) * THREADS;
            x[h] = make_float2(0.f, 0.f);
            r[h] = 1.f;
            if (i < p.total_len) {
                x[h] = p.pos[i];
                r[h] = p.radius[i];
            }
        }
        if constexpr (TP) {
            ntx[k] = make_float2(-x[0].x, -x[1].x);
            nty[k] = make_float2(-x[0].y, -x[1].y);
            rho[k] = make_float2(r[0], r[1]);
        } else {
            ntx[k] = make_float2(-x[0].x, -x[0].x);
            nty[k] = make_float2(-x[0].y, -x[0].y);
            rho[k] = make_float2(r[0], r[0]);
        }
        ax[k] = make_float2(0.f, 0.f);
        ay[k] = make_float2(0.f, 0.f);
    }
    __syncthreads(); /* barrier inits visible to every thread before the first wait */

    /* One source tile against this thread's targets. */
    auto consume_tile = [&](uint32_t k) {
        if (tid == 0 && k + PF < ntile) issue(k + PF);
        const uint32_t s = k % NS;
        mbar_wait(&full[s], (k / NS) & 1);

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
                if constexpr (TP) interact_pack_tp<TR>(xs, ys, gs, ntx, nty, rho, ax, ay);
                else interact_pack<TR, FORM, GROUPED ? kOrdGrouped : kOrdDefault>(xs, ys, gs, ntx, nty, rho, ax, ay);
            }
        }
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty[s]);
    };

    [[maybe_unused]] bool have_run = false;
    if constexpr (!GROUPED) {
        for (uint32_t k = 0; k < ntile; k++) consume_tile(k);
    } else {
        /* the slabs of this group, one after the other: whole source tiles each (p.fold_tiles).  A finished slab's
         * sum even + odd joins the running sum, in slab order, and the chains restart; the group's last slab is
         * folded below, with the epilogue. */
        for (uint32_t k0 = 0; k0 < ntile; k0 += p.fold_tiles) {
            const uint32_t k1 = min(k0 + p.fold_tiles, ntile);
            for (uint32_t k = k0; k < k1; k++) consume_tile(k);
            if (k1 < ntile) {
#pragma unroll
                for (int t = 0; t < T; t++) {
                    float2 a = make_float2(__fadd_rn(ax[t].x, ax[t].y), __fadd_rn(ay[t].x, ay[t].y));
                    if (have_run) {
                        const float2 r = run_s[t * THREADS + tid];
                        a = make_float2(__fadd_rn(r.x, a.x), __fadd_rn(r.y, a.y));
                    }
                    run_s[t * THREADS + tid] = a;
                    ax[t] = make_float2(0.f, 0.f);
                    ay[t] = make_float2(0.f, 0.f);
                }
                have_run = true;
            }
        }
    }

    /* pair loop done: let the next step's grid be scheduled into the slots that free up from here on */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    /* slab partial of each target.  Source-packed: even-source half + odd-source half;
     * target-packed: the two halves are two targets. */
    float2 a[T];
#pragma unroll
    for (int k = 0; k < TR; k++) {
        if constexpr (TP) {
            a[2 * k] = make_float2(ax[k].x, ay[k].x);
            a[2 * k + 1] = make_float2(ax[k].y, ay[k].y);
        } else {
            a[k] = make_float2(__fadd_rn(ax[k].x, ax[k].y), __fadd_rn(ay[k].x, ay[k].y));
            if constexpr (GROUPED) {
                if (have_run) {
                    const float2 r = run_s[k * THREADS + tid];
                    a[k] = make_float2(__fadd_rn(r.x, a[k].x), __fadd_rn(r.y, a[k].y));
                }
            }
        }
    }

    if (p.nslab == 1 || FORM == kAblateReduce) {
#pragma unroll
        for (int k = 0; k < T; k++) {
            const uint32_t i = i0 + k * THREADS;
            if (i < p.total_len) integrate_store(p, i, a[k]);
        }
        return;
    }

    if constexpr (FORM == kFormCluster) {
        cluster_reduce_integrate<THREADS, T>(p, tile, a);
        return;
    }

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
        if (last) p.tile_cnt[tile] = 0; /* ready for the next step */
        s_last = last;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();

    /* last CTA of this target tile: add the slab partials in slab order and integrate.  The loads of
     * 8 slabs of all T targets are in flight together (one L2 round trip per 8 slabs, not per target). */
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
        if (i < p.total_len) integrate_store(p, i, sum[k]);
    }
}

}  // namespace nb
