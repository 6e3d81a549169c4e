/*
 * split_kernel.cuh -- the small-World kernel with a tile's slabs split over a CLUSTER of 2 ... 4 CTAs.
 *
 * small_kernel.cuh gives every tile of 32 targets one CTA.  A World of 2 000 particles is 63 tiles: 63 of 148 SMs
 * work, each issuing all 986 sources against its 32 targets (1.9 us of a 4.0 us step), 85 idle.  Here the CTAs of a
 * cluster share one tile: CTA r of CS stages and computes only slabs [r per, (r + 1) per), per = ceil(nslab / CS),
 * with the same warps-take-slabs scheme, arithmetic and even/odd chains, and leaves each slab's sum in the shared
 * memory of CTA 0 of the cluster (distributed shared memory, st.shared::cluster; CTA 0 stores locally).  One
 * hardware cluster barrier later CTA 0 adds ALL slab sums in slab order and integrates, exactly as
 * small_world_kernel does: same plan, same summation order, same bits
 * (tests/test_gpu_parity.py::test_cluster_split_small_kernel_equals_one_cta_per_tile_bitwise).
 *
 * CS = min(4, SMs / tiles) is part of the launch, not of the plan: it does not touch the order of summation.
 */
#pragma once

#include "force_kernel.cuh"
#include "resident_kernel.cuh" /* cluster_ctarank, cluster_nctarank, cluster_sync_all */
#include "small_kernel.cuh"

namespace nb {

__device__ __forceinline__ void st_cluster_f32x2(uint32_t local, uint32_t rank, float2 v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
    asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(remote), "f"(v.x), "f"(v.y) : "memory");
}

/* shared-memory layout of small_world_kernel (small_smem_bytes), in every CTA of the cluster */
template <int WARPS, int T, int UNR>
__global__ void __launch_bounds__(WARPS * 32, 1) small_split_kernel(const __grid_constant__ StepParams p) {
    constexpr int TILE = 32 * T;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_xy = reinterpret_cast<float4 *>(smem_raw);                       /* [src_packs_pad][2] */
    float4 *s_gm = s_xy + (size_t)p.src_packs_pad * 2;                         /* [src_packs_pad]    */
    float2 *s_part = reinterpret_cast<float2 *>(s_gm + p.src_packs_pad);       /* [nslab][TILE]      */
    uint64_t *bar = reinterpret_cast<uint64_t *>(s_part + (size_t)p.nslab * TILE);

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank(), CS = cluster_nctarank();
    const uint32_t tile = p.tile_first + blockIdx.x / CS;
    const uint32_t t0 = tile * TILE;
    /* this CTA's slabs and the source packs they cover */
    const uint32_t per = (p.nslab + CS - 1) / CS;
    const uint32_t s_first = min(rank * per, p.nslab), s_last = min(s_first + per, p.nslab);
    const uint32_t pk_first = min(s_first * p.slab_packs, p.src_packs_pad);
    const uint32_t pk_last = min(s_last * p.slab_packs, p.src_packs_pad);

    /* Distributed shared memory may only be touched once every CTA of the cluster is running: arrive now, wait
     * before the first remote store. */
    asm volatile("barrier.cluster.arrive.relaxed;" ::: "memory");
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    /* programmatic dependent launch: nothing global before the previous step's writes are visible */
    asm volatile("griddepcontrol.wait;" ::: "memory");

    if (tid == 0) { /* this CTA's sources -> shared memory, at their usual place; bulk copies of at most 32 KB each */
        const uint32_t xy_bytes = (pk_last - pk_first) * 32u, gm_bytes = (pk_last - pk_first) * 16u;
        if (xy_bytes) {
            mbar_expect_tx(bar, xy_bytes + gm_bytes);
            unsigned char *dxy = reinterpret_cast<unsigned char *>(s_xy + (size_t)pk_first * 2);
            const unsigned char *gxy = reinterpret_cast<const unsigned char *>(p.src_xy + (size_t)pk_first * 2);
            for (uint32_t o = 0; o < xy_bytes; o += 32768u) bulk_g2s(dxy + o, gxy + o, min(32768u, xy_bytes - o), bar);
            unsigned char *dgm = reinterpret_cast<unsigned char *>(s_gm + pk_first);
            const unsigned char *ggm = reinterpret_cast<const unsigned char *>(p.src_gm + pk_first);
            for (uint32_t o = 0; o < gm_bytes; o += 32768u) bulk_g2s(dgm + o, ggm + o, min(32768u, gm_bytes - o), bar);
        } else {
            mbar_arrive(bar);
        }
    }

    /* this lane's targets (the same in every warp of every CTA of the cluster); CTA 0 also fetches what it integrates */
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
    const bool finisher = rank == 0 && tid < (uint32_t)TILE && my_i < p.total_len;
    if (finisher) {
        my_v = p.vel[my_i];
        my_x = p.pos[my_i];
    }

    /* (the wait costs ~500 clk even when everybody has arrived: it goes here, under the source copy's flight time) */
    asm volatile("barrier.cluster.wait;" ::: "memory"); /* CTA 0 is there: its shared memory may be written */
    mbar_wait(bar, 0);

    const uint32_t part_local = smem_u32(s_part);
    for (uint32_t s = s_first + warp; s < s_last; s += WARPS) {
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
        for (int k = 0; k < T; k++) {
            const float2 sum = make_float2(__fadd_rn(ax[k].x, ax[k].y), __fadd_rn(ay[k].x, ay[k].y));
            const uint32_t e = (uint32_t)(s * TILE + k * 32 + lane);
            if (rank == 0) s_part[e] = sum;
            else st_cluster_f32x2(part_local + e * (uint32_t)sizeof(float2), 0u, sum); /* CTA 0's copy of s_part */
        }
    }

    /* pair loop done: the next step's grid may be scheduled into the slots that free up from here on */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    cluster_sync_all(); /* every slab sum of the tile is in CTA 0 (also this CTA's own barrier) */

    if (finisher) {
        float2 sum = s_part[tid];
        for (uint32_t s = 1; s < p.nslab; s++) {
            const float2 v = s_part[(size_t)s * TILE + tid];
            sum.x = __fadd_rn(sum.x, v.x);
            sum.y = __fadd_rn(sum.y, v.y);
        }
        integrate_store_vx(p, my_i, sum, my_v, my_x);
    }
}

}  // namespace nb
