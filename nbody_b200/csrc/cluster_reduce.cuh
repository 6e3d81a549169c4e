/*
 * cluster_reduce.cuh -- slab reduction through distributed shared memory (thread-block clusters).
 *
 * The generic path of step_kernel publishes each CTA's slab partial to global memory, draws a ticket, and the
 * last CTA of the target tile adds the nslab partials and integrates: a store + fence + atomic round trip and
 * ceil(nslab / 8) dependent L2 round trips per tile, 3-4 us per step below N = 16K.
 * Here the nslab CTAs of one target tile are ONE cluster (cluster dimension = nslab <= 16, so they are
 * co-scheduled on one GPC): every CTA leaves its partial in its own shared memory, the cluster synchronises,
 * and CTA r of the cluster adds -- in slab order, like the generic path, so the result is bit-identical -- the
 * partials of its 1/nslab share of the tile's targets straight out of the peers' shared memory and integrates
 * them.  No global traffic for the partials, no ticket, and the integration is spread over all CTAs of the tile.
 *
 * STATUS: measured on a B200 in round 1 (profiles/r01_cluster_variant_check.jsonl): bit-identical to the ticket path
 * in every case, but SLOWER -- N = 4 000: 10.4 vs 8.1 us per step (8 slabs), N = 10 000: 28.1 vs 24.9 us, N = 16 384:
 * 61.0 vs 48.6 us (16 slabs).  A cluster of 8-16 CTAs has to be co-scheduled on one GPC and every CTA keeps its slot
 * until the slowest of the cluster has passed the second cluster barrier (each barrier also flushes L1:
 * MEMBAR.ALL.GPU + CCTL.IVALL in the SASS), which costs more than the store + ticket + L2 reads it replaces.
 * Kept selectable with NBODY_VARIANT=c128x2 / c128x4 as the record of that measurement; the planner never picks it.
 */
#pragma once

#include <cooperative_groups.h>

#include "force_kernel.cuh"

namespace nb {

template <int THREADS, int T>
__device__ __forceinline__ void cluster_reduce_integrate(const StepParams &p, uint32_t tile, const float2 (&a)[T]) {
    namespace cg = cooperative_groups;
    constexpr uint32_t TILE = THREADS * T;
    __shared__ float2 red[TILE];
    cg::cluster_group cluster = cg::this_cluster();
    const uint32_t tid = threadIdx.x;
#pragma unroll
    for (int k = 0; k < T; k++) red[k * THREADS + tid] = a[k]; /* slot = target index within the tile */
    cluster.sync();

    const uint32_t C = cluster.num_blocks(); /* == p.nslab; rank == slab (1-D cluster, slab index fastest) */
    const uint32_t r = cluster.block_rank();
    const uint32_t lo = (uint32_t)((uint64_t)TILE * r / C), hi = (uint32_t)((uint64_t)TILE * (r + 1) / C);
    for (uint32_t e = lo + tid; e < hi; e += THREADS) {
        float2 sum = make_float2(0.f, 0.f);
        for (uint32_t s0 = 0; s0 < C; s0 += 8) {
            float2 v[8]; /* 8 remote loads in flight, added strictly in slab order */
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (s0 + u < C) v[u] = cluster.map_shared_rank(red, s0 + u)[e];
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (s0 + u < C) {
                    if (s0 + u == 0) {
                        sum = v[0];
                    } else {
                        sum.x = __fadd_rn(sum.x, v[u].x);
                        sum.y = __fadd_rn(sum.y, v[u].y);
                    }
                }
            }
        }
        const uint32_t i = tile * TILE + e;
        if (i < p.total_len) integrate_store(p, i, sum);
    }
    cluster.sync(); /* nobody leaves while a peer may still read its shared memory */
}

}  // namespace nb
