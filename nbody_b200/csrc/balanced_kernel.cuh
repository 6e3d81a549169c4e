/*
 * balanced_kernel.cuh -- the step kernel of MID-SIZE Worlds (N ~ 5 K ... 30 K): one CTA per SM, every CTA the same
 * number of (target row, slab) units, whatever N is.
 *
 * Between the small Worlds (small_kernel.cuh: one CTA per tile of 32 T targets, at most one CTA per SM) and the
 * throughput regime (step_kernel.cuh: thousands of (tile, slab) CTAs in full waves) the step is quantised by the
 * machine: N = 16 384 is 512 warps of targets for 148 SMs.  The small-World kernel runs it as 128 CTAs (20 SMs idle,
 * 47.7 us), the (tile, slab) kernel as 1 024 CTAs on 888 slots (1.15 waves, 48.3 us); the arithmetic alone is 38 us.
 * N = 10 000: 22.5 us for 14 us of arithmetic.  Here the work is cut into units of one target ROW (32 T targets, T
 * per lane) against one SLAB of sources, rows x nslab units in row-major order, and CTA c of G = min(SMs, units)
 * takes the contiguous unit range [c U / G, (c + 1) U / G): no SM gets more than one unit above the average.
 *
 *   - all sources are staged in shared memory once per CTA (TMA bulk copies, as in small_kernel.cuh); the targets of
 *     the rows a CTA touches (pos, vel, radius) are staged beside them at kernel entry, off the critical path;
 *   - the CTA's range is cut into WARPS contiguous runs, one per warp.  A warp walks its run with the same pair
 *     arithmetic and the same even/odd chains as the other kernels (interact_pack); consecutive slabs of a row go
 *     back to back with the targets in registers, so a unit boundary costs a slab-sum store and nothing else, and
 *     slabs can be short (the plan picks ~10 units per warp: the run lengths differ by at most one unit);
 *   - after one __syncthreads one warp per row finishes it: a row whose nslab units all belong to this CTA is added
 *     up in slab order from shared memory and integrated; a row that straddles a CTA boundary (at most two per CTA)
 *     goes through the global partial buffer and an arrival ticket exactly like step_kernel.cuh -- this CTA's slab
 *     sums are stored, and the last of the row's CTAs to arrive adds ALL nslab partials in slab order and integrates.
 *
 * Two earlier versions are on record (profiles/r02_balanced_v1_sweep.txt, r02_balanced_v2_modes.txt,
 * r02_step_timelines_v2.txt): units handed out one at a time -- round-robin, or from a shared-memory queue with
 * per-row completion counters and no barrier -- paid ~2 600 clk per unit at the unit boundaries (all warps of an SM
 * reach them together, so nobody hides the target fetch, the index arithmetic, the fence and the atomics) and lost
 * to the shipped plans at most sizes.
 *
 * The summation order is the plan's alone (slab boundaries, even/odd chains, slabs in order): which CTA computed
 * which unit does not enter, so the results are bit-identical to step_kernel.cuh and small_kernel.cuh with the same
 * slab length, for any launch range and any number of GPUs (tests/test_gpu_parity.py).
 */
#pragma once

#include "force_kernel.cuh"
#include "small_kernel.cuh"

namespace nb {

struct BalancedGeom {
    uint32_t units;   /* rows x nslab                                        */
    uint32_t upc_max; /* most units any CTA gets                             */
    uint32_t rpc_max; /* most rows any CTA touches (whole or partly)         */
};

/* ROWS target rows against NSLAB slabs over CTAS CTAs (CTAS <= units) */
__host__ __device__ inline BalancedGeom balanced_geom(uint32_t rows, uint32_t nslab, uint32_t ctas) {
    BalancedGeom g;
    g.units = rows * nslab;
    g.upc_max = (g.units + ctas - 1) / ctas;
    g.rpc_max = (g.upc_max + nslab - 2) / nslab + 1;
    return g;
}

/* dynamic shared memory: sources [src_packs_pad] x 48 B | unit sums [upc_max][32 T] float2 |
 * staged targets [rpc_max][32 T] x (pos, vel: float2; radius: float) | mbarrier */
__host__ __device__ inline size_t balanced_smem_bytes(uint32_t src_packs_pad, uint32_t nslab, int T, uint32_t rows,
                                                      uint32_t ctas) {
    const BalancedGeom g = balanced_geom(rows, nslab, ctas);
    return (size_t)src_packs_pad * 48 + (size_t)g.upc_max * 32 * T * sizeof(float2) + (size_t)g.rpc_max * 32 * T * 20 + 16;
}

/* integrate_store_vx with every field of the parameter block already in registers (RowOut).  Inside the out-of-line
 * balanced_finish_row the block is reached through a generic pointer, and the compiler must assume that each of our
 * stores may have changed it: with the fields read through that pointer where they are used, every store was followed
 * by dependent re-loads of part / part_stride / pos / vel / ... (measured: +4.7 us per step at T = 2, +8.5 us at T = 4). */
struct RowOut {
    float2 *pos, *vel, *acc;
    float *dst0;
    float *const *dst_xy;
    int ndst;
    uint32_t total_len, mass_len;
    float dt;
};
__device__ __forceinline__ void row_integrate_store(const RowOut &o, uint32_t i, float2 a, float2 v, float2 x) {
    v.x = __fadd_rn(v.x, __fmul_rn(a.x, o.dt));
    v.y = __fadd_rn(v.y, __fmul_rn(a.y, o.dt));
    x.x = __fadd_rn(x.x, __fmul_rn(v.x, o.dt));
    x.y = __fadd_rn(x.y, __fmul_rn(v.y, o.dt));
    o.acc[i] = a;
    o.vel[i] = v;
    o.pos[i] = x;
    if (i < o.mass_len) {
        const uint32_t w = (i >> 2) * 8u + (i & 3u);
        o.dst0[w] = x.x;
        o.dst0[w + 4] = x.y;
#pragma unroll 1
        for (int d = 1; d < o.ndst; d++) { /* peer GPUs */
            float *dst = o.dst_xy[d];
            dst[w] = x.x;
            dst[w + 4] = x.y;
        }
    }
}

/* Finish one target row (the caller's warp just completed the last of this CTA's units of it): kept out of line so
 * that its registers -- the partial loads of the straddling case -- do not weigh on the pair loop. */
template <int T>
__device__ __noinline__ void balanced_finish_row(const StepParams &p, const float2 *s_part, const float2 *row_vel,
                                                 const float2 *row_pos, uint32_t row, uint32_t u0, uint32_t u1, uint32_t G,
                                                 uint32_t units, uint32_t lane) {
    constexpr int ROW = 32 * T;
    /* the parameter block, once */
    const RowOut out = {p.pos, p.vel, p.acc, p.dst_xy[0], p.dst_xy, p.ndst, p.total_len, p.mass_len, p.dt};
    float2 *const part = p.part;
    unsigned *const tile_cnt = p.tile_cnt;
    const uint32_t nslab = p.nslab, part_stride = p.part_stride;
    const uint32_t ub = row * nslab;                         /* the row's unit of slab 0           */
    const uint32_t s_lo = ub >= u0 ? 0u : u0 - ub;           /* this CTA holds slabs [s_lo, s_hi)  */
    const uint32_t s_hi = min(nslab, u1 - ub);
    const uint32_t tile = p.tile_first + row;
    const uint32_t i0 = tile * ROW + lane;
    /* this CTA's sum of slab s (s_lo <= s < s_hi) of the row, target k * 32 + lane */
    auto mine = [&](uint32_t sl, int k) { return s_part[(size_t)(ub + sl - u0) * ROW + k * 32 + lane]; };
    if (s_lo == 0 && s_hi == nslab) { /* the whole row is here: slab order, from shared memory */
#pragma unroll
        for (int k = 0; k < T; k++) {
            float2 sum = mine(0, k);
            for (uint32_t sl = 1; sl < nslab; sl++) {
                const float2 v = mine(sl, k);
                sum.x = __fadd_rn(sum.x, v.x);
                sum.y = __fadd_rn(sum.y, v.y);
            }
            const uint32_t i = i0 + k * 32;
            if (i < out.total_len) row_integrate_store(out, i, sum, row_vel[k * 32 + lane], row_pos[k * 32 + lane]);
        }
        return;
    }
    /* the row straddles a CTA boundary: publish this CTA's slab sums, take a ticket */
    for (uint32_t sl = s_lo; sl < s_hi; sl++) {
#pragma unroll
        for (int k = 0; k < T; k++) __stcg(&part[(size_t)sl * part_stride + i0 + k * 32], mine(sl, k));
    }
    __threadfence();
    __syncwarp();
    int last = 0;
    if (lane == 0) {
        /* unit u belongs to CTA ceil((u + 1) G / U) - 1 */
        const uint32_t c_first = (uint32_t)((((uint64_t)ub + 1) * G + units - 1) / units) - 1;
        const uint32_t c_last = (uint32_t)((((uint64_t)ub + nslab) * G + units - 1) / units) - 1;
        const unsigned ticket = atomicAdd(&tile_cnt[tile], 1u);
        last = (ticket == c_last - c_first);
        if (last) tile_cnt[tile] = 0; /* ready for the next step */
    }
    last = __shfl_sync(0xffffffffu, last, 0);
    if (!last) return;
    __threadfence();
    /* last CTA of this row: all slab partials in slab order; RB slabs of all T targets in flight together */
    constexpr int RB = T <= 2 ? 16 : 8;
    float2 sum[T];
    for (uint32_t s0 = 0; s0 < nslab; s0 += RB) {
        float2 v[T][RB];
#pragma unroll
        for (int k = 0; k < T; k++) {
#pragma unroll
            for (int w = 0; w < RB; w++)
                if (s0 + w < nslab) v[k][w] = __ldcg(&part[(size_t)(s0 + w) * part_stride + i0 + k * 32]);
        }
#pragma unroll
        for (int k = 0; k < T; k++) {
#pragma unroll
            for (int w = 0; w < RB; w++) {
                if (s0 + w < nslab) {
                    if (s0 + w == 0) {
                        sum[k] = v[k][0];
                    } else {
                        sum[k].x = __fadd_rn(sum[k].x, v[k][w].x);
                        sum[k].y = __fadd_rn(sum[k].y, v[k][w].y);
                    }
                }
            }
        }
    }
#pragma unroll
    for (int k = 0; k < T; k++) {
        const uint32_t i = i0 + k * 32;
        if (i < out.total_len) row_integrate_store(out, i, sum[k], row_vel[k * 32 + lane], row_pos[k * 32 + lane]);
    }
}

/* StepParams as the other kernels see it; fold_tiles carries the number of target rows of this launch. */
/* MODE bit 3: record the per-warp timeline (tools/step_timeline.py). */
template <int WARPS, int T, int UNR, int MODE = 0>
__global__ void __launch_bounds__(WARPS * 32, 1) balanced_kernel(const __grid_constant__ StepParams p) {
    constexpr int ROW = 32 * T;
    const uint32_t G = gridDim.x;
    const BalancedGeom geo = balanced_geom(p.fold_tiles, p.nslab, G);

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_xy = reinterpret_cast<float4 *>(smem_raw);                  /* [src_packs_pad][2] */
    float4 *s_gm = s_xy + (size_t)p.src_packs_pad * 2;                    /* [src_packs_pad]    */
    float2 *s_part = reinterpret_cast<float2 *>(s_gm + p.src_packs_pad);  /* [upc_max][ROW]     */
    float2 *s_tpos = s_part + (size_t)geo.upc_max * ROW;                  /* [rpc_max][ROW]     */
    float2 *s_tvel = s_tpos + (size_t)geo.rpc_max * ROW;
    float *s_trad = reinterpret_cast<float *>(s_tvel + (size_t)geo.rpc_max * ROW);
    uint64_t *bar = reinterpret_cast<uint64_t *>(s_trad + (size_t)geo.rpc_max * ROW);

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    /* this CTA's units, and the rows (launch-local numbers) they belong to */
    const uint32_t u0 = (uint32_t)((uint64_t)blockIdx.x * geo.units / G);
    const uint32_t u1 = (uint32_t)((uint64_t)(blockIdx.x + 1) * geo.units / G);
    const uint32_t L = u1 - u0;
    const uint32_t row_first = u0 / p.nslab;
    const uint32_t nrow = (u1 - 1) / p.nslab - row_first + 1;

    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    /* programmatic dependent launch: nothing global before the previous step's writes are visible */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    constexpr bool TL = (MODE & 8) != 0;
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

    /* the targets of every row this CTA touches: one round trip, concurrent with the source copy */
    const uint32_t i_first = (p.tile_first + row_first) * ROW;
    for (uint32_t k = tid; k < nrow * ROW; k += WARPS * 32) {
        const uint32_t i = i_first + k;
        float2 x = make_float2(0.f, 0.f), v = make_float2(0.f, 0.f);
        float r = 1.f;
        if (i < p.total_len) {
            x = p.pos[i];
            v = p.vel[i];
            r = p.radius[i];
        }
        s_tpos[k] = x;
        s_tvel[k] = v;
        s_trad[k] = r;
    }
    __syncthreads();
    mbar_wait(bar, 0);
    timeline_mark(TL, tl_slot, 1);

    /* Warp w walks the contiguous run [w L / WARPS, (w + 1) L / WARPS) of the CTA's units: consecutive slabs of a row
     * back to back with the targets in registers -- between two units only the slab sum is stored and the chains
     * restart -- and a fresh set of targets where the run crosses into the next row. */
    {
        const uint32_t a = (uint32_t)((uint64_t)warp * L / WARPS), b = (uint32_t)((uint64_t)(warp + 1) * L / WARPS);
        uint32_t row = (u0 + a) / p.nslab, s = (u0 + a) - row * p.nslab;
        float2 ntx[T], nty[T], rho[T], ax[T], ay[T];
        bool fresh = true;
        for (uint32_t ul = a; ul < b; ul++) {
            if (fresh) {
                const uint32_t t = (row - row_first) * ROW + lane;
#pragma unroll
                for (int k = 0; k < T; k++) {
                    const float2 x = s_tpos[t + k * 32];
                    const float r = s_trad[t + k * 32];
                    ntx[k] = make_float2(-x.x, -x.x);
                    nty[k] = make_float2(-x.y, -x.y);
                    rho[k] = make_float2(r, r);
                }
            }
#pragma unroll
            for (int k = 0; k < T; k++) {
                ax[k] = make_float2(0.f, 0.f);
                ay[k] = make_float2(0.f, 0.f);
            }
            const uint32_t pk0 = s * p.slab_packs;
            const uint32_t pk1 = min(pk0 + p.slab_packs, p.src_packs_pad);
#pragma unroll 1
            for (uint32_t pk = pk0; pk < pk1; pk += UNR) {
#pragma unroll
                for (int w = 0; w < UNR; w++) {
                    const float4 xs = s_xy[2 * (pk + w)];
                    const float4 ys = s_xy[2 * (pk + w) + 1];
                    const float4 gs = s_gm[pk + w];
                    interact_pack<T, kFormPacked>(xs, ys, gs, ntx, nty, rho, ax, ay);
                }
            }
#pragma unroll
            for (int k = 0; k < T; k++)
                s_part[(size_t)ul * ROW + k * 32 + lane] =
                    make_float2(__fadd_rn(ax[k].x, ax[k].y), __fadd_rn(ay[k].x, ay[k].y));
            fresh = ++s == p.nslab;
            if (fresh) {
                s = 0;
                row++;
            }
        }
    }
    timeline_mark(TL, tl_slot, 2);

    /* pair loop done: the next step's grid may be scheduled into the slots that free up from here on */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();

    /* finish the rows, one warp per row; the rows shared with a neighbour CTA (the first and the last of the range)
     * go first: theirs is the longer chain (stores, ticket, partial loads) */
    for (uint32_t k = warp; k < nrow; k += WARPS) {
        const uint32_t rc = k == 0 ? 0u : k == 1 ? nrow - 1 : k - 1;
        balanced_finish_row<T>(p, s_part, s_tvel + rc * ROW, s_tpos + rc * ROW, row_first + rc, u0, u1, G, geo.units, lane);
    }
    timeline_mark(TL, tl_slot, 3);
}

}  // namespace nb
