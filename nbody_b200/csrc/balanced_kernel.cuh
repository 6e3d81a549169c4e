/*
 * balanced_kernel.cuh -- the step kernel of MID-SIZE Worlds (N ~ 5 K ... 30 K): one CTA per SM, every CTA the same
 * number of (target row, slab) units, whatever N is.  Planned (variant b16x2, sim_cuda.cu balanced_variant_for) for the
 * sizes below N = 24 000 that the small-World kernel leaves out -- 9 473 ... 11 100, 14 209 ... 14 800, 18 945 ... 24 000
 * at M ~ N / 2 -- where it is 3 ... 13 % faster than the (tile, slab) kernel (profiles/r02_balanced_v4_bands.txt);
 * elsewhere the other two kernels win.
 *
 * Between the small Worlds (small_kernel.cuh: one CTA per tile of 32 T targets, at most one CTA per SM) and the
 * throughput regime (step_kernel.cuh: thousands of (tile, slab) CTAs in full waves) the step is quantised by the
 * machine: N = 16 384 is 512 warps of targets for 148 SMs.  The small-World kernel runs it as 128 CTAs (20 SMs idle,
 * 47.7 us), the (tile, slab) kernel as 1 024 CTAs on 888 slots (1.15 waves, 48.3 us); the arithmetic alone is 38 us.
 * N = 10 000: 22.5 us for 14 us of arithmetic.  Here the work is cut into units of one target ROW (32 T targets, T
 * per lane) against one SLAB of sources, and shared out over G = min(SMs, units) CTAs like this:
 *
 *   - every CTA gets w = floor(rows / G) WHOLE rows (all nslab units of each);
 *   - the r = rows - w G rows left over are SHARED: their r nslab units, in row-major order, are cut into G equal
 *     contiguous pieces, one per CTA (less than one row's worth each).
 *
 * So every CTA has the same number of units to within one.  Inside the CTA
 *
 *   - all sources are staged in shared memory once (TMA bulk copies, as in small_kernel.cuh); the targets of the rows
 *     the CTA touches (pos, vel, radius) are staged beside them at kernel entry, off the critical path;
 *   - each of the CW compute warps walks two contiguous runs of units: 1 / CW of the CTA's piece of the shared rows
 *     FIRST, then 1 / CW of its whole rows (cut so that the totals differ by one unit at most), with the same pair
 *     arithmetic and the same even/odd chains as the other kernels (interact_pack); consecutive slabs of a row go
 *     back to back with the targets in registers, so a unit boundary costs a slab-sum store and nothing else, and
 *     slabs can be short;
 *   - one more warp, the FINISHER, computes no pairs.  It sleeps on an mbarrier until the compute warps have left the
 *     shared part of the list (a few us into the step), then publishes this CTA's slab sums of the shared rows in the
 *     global partial buffer and takes an arrival ticket per row, exactly like step_kernel.cuh; where it is the last
 *     of a row's CTAs to arrive it adds ALL nslab partials in slab order and integrates.  That chain of L2 round trips
 *     (stores, fence, ticket, ceil(nslab / 16) rounds of loads: 3 ... 6 us) runs beside the pair loops of the whole
 *     rows instead of after them;
 *   - after one __syncthreads the whole rows are added up in slab order from shared memory and integrated, 32 targets
 *     per warp, all warps at once.
 *
 * The summation order is the plan's alone (slab boundaries, even/odd chains, slabs in order): which CTA or warp
 * computed which unit does not enter, so the results are bit-identical to step_kernel.cuh and small_kernel.cuh with
 * the same slab length, for any launch range and any number of GPUs (tests/test_gpu_parity.py).
 *
 * Three earlier versions are on record (profiles/r02_balanced_v1_sweep.txt, r02_balanced_v2_modes.txt,
 * r02_step_timelines_v2.txt, r02_step_timelines_v3.txt, r02_balanced_v3_sweep.txt).  Units handed out one at a time
 * -- round-robin, or from a shared-memory queue with per-row completion counters -- paid ~2 600 clk per unit at the
 * unit boundaries: all warps of an SM reach them together, so nobody hides the target fetch, the index arithmetic,
 * the fence and the atomics.  Contiguous runs fixed that, but with every CTA's range starting and ending inside a row
 * the ticket chain of those two rows came after the pair loop: 5.7 us of a 48.7 us step at N = 16 384 (T = 2,
 * 46 slabs), 14 us at T = 4 with 73 slabs.
 */
#pragma once

#include "force_kernel.cuh"
#include "small_kernel.cuh"

namespace nb {

struct BalancedGeom {
    uint32_t whole;        /* w: whole rows per CTA                                        */
    uint32_t shared_rows;  /* r: rows whose units are spread over all CTAs                  */
    uint32_t shared_units; /* r x nslab                                                     */
    uint32_t spc_max;      /* most shared units any CTA gets                                */
    uint32_t upc_max;      /* most units any CTA gets: spc_max + w x nslab                  */
    uint32_t rpc_max;      /* most rows any CTA touches: its shared rows + its whole rows   */
};

/* ROWS target rows against NSLAB slabs over CTAS CTAs (CTAS <= ROWS x NSLAB) */
__host__ __device__ inline BalancedGeom balanced_geom(uint32_t rows, uint32_t nslab, uint32_t ctas) {
    BalancedGeom g;
    g.whole = rows / ctas;
    g.shared_rows = rows - g.whole * ctas;
    g.shared_units = g.shared_rows * nslab;
    g.spc_max = (g.shared_units + ctas - 1) / ctas;
    g.upc_max = g.spc_max + g.whole * nslab;
    g.rpc_max = (g.spc_max ? (g.spc_max + nslab - 2) / nslab + 1 : 0) + g.whole;
    return g;
}

/* dynamic shared memory: sources [src_packs_pad] x 48 B | unit sums [upc_max][32 T] float2 |
 * staged targets [rpc_max][32 T] x (pos, vel: float2; radius: float) | 2 mbarriers */
__host__ __device__ inline size_t balanced_smem_bytes(uint32_t src_packs_pad, uint32_t nslab, int T, uint32_t rows,
                                                      uint32_t ctas) {
    const BalancedGeom g = balanced_geom(rows, nslab, ctas);
    return (size_t)src_packs_pad * 48 + (size_t)g.upc_max * 32 * T * sizeof(float2) + (size_t)g.rpc_max * 32 * T * 20 + 16;
}

/* integrate_store_vx with every field of the parameter block already in registers (RowOut).  Inside the out-of-line
 * finishers the block is reached through a generic pointer, and the compiler must assume that each of our stores may
 * have changed it: with the fields read through that pointer where they are used, every store was followed by
 * dependent re-loads of part / part_stride / pos / vel / ... (measured: +4.7 us per step at T = 2, +8.5 us at T = 4). */
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

/*
 * The finisher's work on one SHARED row: this CTA holds the sums of slabs [s_lo, s_hi) of it (MINE points at the sum
 * of slab s_lo; consecutive slabs are 32 T float2 apart).  Out of line, so that its registers -- the partial loads
 * -- do not weigh on the pair loop.  CONTRIBUTORS CTAs hold units of the row.
 */
template <int T>
__device__ __noinline__ void balanced_finish_shared_row(const StepParams &p, const float2 *mine, const float2 *row_vel,
                                                        const float2 *row_pos, uint32_t tile, uint32_t s_lo, uint32_t s_hi,
                                                        uint32_t contributors, uint32_t lane) {
    constexpr int ROW = 32 * T;
    /* the parameter block, once */
    const RowOut out = {p.pos, p.vel, p.acc, p.dst_xy[0], p.dst_xy, p.ndst, p.total_len, p.mass_len, p.dt};
    float2 *const part = p.part;
    unsigned *const tile_cnt = p.tile_cnt;
    const uint32_t nslab = p.nslab, part_stride = p.part_stride;
    const uint32_t i0 = tile * ROW + lane;
    float2 sum[T];
    if (contributors > 1) {
        /* publish this CTA's slab sums, take a ticket */
        for (uint32_t sl = s_lo; sl < s_hi; sl++) {
#pragma unroll
            for (int k = 0; k < T; k++)
                __stcg(&part[(size_t)sl * part_stride + i0 + k * 32], mine[(size_t)(sl - s_lo) * ROW + k * 32 + lane]);
        }
        __threadfence();
        __syncwarp();
        int last = 0;
        if (lane == 0) {
            const unsigned ticket = atomicAdd(&tile_cnt[tile], 1u);
            last = (ticket == contributors - 1);
            if (last) tile_cnt[tile] = 0; /* ready for the next step */
        }
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last) return;
        __threadfence();
        /* last CTA of this row: all slab partials in slab order; RB slabs of all T targets in flight together */
        constexpr int RB = T <= 2 ? 16 : 8;
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
    } else { /* the whole row happens to be in this CTA's piece */
#pragma unroll
        for (int k = 0; k < T; k++) sum[k] = mine[k * 32 + lane];
        for (uint32_t sl = 1; sl < nslab; sl++) {
#pragma unroll
            for (int k = 0; k < T; k++) {
                const float2 v = mine[(size_t)sl * ROW + k * 32 + lane];
                sum[k].x = __fadd_rn(sum[k].x, v.x);
                sum[k].y = __fadd_rn(sum[k].y, v.y);
            }
        }
    }
#pragma unroll
    for (int k = 0; k < T; k++) {
        const uint32_t i = i0 + k * 32;
        if (i < out.total_len) row_integrate_store(out, i, sum[k], row_vel[k * 32 + lane], row_pos[k * 32 + lane]);
    }
}

/* One target of a WHOLE row: slab sums MINE[s * stride] for s < nslab, in slab order, then the integrator. */
__device__ __noinline__ void balanced_finish_whole(const StepParams &p, const float2 *mine, uint32_t stride, float2 v0,
                                                   float2 x0, uint32_t i) {
    const RowOut out = {p.pos, p.vel, p.acc, p.dst_xy[0], p.dst_xy, p.ndst, p.total_len, p.mass_len, p.dt};
    const uint32_t nslab = p.nslab;
    float2 sum = mine[0];
    uint32_t sl = 1;
    for (; sl + 8 <= nslab; sl += 8) {
        float2 v[8];
#pragma unroll
        for (int w = 0; w < 8; w++) v[w] = mine[(size_t)(sl + w) * stride];
#pragma unroll
        for (int w = 0; w < 8; w++) {
            sum.x = __fadd_rn(sum.x, v[w].x);
            sum.y = __fadd_rn(sum.y, v[w].y);
        }
    }
    for (; sl < nslab; sl++) {
        const float2 v = mine[(size_t)sl * stride];
        sum.x = __fadd_rn(sum.x, v.x);
        sum.y = __fadd_rn(sum.y, v.y);
    }
    if (i < out.total_len) row_integrate_store(out, i, sum, v0, x0);
}

/*
 * CW compute warps + 1 finisher warp.  StepParams as the other kernels see it; fold_tiles carries the number of target
 * rows of this launch.  TL: record the per-warp timeline (tools/step_timeline.py).
 */
template <int CW, int T, int UNR, bool TL = false>
__global__ void __launch_bounds__((CW + 1) * 32, 1) balanced_kernel(const __grid_constant__ StepParams p) {
    constexpr int ROW = 32 * T;
    constexpr int THREADS = (CW + 1) * 32;
    const uint32_t G = gridDim.x, c = blockIdx.x;
    const BalancedGeom geo = balanced_geom(p.fold_tiles, p.nslab, G);

    extern __shared__ __align__(128) unsigned char smem_raw[];
    float4 *s_xy = reinterpret_cast<float4 *>(smem_raw);                  /* [src_packs_pad][2] */
    float4 *s_gm = s_xy + (size_t)p.src_packs_pad * 2;                    /* [src_packs_pad]    */
    float2 *s_part = reinterpret_cast<float2 *>(s_gm + p.src_packs_pad);  /* [upc_max][ROW]     */
    float2 *s_tpos = s_part + (size_t)geo.upc_max * ROW;                  /* [rpc_max][ROW]     */
    float2 *s_tvel = s_tpos + (size_t)geo.rpc_max * ROW;
    float *s_trad = reinterpret_cast<float *>(s_tvel + (size_t)geo.rpc_max * ROW);
    uint64_t *bar = reinterpret_cast<uint64_t *>(s_trad + (size_t)geo.rpc_max * ROW); /* [0] sources, [1] shared part done */

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    /* this CTA's piece [us0, us1) of the shared units and the shared rows (numbered from 0) it lies in */
    const uint32_t us0 = (uint32_t)((uint64_t)c * geo.shared_units / G);
    const uint32_t us1 = (uint32_t)((uint64_t)(c + 1) * geo.shared_units / G);
    const uint32_t Ls = us1 - us0;
    const uint32_t srow_first = Ls ? us0 / p.nslab : 0u;
    const uint32_t nsr = Ls ? (us1 - 1) / p.nslab - srow_first + 1 : 0u;
    const uint32_t L = Ls + geo.whole * p.nslab;   /* the unit list: shared piece, then whole rows */
    const uint32_t nrow = nsr + geo.whole;         /* staged rows: shared rows, then whole rows    */
    /* launch-local row number of staged row j */
    auto row_of = [&](uint32_t j) { return j < nsr ? geo.whole * G + srow_first + j : c * geo.whole + (j - nsr); };
    /* Every compute warp walks TWO contiguous runs: its 1 / CW of the shared piece first, then its 1 / CW of the
     * whole rows -- cut so that the two together are 1 / CW of the list to within one unit.  The shared piece is
     * therefore behind all warps after Ls / L of the pair time, and the finisher has the rest of it for the ticket
     * chain.  run_cut(w) = units of the whole rows that belong to warps below w. */
    const uint32_t Lw = L - Ls;
    auto run_cut = [&](uint32_t w) {
        uint32_t cut = 0;
        for (uint32_t v = 1; v <= w; v++) { /* monotone by construction, ends at Lw */
            const uint32_t want = (uint32_t)((uint64_t)v * L / CW) - (uint32_t)((uint64_t)v * Ls / CW);
            cut = min(max(cut, want), Lw);
        }
        return w == (uint32_t)CW ? Lw : cut;
    };
    uint32_t leavers = 0; /* compute warps with at least one shared unit: they tell the finisher when it is done */
    for (uint32_t w = 0; w < (uint32_t)CW; w++) leavers += (uint32_t)((uint64_t)(w + 1) * Ls / CW) > (uint32_t)((uint64_t)w * Ls / CW);

    if (tid == 0) {
        mbar_init(&bar[0], 1);
        mbar_init(&bar[1], leavers ? leavers : 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    /* programmatic dependent launch: nothing global before the previous step's writes are visible */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const uint32_t tl_slot = c * (CW + 1) + warp;
    timeline_mark(TL, tl_slot, 0);

    if (tid == 0) { /* all sources -> shared memory; bulk copies of at most 32 KB each */
        const uint32_t xy_bytes = p.src_packs_pad * 32u, gm_bytes = p.src_packs_pad * 16u;
        mbar_expect_tx(&bar[0], xy_bytes + gm_bytes);
        for (uint32_t o = 0; o < xy_bytes; o += 32768u)
            bulk_g2s(reinterpret_cast<unsigned char *>(s_xy) + o, reinterpret_cast<const unsigned char *>(p.src_xy) + o,
                     min(32768u, xy_bytes - o), &bar[0]);
        for (uint32_t o = 0; o < gm_bytes; o += 32768u)
            bulk_g2s(reinterpret_cast<unsigned char *>(s_gm) + o, reinterpret_cast<const unsigned char *>(p.src_gm) + o,
                     min(32768u, gm_bytes - o), &bar[0]);
    }

    /* the targets of every row this CTA touches: one round trip, concurrent with the source copy */
    for (uint32_t k = tid; k < nrow * ROW; k += THREADS) {
        const uint32_t j = k / ROW;
        const uint32_t i = (p.tile_first + row_of(j)) * ROW + (k - j * ROW);
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

    if (warp < (uint32_t)CW) {
        mbar_wait(&bar[0], 0);
        timeline_mark(TL, tl_slot, 1);
        float2 ntx[T], nty[T], rho[T], ax[T], ay[T];
#pragma unroll 1
        for (int phase = 0; phase < 2; phase++) {
            /* positions [a, b) of the unit list, staged row j and slab s of unit a */
            uint32_t a, b, j, s;
            if (phase == 0) {
                a = (uint32_t)((uint64_t)warp * Ls / CW);
                b = (uint32_t)((uint64_t)(warp + 1) * Ls / CW);
                j = (us0 + a) / p.nslab - srow_first;
                s = (us0 + a) - (srow_first + j) * p.nslab;
            } else {
                const uint32_t wa = run_cut(warp), wb = run_cut(warp + 1);
                a = Ls + wa;
                b = Ls + wb;
                j = nsr + wa / p.nslab;
                s = wa - (j - nsr) * p.nslab;
            }
            bool fresh = true;
            for (uint32_t ul = a; ul < b; ul++) {
                if (fresh) {
                    const uint32_t t = j * ROW + lane;
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
                fresh = ++s == p.nslab; /* next: the next slab of this row, or the next row */
                if (fresh) {
                    s = 0;
                    j++;
                }
            }
            if (phase == 0 && b > a) { /* this warp's shared units are done: tell the finisher */
                __syncwarp();
                if (lane == 0) mbar_arrive(&bar[1]);
            }
        }
        timeline_mark(TL, tl_slot, 2);
    } else if (Ls) {
        /* The finisher: once the shared piece is computed, publish it and finish the rows this CTA arrives last at. */
        mbar_wait(&bar[1], 0);
        timeline_mark(TL, tl_slot, 1);
        for (uint32_t j = 0; j < nsr; j++) {
            const uint32_t ub = (srow_first + j) * p.nslab;          /* the row's unit of slab 0, in shared units */
            const uint32_t s_lo = ub >= us0 ? 0u : us0 - ub;         /* this CTA holds slabs [s_lo, s_hi)         */
            const uint32_t s_hi = min(p.nslab, us1 - ub);
            /* unit u of the shared units belongs to CTA ceil((u + 1) G / U) - 1 */
            const uint32_t c_first = (uint32_t)((((uint64_t)ub + 1) * G + geo.shared_units - 1) / geo.shared_units) - 1;
            const uint32_t c_last = (uint32_t)((((uint64_t)ub + p.nslab) * G + geo.shared_units - 1) / geo.shared_units) - 1;
            balanced_finish_shared_row<T>(p, s_part + (size_t)(ub + s_lo - us0) * ROW, s_tvel + j * ROW, s_tpos + j * ROW,
                                          p.tile_first + row_of(j), s_lo, s_hi, c_last - c_first + 1, lane);
        }
        timeline_mark(TL, tl_slot, 2);
    }

    /* pair loops done: the next step's grid may be scheduled into the slots that free up from here on */
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();

    /* the whole rows: 32 targets per warp, all warps at once */
    for (uint32_t item = warp; item < geo.whole * T; item += CW + 1) {
        const uint32_t wr = item / T, k = item - wr * T;
        const uint32_t j = nsr + wr;
        const uint32_t t = j * ROW + k * 32 + lane;
        balanced_finish_whole(p, s_part + (size_t)(Ls + wr * p.nslab) * ROW + k * 32 + lane, ROW, s_tvel[t], s_tpos[t],
                              (p.tile_first + row_of(j)) * ROW + k * 32 + lane);
    }
    timeline_mark(TL, tl_slot, 3);
}

}  // namespace nb
