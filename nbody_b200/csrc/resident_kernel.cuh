/*
 * resident_kernel.cuh -- n chained steps of a TINY World (N <= 640) in ONE launch of ONE thread-block cluster,
 * with the whole state of the World resident on chip between steps.
 *
 * UpdateWorld_GPU(w, dt, n) at the reference's own benchmark sizes N = 250 ... 1 200 (ref src/bench.c:38, 100 steps
 * per call) is not arithmetic: the small-World kernel replayed as CUDA-graph chunks takes 2.5 ... 3.2 us per step, of
 * which < 0.2 us is the pair loop -- the rest is what a kernel boundary costs (grid launch, dependency wait, the
 * first L2 round trips for sources and targets, stores, flush).  The persistent kernel of persist_kernel.cuh tried to
 * avoid the boundary with a software grid barrier over 148 SMs and was slower.  A World this small does not need 148
 * SMs: here ONE cluster of C <= 16 CTAs runs all n steps,
 *
 *   - every CTA keeps ALL sources in its shared memory (two position buffers, ping-pong, and G*m) and owns a
 *     contiguous range of target rows (32 targets each); thread t of the CTA owns target t of that range for the
 *     whole launch, with its position and velocity in registers;
 *   - per step: the warps take (row, slab) units round-robin -- the same pair arithmetic, slab boundaries and even/odd
 *     chains as small_kernel.cuh (interact_pack, T = 1) -- and leave the slab sums in shared memory; after a
 *     __syncthreads the owner of each target adds its slab sums IN SLAB ORDER, integrates, and stores the new
 *     position of a massive target into the next source buffer of EVERY CTA of the cluster through distributed shared
 *     memory (mapa + st.shared::cluster); one hardware cluster barrier (barrier.cluster arrive.release /
 *     wait.acquire, ~0.2 us) ends the step;
 *   - global memory is read once at the start and written once at the end (pos, vel, acc of the last step, and the
 *     source buffer the next launch will read).
 *
 * Same plan, same summation order: bit-identical to n launches of small_world_kernel
 * (tests/test_gpu_parity.py::test_resident_cluster_kernel_equals_launch_per_step_bitwise).
 *
 * MEASURED ON B200 (round 2, profiles/r02_resident_kernel.txt; nbody-bench protocol, us per step, 100 steps per call):
 *       N                 250    500    800   1000   1200   1536   2000
 *   resident cluster      1.3    1.8    3.0    3.5    4.8    5.9    8.7
 *   CUDA-graph chunks     2.4    2.5    2.9    3.2    3.2    3.4    4.0
 *   launch by launch      3.3    3.3    3.4    3.3    3.3    3.4    3.9
 * (first version; with four packs per loop iteration and 8 / 16 / 32 warps per CTA: 1.1 / 1.7 us at N = 250 / 500.)
 * Where a step's clocks go (tools/resident_breakdown.py, profiles/r02_resident_breakdown.txt; thread 0 of a CTA,
 * 1 000 steps per call): N = 250: 1 835 clk = pair loop 683 + slab sums, integrator, remote stores 493 + cluster
 * barrier 645; N = 500: 2 918 = 1 139 + 833 + 925; N = 640: 4 657 = 2 749 + 913 + 982.  The pair loop of one 16-source
 * unit is a latency chain at N = 250 and already the arithmetic of 16 SMs at N = 500 (7.8 K pairs per CTA: 650 clk at
 * the large kernel's rate), which overtakes the 148-SM paths near N = 750: the planner uses the cluster up to N = 640
 * (sim_cuda.cu kResidentMaxTargets).  A second build that sent the positions as one bulk copy per peer CTA
 * (cp.async.bulk shared::cta -> shared::cluster, complete_tx on alternating mbarriers, no cluster barrier) was
 * bit-identical too and 0.2 ... 1.1 us per step slower at every size (1.5 / 2.2 / 3.7 / 5.6 us at N = 250 / 500 / 800 /
 * 1 200): at these sizes the exchange is latency, not bandwidth, and a remote store is the shorter path.
 */
#pragma once

#include "force_kernel.cuh"
#include "small_kernel.cuh" /* nb_timeline */

namespace nb {

struct ResidentParams {
    StepParams s;             /* s.src_xy: positions read by the first step; s.dst_xy[0]: the buffer the positions of  */
                              /* the LAST step go to (the other one if nsteps is odd, the same one if it is even)      */
    const float4 *src_xy_alt; /* the other global position buffer: only its padding lanes are used                      */
    uint32_t nsteps;          /* >= 1                                                                                   */
    uint32_t rows_per_cta;    /* target rows (32 targets) per CTA; rows_per_cta * 32 <= threads per CTA                */
};

constexpr int kResidentWarps = 32; /* at most; a World with few (row, slab) units per CTA runs with 8 or 16 warps: every */
                                   /* warp of the cluster takes part in the barrier that ends a step                */
constexpr int kResidentMaxCtas = 16; /* largest (non-portable) cluster on sm_100 */

/* dynamic shared memory: xy [2][P][2] float4 | gm [P] float4 | slab sums [rows_per_cta * nslab][32] float2 |
 * target positions [rows_per_cta * 32] float2 | radii [rows_per_cta * 32] float */
__host__ __device__ inline size_t resident_smem_bytes(uint32_t src_packs_pad, uint32_t nslab, uint32_t rows_per_cta) {
    return (size_t)src_packs_pad * 80 + (size_t)rows_per_cta * nslab * 32 * sizeof(float2) + (size_t)rows_per_cta * 32 * 12;
}

__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t cluster_nctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
/* store V at the shared-memory address LOCAL (of this CTA's window) in CTA RANK of the cluster */
__device__ __forceinline__ void st_cluster_f32(uint32_t local, uint32_t rank, float v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(rank));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}

/* TL (diagnostics, NBODY_RESIDENT_TL=1, tools/resident_breakdown.py): thread 0 of every CTA adds up the clocks of the
 * four parts of a step -- its own pair loop, the __syncthreads, slab sums + integrator + remote stores, the cluster
 * barrier -- and leaves the totals in nb_timeline[rank][0..3]. */
template <int UNR, int WARPS, bool TL = false>
__global__ void __launch_bounds__(WARPS * 32, 1) resident_kernel(const __grid_constant__ ResidentParams rp) {
    const StepParams &p = rp.s;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const uint32_t P = p.src_packs_pad;
    float4 *s_xy = reinterpret_cast<float4 *>(smem_raw);                      /* [2][P][2] */
    float4 *s_gm = s_xy + (size_t)P * 4;                                      /* [P]       */
    float2 *s_part = reinterpret_cast<float2 *>(s_gm + P);                    /* [rows_per_cta * nslab][32] */
    float2 *s_tpos = s_part + (size_t)rp.rows_per_cta * p.nslab * 32;         /* [rows_per_cta * 32] */
    float *s_trad = reinterpret_cast<float *>(s_tpos + (size_t)rp.rows_per_cta * 32);

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t rank = cluster_ctarank(), C = cluster_nctarank();
    const uint32_t rows = (p.total_len + 31u) / 32u;
    const uint32_t row0 = min(rank * rp.rows_per_cta, rows);
    const uint32_t nrow = min(rows - row0, rp.rows_per_cta);                  /* may be 0: the CTA still keeps the barriers */

    asm volatile("griddepcontrol.wait;" ::: "memory");

    /* everything this CTA will ever read from global memory */
    for (uint32_t k = tid; k < P * 2; k += WARPS * 32) {
        s_xy[k] = p.src_xy[k];
        s_xy[(size_t)P * 2 + k] = rp.src_xy_alt[k]; /* padding lanes; the real lanes are rewritten by step 1 before use */
    }
    for (uint32_t k = tid; k < P; k += WARPS * 32) s_gm[k] = p.src_gm[k];
    const uint32_t my_i = row0 * 32 + tid;
    const bool mine = tid < nrow * 32 && my_i < p.total_len;
    float2 my_x = make_float2(0.f, 0.f), my_v = make_float2(0.f, 0.f), my_a = make_float2(0.f, 0.f);
    if (tid < nrow * 32) {
        float r = 1.f;
        if (mine) {
            my_x = p.pos[my_i];
            my_v = p.vel[my_i];
            r = p.radius[my_i];
        }
        s_tpos[tid] = my_x;
        s_trad[tid] = r;
    }
    /* nobody stores into a neighbour's buffers before that neighbour has filled them */
    cluster_sync_all();

    const uint32_t units = nrow * p.nslab;
    const uint32_t row_w = warp / p.nslab, s_w = warp - row_w * p.nslab;
    const uint32_t xy_local = smem_u32(s_xy);
    uint32_t cur = 0;
    [[maybe_unused]] long long tl[4] = {0, 0, 0, 0}, tl_t = 0;
    auto tl_lap = [&](int k) {
        if (TL && tid == 0) {
            const long long now = clock64();
            if (k >= 0) tl[k] += now - tl_t;
            tl_t = now;
        }
    };
    for (uint32_t step = 0; step < rp.nsteps; step++, cur ^= 1u) {
        const float4 *xy = s_xy + (size_t)cur * P * 2;
        tl_lap(-1);
        uint32_t row = row_w, s = s_w; /* of unit `warp`; the next one is WARPS units on (no division in the loop) */
        for (uint32_t u = warp; u < units; u += WARPS) {
            const float2 x = s_tpos[row * 32 + lane];
            const float r = s_trad[row * 32 + lane];
            float2 ntx[1] = {make_float2(-x.x, -x.x)}, nty[1] = {make_float2(-x.y, -x.y)}, rho[1] = {make_float2(r, r)};
            float2 ax[1] = {make_float2(0.f, 0.f)}, ay[1] = {make_float2(0.f, 0.f)};
            const uint32_t pk0 = s * p.slab_packs;
            const uint32_t pk1 = min(pk0 + p.slab_packs, P);
#pragma unroll 1
            for (uint32_t q = pk0; q < pk1; q += UNR) {
#pragma unroll
                for (int w = 0; w < UNR; w++) {
                    const float4 xs = xy[2 * (q + w)];
                    const float4 ys = xy[2 * (q + w) + 1];
                    const float4 gs = s_gm[q + w];
                    interact_pack<1, kFormPacked>(xs, ys, gs, ntx, nty, rho, ax, ay);
                }
            }
            s_part[(size_t)u * 32 + lane] = make_float2(__fadd_rn(ax[0].x, ax[0].y), __fadd_rn(ay[0].x, ay[0].y));
            s += WARPS;
            while (s >= p.nslab) {
                s -= p.nslab;
                row++;
            }
        }
        tl_lap(0);
        __syncthreads();
        tl_lap(1);

        if (tid < nrow * 32) {
            const uint32_t row = tid >> 5;
            const float2 *sums = s_part + (size_t)row * p.nslab * 32 + lane;
            float2 sum = sums[0];
            uint32_t s = 1;
            for (; s + 8 <= p.nslab; s += 8) { /* eight loads in flight, added in slab order */
                float2 v[8];
#pragma unroll
                for (int w = 0; w < 8; w++) v[w] = sums[(size_t)(s + w) * 32];
#pragma unroll
                for (int w = 0; w < 8; w++) {
                    sum.x = __fadd_rn(sum.x, v[w].x);
                    sum.y = __fadd_rn(sum.y, v[w].y);
                }
            }
            for (; s < p.nslab; s++) {
                const float2 v = sums[(size_t)s * 32];
                sum.x = __fadd_rn(sum.x, v.x);
                sum.y = __fadd_rn(sum.y, v.y);
            }
            /* the integrator of integrate_store (force_kernel.cuh), state in registers */
            my_a = sum;
            my_v.x = __fadd_rn(my_v.x, __fmul_rn(sum.x, p.dt));
            my_v.y = __fadd_rn(my_v.y, __fmul_rn(sum.y, p.dt));
            my_x.x = __fadd_rn(my_x.x, __fmul_rn(my_v.x, p.dt));
            my_x.y = __fadd_rn(my_x.y, __fmul_rn(my_v.y, p.dt));
            s_tpos[tid] = my_x;
            if (mine && my_i < p.mass_len && step + 1 < rp.nsteps) {
                /* the next step's source buffer of every CTA of the cluster */
                const uint32_t o = ((cur ^ 1u) * P * 8u + (my_i >> 2) * 8u + (my_i & 3u)) * 4u;
                for (uint32_t c = 0; c < C; c++) {
                    st_cluster_f32(xy_local + o, c, my_x.x);
                    st_cluster_f32(xy_local + o + 16u, c, my_x.y);
                }
            }
        }
        tl_lap(2);
        if (step + 1 < rp.nsteps) cluster_sync_all(); /* all new positions have landed everywhere; also this CTA's barrier */
        tl_lap(3);
    }
    if (TL && tid == 0)
        for (int k = 0; k < 4; k++) nb_timeline[rank * 4 + k] = (unsigned long long)tl[k];

    /* the state after the last step goes back to global memory */
    if (mine) {
        p.acc[my_i] = my_a;
        p.vel[my_i] = my_v;
        p.pos[my_i] = my_x;
        if (my_i < p.mass_len) {
            const uint32_t o = (my_i >> 2) * 8u + (my_i & 3u);
#pragma unroll 1
            for (int d = 0; d < p.ndst; d++) {
                p.dst_xy[d][o] = my_x.x;
                p.dst_xy[d][o + 4] = my_x.y;
            }
        }
    }
}

}  // namespace nb
