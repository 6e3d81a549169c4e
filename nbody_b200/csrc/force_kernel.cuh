/*
 * force_kernel.cuh -- the hot path: all-pairs 2D softened gravity + semi-implicit Euler,
 * hand-written for sm_100a (B200).
 *
 * Replaces, in one kernel: the reference compute shader (ref src/shader/particle_cs.glsl:28-55),
 * its CPU twin PackedUpdate (ref src/lib/sim_cpu.c:156-194) and the per-step buffer copy of
 * PerformSimUpdate (ref src/lib/sim_gpu.c:314-333).
 *
 * Work decomposition.  Targets are cut into tiles of THREADS*T particles; sources [0, mass_len)
 * are cut into `nslab` slabs.  One CTA = one (target tile, slab) item: it streams the slab through
 * shared memory in tiles of TJP packs (1 pack = 4 sources: x[4], y[4], G*m[4]) with
 * cp.async.bulk (TMA 1-D bulk copy) into an NS-deep mbarrier ring, and every thread keeps T
 * targets in registers.  The pair arithmetic is issued as packed f32x2 instructions
 * (FADD2/FFMA2/FMUL2: two sources per instruction against one target) so that the FP32 pipe can be
 * saturated with about half of the issue slots, leaving the rest for MUFU.RSQ and LDS.128:
 *
 *     d   = s - t                      2 FADD2         (targets stored negated)
 *     r2  = dx*dx + (dy*dy + rho_t)    2 FFMA2         (softening = target radius, ref sim_cpu.c:176)
 *     inv = rsqrt(r2)                  2 MUFU.RSQ
 *     f   = (inv*inv) * (Gm*inv)       3 FMUL2
 *     a  += d*f                        2 FFMA2
 *
 * = 9 FP32-pipe instructions + 2 MUFU per TWO pairs, 14 algorithmic FLOP per pair.
 *
 * Summation order (what makes results independent of the number of GPUs): within a slab a thread
 * adds even-numbered and odd-numbered sources of its target into the two halves of an f32x2
 * accumulator, in source order; the slab partial is even+odd; slab partials are added in slab
 * order.  With nslab > 1 partials go through a scratch buffer and the LAST CTA to finish a target
 * tile (atomic ticket) adds them and integrates -- the integrator stays fused in this kernel.
 *
 * Integrator (ref sim_cpu.c:191-193, particle_cs.glsl:51-54): acc stored; vel += acc*dt;
 * pos += vel*dt with the new vel; each product and sum individually rounded like the reference's
 * non-FMA CPU build.  New positions of massive particles are also written to the NEXT source
 * buffer of every destination in dst_xy[] (this GPU, and peer GPUs over NVLink when they are
 * mapped), which is both the reference's storage[1]->storage[0] copy and the all-gather.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

/* Source-order knobs of the pair arithmetic.  Every setting computes bit-identical results; they only change the
 * order in which ptxas sees the operations, and with it the schedule of the pair loop, which is worth up to 2.8 %
 * (tools/schedule_lottery.py builds all 144 settings, 46 distinct schedules; profiles/r01_schedule_lottery.txt:
 * 156.3 .. 160.7 ms at N = 1M).  The defaults are the measured best. */
#ifndef NB_ORD_ACC
#define NB_ORD_ACC 1  /* 0: fma(dx, f, ax); fma(dy, f, ay)   1: fma(f, dx, ax); fma(f, dy, ay)   2: ay before ax */
#endif
#ifndef NB_ORD_R2
#define NB_ORD_R2 0   /* 0: r2 = dx*dx + (dy*dy + rho)   1: r2 = dy*dy + (dx*dx + rho) -- NOT bit-identical to 0  */
#endif
#ifndef NB_ORD_GI
#define NB_ORD_GI 0   /* 0: g * inv   1: inv * g */
#endif
#ifndef NB_ORD_F
#define NB_ORD_F 0    /* 0: inv2 * gi   1: gi * inv2 */
#endif
#ifndef NB_ORD_SUB
#define NB_ORD_SUB 0  /* 0: dx, dy   1: dy, dx   2: operands swapped */
#endif
#ifndef NB_ORD_LOOP
#define NB_ORD_LOOP 1 /* 0: per target {sources 0,1; sources 2,3}   1: sources 0,1 for all targets, then 2,3 */
#endif

/* The same knobs for the GROUPED build of the kernel (step_kernel.cuh): its loop is scheduled by ptxas on its own,
 * so its best source order is measured on its own (tools/schedule_lottery.py --grouped: 24 distinct schedules,
 * 158.7 .. 163.2 ms at N = 1M against 156.1 ms for the ungrouped kernel, profiles/r02_schedule_lottery_grouped.txt;
 * the ungrouped kernel's own best order gives the grouped one 161.1 ms). */
#ifndef NB_ORDG_ACC
#define NB_ORDG_ACC 0
#endif
#ifndef NB_ORDG_GI
#define NB_ORDG_GI 0
#endif
#ifndef NB_ORDG_F
#define NB_ORDG_F 1
#endif
#ifndef NB_ORDG_SUB
#define NB_ORDG_SUB 0
#endif
#ifndef NB_ORDG_LOOP
#define NB_ORDG_LOOP 0
#endif

namespace nb {

/* a source order as one template argument: decimal digits ACC R2 GI LOOP F SUB */
__host__ __device__ constexpr int ord_pack(int acc, int r2, int gi, int loop, int f, int sub) {
    return ((((acc * 10 + r2) * 10 + gi) * 10 + loop) * 10 + f) * 10 + sub;
}
constexpr int kOrdDefault = ord_pack(NB_ORD_ACC, NB_ORD_R2, NB_ORD_GI, NB_ORD_LOOP, NB_ORD_F, NB_ORD_SUB);
/* R2 changes the rounding, so the grouped build shares it with the default */
constexpr int kOrdGrouped = ord_pack(NB_ORDG_ACC, NB_ORD_R2, NB_ORDG_GI, NB_ORDG_LOOP, NB_ORDG_F, NB_ORDG_SUB);
__host__ __device__ constexpr int ord_acc(int o) { return o / 100000 % 10; }
__host__ __device__ constexpr int ord_r2(int o) { return o / 10000 % 10; }
__host__ __device__ constexpr int ord_gi(int o) { return o / 1000 % 10; }
__host__ __device__ constexpr int ord_loop(int o) { return o / 100 % 10; }
__host__ __device__ constexpr int ord_f(int o) { return o / 10 % 10; }
__host__ __device__ constexpr int ord_sub(int o) { return o % 10; }

constexpr int kPack = 4;            /* sources per pack                                  */
constexpr int kMaxDst = 8;          /* destinations for new source positions (GPUs)      */
constexpr float kPadPos = 1.0e15f;  /* position of padding sources (G*m = 0): r2 finite,  */
                                    /* inv^3 finite, contribution exactly +-0              */

struct StepParams {
    const float4 *src_xy;     /* [src_packs_pad][2]  xs, ys of 4 sources (current step)  */
    const float4 *src_gm;     /* [src_packs_pad]     G*mass of 4 sources                  */
    float *dst_xy[kMaxDst];   /* next-step source buffers (as floats), self first         */
    int ndst;
    uint32_t fold_tiles;      /* grouped kernels only (step_kernel GROUPED): source tiles per slab; a CTA walks */
                              /* several consecutive slabs and folds each slab's sum into a running sum.  Sits   */
                              /* in what used to be alignment padding: the layout the other kernels see is       */
                              /* unchanged (an extra parameter elsewhere re-ordered their pair loop)  */
    float2 *pos;              /* [total_len] target positions, updated in place           */
    float2 *vel;              /* [total_len]                                              */
    float2 *acc;              /* [total_len]                                              */
    const float *radius;      /* [total_len]                                              */
    float2 *part;             /* slab partial sums (nslab > 1): part[slab * part_stride + i] for particle i.  The host */
                              /* biases the pointer so that the first particle of the launch lands in slot 0      */
    unsigned *tile_cnt;       /* [tiles] arrival tickets, zero between steps              */
    uint32_t total_len, mass_len;
    uint32_t part_stride;     /* targets this GPU computes, in whole tiles                */
    uint32_t src_packs_pad;   /* source packs incl. padding, multiple of TJP              */
    uint32_t slab_packs;      /* packs per CTA: one slab, or one group of slabs (GROUPED) */
    uint32_t nslab;           /* CTAs per target tile: slabs, or groups of slabs (GROUPED) */
    uint32_t tile_first;      /* first target tile of this launch                         */
    float dt;
};

/* ---- PTX helpers: mbarrier, bulk copy, packed math ---------------------------------- */

__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NB_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NB_DONE_%=;\n"
        "bra NB_WAIT_%=;\n"
        "NB_DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
/* TMA 1-D bulk copy global -> shared, completion counted in bytes on an mbarrier. */
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ float rsqrt_mufu(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

/*
 * Instruction forms of the pair arithmetic (same operations, same rounding, same summation order --
 * results are bit-identical between forms; only the SASS differs):
 *   kFormPacked      everything as f32x2 (FADD2/FFMA2/FMUL2)
 *   kFormScalarAcc   packed up to f, the two accumulate FMAs as four scalar FFMA: an FFMA2 that reads
 *                    three distinct register pairs (dx, f, acc) costs the SM sub-partition three
 *                    dispatch cycles instead of two (tools/tune.cu `pipes`)
 *   kFormScalar      everything scalar
 */
enum {
    kFormPacked = 0,
    kFormTargetPacked = 1,
    kFormScalarAcc = 2,
    kFormScalar = 3,
    /* ablations for tools/tune.cu ONLY (wrong results on purpose; they isolate one cost of the dispatch
     * model of DESIGN.md section 5): */
    kAblateMufu = 4,     /* inv = r2 instead of rsqrt(r2): removes the 2 MUFU dispatch cycles          */
    kAblateAccRead = 5,  /* ax += f*f, ay += dy*dy: each accumulate reads 2 register pairs, not 3       */
    kAblateBoth = 6,
    kAblateReduce = 7,   /* every slab CTA integrates its own partial: no partial store, ticket, reduction */
    /* packed arithmetic; the slab CTAs of a target tile form one thread-block cluster and add their partials
     * through distributed shared memory instead of global memory + ticket (cluster_reduce.cuh).
     * NBODY_VARIANT only: measured slower than the ticket path at every size tried (cluster_reduce.cuh). */
    kFormCluster = 8
};

template <int FORM, int ORD = kOrdDefault>
__device__ __forceinline__ void interact_two(const float2 x, const float2 y, const float2 g, const float2 ntx,
                                             const float2 nty, const float2 rho, float2 &ax, float2 &ay) {
    if constexpr (FORM == kFormScalar) {
        const float dx0 = __fadd_rn(x.x, ntx.x), dx1 = __fadd_rn(x.y, ntx.y);
        const float dy0 = __fadd_rn(y.x, nty.x), dy1 = __fadd_rn(y.y, nty.y);
        float r20 = __fmaf_rn(dy0, dy0, rho.x), r21 = __fmaf_rn(dy1, dy1, rho.y);
        r20 = __fmaf_rn(dx0, dx0, r20);
        r21 = __fmaf_rn(dx1, dx1, r21);
        const float i0 = rsqrt_mufu(r20), i1 = rsqrt_mufu(r21);
        const float f0 = __fmul_rn(__fmul_rn(i0, i0), __fmul_rn(g.x, i0));
        const float f1 = __fmul_rn(__fmul_rn(i1, i1), __fmul_rn(g.y, i1));
        ax.x = __fmaf_rn(dx0, f0, ax.x);
        ax.y = __fmaf_rn(dx1, f1, ax.y);
        ay.x = __fmaf_rn(dy0, f0, ay.x);
        ay.y = __fmaf_rn(dy1, f1, ay.y);
    } else {
        float2 dx, dy;
        if constexpr (ord_sub(ORD) == 1) {
            dy = __fadd2_rn(y, nty);
            dx = __fadd2_rn(x, ntx);
        } else if constexpr (ord_sub(ORD) == 2) {
            dx = __fadd2_rn(ntx, x);
            dy = __fadd2_rn(nty, y);
        } else {
            dx = __fadd2_rn(x, ntx);
            dy = __fadd2_rn(y, nty);
        }
        float2 r2;
        if constexpr (ord_r2(ORD) == 1) {
            r2 = __ffma2_rn(dx, dx, rho);
            r2 = __ffma2_rn(dy, dy, r2);
        } else {
            r2 = __ffma2_rn(dy, dy, rho);
            r2 = __ffma2_rn(dx, dx, r2);
        }
        float2 inv;
        if constexpr (FORM == kAblateMufu || FORM == kAblateBoth) {
            inv = r2;
        } else {
            inv.x = rsqrt_mufu(r2.x);
            inv.y = rsqrt_mufu(r2.y);
        }
        const float2 inv2 = __fmul2_rn(inv, inv);
        float2 gi, f;
        if constexpr (ord_gi(ORD) == 1) gi = __fmul2_rn(inv, g);
        else gi = __fmul2_rn(g, inv);
        if constexpr (ord_f(ORD) == 1) f = __fmul2_rn(gi, inv2);
        else f = __fmul2_rn(inv2, gi);
        if constexpr (FORM == kAblateAccRead || FORM == kAblateBoth) {
            /* same instruction count and dependences on dx (through r2), dy and f, but each accumulate
             * FFMA2 reads two distinct register pairs instead of three */
            ax = __ffma2_rn(f, f, ax);
            ay = __ffma2_rn(dy, dy, ay);
        } else if constexpr (FORM == kFormScalarAcc) {
            ax.x = __fmaf_rn(dx.x, f.x, ax.x);
            ax.y = __fmaf_rn(dx.y, f.y, ax.y);
            ay.x = __fmaf_rn(dy.x, f.x, ay.x);
            ay.y = __fmaf_rn(dy.y, f.y, ay.y);
        } else {
            if constexpr (ord_acc(ORD) == 1) {
                ax = __ffma2_rn(f, dx, ax);
                ay = __ffma2_rn(f, dy, ay);
            } else if constexpr (ord_acc(ORD) == 2) {
                ay = __ffma2_rn(dy, f, ay);
                ax = __ffma2_rn(dx, f, ax);
            } else {
                ax = __ffma2_rn(dx, f, ax);
                ay = __ffma2_rn(dy, f, ay);
            }
        }
    }
}

/* 4 sources (one pack) against T register-resident targets. */
template <int T, int FORM = kFormPacked, int ORD = kOrdDefault>
__device__ __forceinline__ void interact_pack(const float4 xs, const float4 ys, const float4 gs,
                                              const float2 (&ntx)[T], const float2 (&nty)[T],
                                              const float2 (&rho)[T], float2 (&ax)[T], float2 (&ay)[T]) {
    const float2 x01 = make_float2(xs.x, xs.y), x23 = make_float2(xs.z, xs.w);
    const float2 y01 = make_float2(ys.x, ys.y), y23 = make_float2(ys.z, ys.w);
    const float2 g01 = make_float2(gs.x, gs.y), g23 = make_float2(gs.z, gs.w);
    if constexpr (ord_loop(ORD) == 1) {
        /* sources 0,1 against every target, then sources 2,3: per target the same order either way */
#pragma unroll
        for (int t = 0; t < T; t++) interact_two<FORM, ORD>(x01, y01, g01, ntx[t], nty[t], rho[t], ax[t], ay[t]);
#pragma unroll
        for (int t = 0; t < T; t++) interact_two<FORM, ORD>(x23, y23, g23, ntx[t], nty[t], rho[t], ax[t], ay[t]);
    } else {
#pragma unroll
        for (int t = 0; t < T; t++) {
            interact_two<FORM, ORD>(x01, y01, g01, ntx[t], nty[t], rho[t], ax[t], ay[t]);
            interact_two<FORM, ORD>(x23, y23, g23, ntx[t], nty[t], rho[t], ax[t], ay[t]);
        }
    }
}

/*
 * Target-packed form: an f32x2 holds TWO targets against ONE source (the source is a warp-uniform
 * scalar operand).  Same 9 packed instructions + 2 MUFU per two pairs, but one accumulator pair per
 * two targets (half the accumulator registers of the source-packed form, so more targets fit per
 * thread and every LDS.128 of sources feeds twice as many pairs) and each target sums its sources
 * in plain source order.
 */
template <int T2>
__device__ __forceinline__ void interact_pack_tp(const float4 xs, const float4 ys, const float4 gs,
                                                 const float2 (&ntx)[T2], const float2 (&nty)[T2],
                                                 const float2 (&rho)[T2], float2 (&ax)[T2], float2 (&ay)[T2]) {
    const float sx[4] = {xs.x, xs.y, xs.z, xs.w};
    const float sy[4] = {ys.x, ys.y, ys.z, ys.w};
    const float sg[4] = {gs.x, gs.y, gs.z, gs.w};
#pragma unroll
    for (int l = 0; l < 4; l++) {
        const float2 x2 = make_float2(sx[l], sx[l]);
        const float2 y2 = make_float2(sy[l], sy[l]);
        const float2 g2 = make_float2(sg[l], sg[l]);
#pragma unroll
        for (int t = 0; t < T2; t++) interact_two<kFormPacked>(x2, y2, g2, ntx[t], nty[t], rho[t], ax[t], ay[t]);
    }
}

/* Integrate one target and publish it (ref sim_cpu.c:191-193). */
__device__ __forceinline__ void integrate_store(const StepParams &p, uint32_t i, float2 a) {
    float2 v = p.vel[i];
    float2 x = p.pos[i];
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

}  // namespace nb
