/*
 * tools/tune.cu -- measurement harness for the hot kernel (development tool, not product).
 *
 *   tune pipes              FP32 / MUFU pipe micro-benchmarks (what the roofline rests on)
 *   tune sweep [N] [M]      time every compiled step_kernel variant on random particles
 *
 * Numbers are CUDA-event times after warm-up; the effective SM clock is derived in-kernel from
 * clock64() against %globaltimer so that fractions of peak can be quoted at the clock that was
 * actually running.
 */
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "step_kernel.cuh"
#include "util.cuh"

using namespace nb;

/* ------------------------------------------------------------------------------------------ */
/* pipe micro-benchmarks                                                                        */

__device__ __forceinline__ unsigned long long gtimer() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

struct ClockRec {
    unsigned long long cycles, ns;
};

/* MODE 0: FFMA (scalar) x ILP chains; 1: FFMA2; 2: MUFU.RSQ; 3: mix 9 packed FP32 : 2 MUFU (the
 * kernel's ratio); 4: FFMA2 + FMUL2 + FADD2 mix 4:3:2 without MUFU */
template <int MODE>
__global__ void __launch_bounds__(256) pipe_kernel(float *out, int iters, ClockRec *rec) {
    float2 a[8], b, c, bb[8], cc[8];
    for (int k = 0; k < 8; k++) a[k] = make_float2(threadIdx.x * 0.001f + k, threadIdx.x * 0.002f - k);
    for (int k = 0; k < 8; k++) {
        bb[k] = make_float2(1.0f + 1e-6f * (threadIdx.x + k), 1.0f - 1e-6f * (threadIdx.x + k));
        cc[k] = make_float2(1e-5f * (threadIdx.x - k), -1e-5f * (threadIdx.x + 2 * k));
    }
    unsigned ia = threadIdx.x, ib = blockIdx.x;
    b = make_float2(1.0001f, 0.9999f);
    c = make_float2(0.0001f, -0.0001f);
    const unsigned long long t0 = gtimer();
    const long long c0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            if (MODE == 0) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    a[k].x = fmaf(a[k].x, b.x, c.x);
                    a[k].y = fmaf(a[k].y, b.y, c.y);
                }
            } else if (MODE == 1) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], b, c);
            } else if (MODE == 2) {
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    a[k].x = rsqrt_mufu(a[k].x);
                    a[k].y = rsqrt_mufu(a[k].y);
                }
            } else if (MODE == 5) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __fmul2_rn(a[k], b);
            } else if (MODE == 6) {
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __fadd2_rn(a[k], c);
            } else if (MODE == 7) {
                /* three distinct register pairs per instruction, nothing for the reuse cache */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], bb[k], cc[k]);
            } else if (MODE == 8) {
                /* two distinct pairs + one shared */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], bb[k], c);
            } else if (MODE == 12) {
                /* scalar FFMA, three distinct registers */
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    a[k].x = fmaf(a[k].x, bb[k].x, cc[k].x);
                    a[k].y = fmaf(a[k].y, bb[k].y, cc[k].y);
                }
            } else if (MODE == 13) {
                /* the kernel's accumulate pair: two FFMA2 sharing one operand (3 + 2 distinct pairs) */
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    a[k] = __ffma2_rn(bb[k], cc[k], a[k]);
                    a[k + 4] = __ffma2_rn(bb[k + 4], cc[k], a[k + 4]);
                }
            } else if (MODE == 14) {
                /* same work as four scalar FFMA */
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    a[k].x = fmaf(bb[k].x, cc[k].x, a[k].x);
                    a[k].y = fmaf(bb[k].y, cc[k].y, a[k].y);
                    a[k + 4].x = fmaf(bb[k + 4].x, cc[k].x, a[k + 4].x);
                    a[k + 4].y = fmaf(bb[k + 4].y, cc[k].y, a[k + 4].y);
                }
            } else if (MODE == 15) {
                /* FFMA2 reading two register pairs and one scalar broadcast to both halves */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(bb[k], make_float2(cc[k].x, cc[k].x), a[k]);
            } else if (MODE == 16) {
                /* "transposed" pair arithmetic: an f32x2 is (x, y) of ONE pair; scalar r2 / f chain; one
                 * FFMA2 (d, f broadcast, acc) accumulates both components.  8 pairs per round. */
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const float2 d = __fadd2_rn(a[k], b);                  /* (dx, dy)            */
                    float r2 = fmaf(d.y, d.y, c.x);
                    r2 = fmaf(d.x, d.x, r2);
                    const float inv = rsqrt_mufu(r2);
                    const float f = (inv * inv) * (c.y * inv);
                    a[k] = __ffma2_rn(d, make_float2(f, f), a[k]);
                }
            } else if (MODE == 17) {
                /* same without the MUFU */
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const float2 d = __fadd2_rn(a[k], b);
                    float r2 = fmaf(d.y, d.y, c.x);
                    r2 = fmaf(d.x, d.x, r2);
                    const float f = (r2 * r2) * (c.y * r2);
                    a[k] = __ffma2_rn(d, make_float2(f, f), a[k]);
                }
            } else if (MODE == 9) {
                /* 8 FFMA2 + 2 independent MUFU: does a MUFU issue cost the FMA pipe a cycle? */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], b, c);
                cc[0].x = rsqrt_mufu(cc[0].x);
                cc[1].x = rsqrt_mufu(cc[1].x);
            } else if (MODE == 10) {
                /* 8 FFMA2 + 2 independent integer ALU ops */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], b, c);
                ia = (ia ^ (ia >> 3)) + 0x9e3779b9u;
                ib = (ib ^ (ib << 5)) + 0x7f4a7c15u;
            } else if (MODE == 11) {
                /* 8 FFMA2 + 4 MUFU */
#pragma unroll
                for (int k = 0; k < 8; k++) a[k] = __ffma2_rn(a[k], b, c);
                cc[0].x = rsqrt_mufu(cc[0].x);
                cc[1].x = rsqrt_mufu(cc[1].x);
                cc[2].x = rsqrt_mufu(cc[2].x);
                cc[3].x = rsqrt_mufu(cc[3].x);
            } else if (MODE == 3) {
                /* per chain pair: 9 packed + 2 MUFU, four independent chains */
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    float2 dx = __fadd2_rn(a[k], b);
                    float2 dy = __fadd2_rn(a[k + 4], c);
                    float2 r2 = __ffma2_rn(dy, dy, b);
                    r2 = __ffma2_rn(dx, dx, r2);
                    float2 inv;
                    inv.x = rsqrt_mufu(r2.x);
                    inv.y = rsqrt_mufu(r2.y);
                    float2 i2 = __fmul2_rn(inv, inv);
                    float2 gi = __fmul2_rn(c, inv);
                    float2 f = __fmul2_rn(i2, gi);
                    a[k] = __ffma2_rn(dx, f, a[k]);
                    a[k + 4] = __ffma2_rn(dy, f, a[k + 4]);
                }
            } else {
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    float2 dx = __fadd2_rn(a[k], b);
                    float2 dy = __fadd2_rn(a[k + 4], c);
                    float2 r2 = __ffma2_rn(dy, dy, b);
                    r2 = __ffma2_rn(dx, dx, r2);
                    float2 i2 = __fmul2_rn(r2, r2);
                    float2 gi = __fmul2_rn(c, r2);
                    float2 f = __fmul2_rn(i2, gi);
                    a[k] = __ffma2_rn(dx, f, a[k]);
                    a[k + 4] = __ffma2_rn(dy, f, a[k + 4]);
                }
            }
        }
    }
    const long long c1 = clock64();
    const unsigned long long t1 = gtimer();
    float s = 0.f;
    for (int k = 0; k < 8; k++) s += a[k].x + a[k].y;
    if (MODE >= 7)
        for (int k = 0; k < 8; k++) s += bb[k].x * cc[k].y + cc[k].x;
    s += (float)(ia + ib);
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) {
        rec[blockIdx.x].cycles = (unsigned long long)(c1 - c0);
        rec[blockIdx.x].ns = t1 - t0;
    }
}

template <int MODE>
static void run_pipe(const char *name, double ops_per_iter_per_thread, const char *unit, int sms, int ctas_per_sm) {
    const int grid = sms * ctas_per_sm, block = 256, iters = 20000;
    float *out;
    ClockRec *rec;
    NB_CUDA(cudaMalloc(&out, (size_t)grid * block * sizeof(float)));
    NB_CUDA(cudaMalloc(&rec, (size_t)grid * sizeof(ClockRec)));
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    pipe_kernel<MODE><<<grid, block>>>(out, 2000, rec);
    NB_CUDA(cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 3; r++) {
        cudaEventRecord(e0);
        pipe_kernel<MODE><<<grid, block>>>(out, iters, rec);
        cudaEventRecord(e1);
        NB_CUDA(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = std::min(best, ms);
    }
    std::vector<ClockRec> h(grid);
    NB_CUDA(cudaMemcpy(h.data(), rec, grid * sizeof(ClockRec), cudaMemcpyDeviceToHost));
    double cyc = 0, ns = 0;
    for (auto &r : h) {
        cyc += r.cycles;
        ns += r.ns;
    }
    const double mhz = cyc / ns * 1e3;
    const double total = ops_per_iter_per_thread * iters * (double)grid * block;
    const double per_clk_sm = total / (best * 1e-3 * mhz * 1e6) / sms; /* ops per SM clock per SM */
    printf("  %-34s ctas/sm %d  %8.3f ms  %9.2f G%s/s  %7.2f %s/clk/SM  (SM clock %.0f MHz)\n", name, ctas_per_sm, best,
           total / best * 1e-6, unit, per_clk_sm, unit, mhz);
    cudaFree(out);
    cudaFree(rec);
}

static void pipes(int sms) {
    printf("pipe micro-benchmarks (256 threads/CTA):\n");
    for (int c : {2, 8}) {
        run_pipe<0>("FFMA scalar (fma lanes)", 4 * 16.0, "fma", sms, c);
        run_pipe<1>("FFMA2 packed (fma lanes)", 4 * 16.0, "fma", sms, c);
        run_pipe<5>("FMUL2 packed (lanes)", 4 * 16.0, "mul", sms, c);
        run_pipe<6>("FADD2 packed (lanes)", 4 * 16.0, "add", sms, c);
        run_pipe<7>("FFMA2, 3 distinct register pairs", 4 * 16.0, "fma", sms, c);
        run_pipe<8>("FFMA2, 2 distinct pairs + 1 shared", 4 * 16.0, "fma", sms, c);
        run_pipe<12>("FFMA scalar, 3 distinct registers", 4 * 16.0, "fma", sms, c);
        run_pipe<13>("FFMA2 pairs sharing one operand", 4 * 16.0, "fma", sms, c);
        run_pipe<14>("same as scalar FFMA x4", 4 * 16.0, "fma", sms, c);
        run_pipe<15>("FFMA2, 2 pairs + 1 scalar broadcast", 4 * 16.0, "fma", sms, c);
        run_pipe<16>("transposed mix incl. MUFU (pairs)", 4 * 8.0, "pair", sms, c);
        run_pipe<17>("transposed mix without MUFU (pairs)", 4 * 8.0, "pair", sms, c);
        run_pipe<9>("8 FFMA2 + 2 MUFU (fma lanes)", 4 * 16.0, "fma", sms, c);
        run_pipe<11>("8 FFMA2 + 4 MUFU (fma lanes)", 4 * 16.0, "fma", sms, c);
        run_pipe<10>("8 FFMA2 + 2 int ALU (fma lanes)", 4 * 16.0, "fma", sms, c);
        run_pipe<2>("MUFU.RSQ", 4 * 16.0, "rsq", sms, c);
        run_pipe<4>("FP32 mix 4 FFMA2:3 FMUL2:2 FADD2 (pairs)", 4 * 8.0, "pair", sms, c);
        run_pipe<3>("kernel mix 9 packed + 2 MUFU (pairs)", 4 * 8.0, "pair", sms, c);
    }
}

/* ------------------------------------------------------------------------------------------ */
/* step_kernel variant sweep                                                                    */

struct Buffers {
    float4 *src_xy[2], *src_gm;
    float2 *pos, *vel, *acc, *part;
    float *radius;
    unsigned *cnt;
    uint32_t N, M, src_packs, src_packs_pad;
    size_t part_elems;
};

static float frand() { return (float)rand() / (float)RAND_MAX; }

static Buffers make_buffers(uint32_t N, uint32_t M) {
    Buffers b;
    b.N = N;
    b.M = M;
    b.src_packs = (M + 3) / 4;
    b.src_packs_pad = (b.src_packs + 255) / 256 * 256;
    std::vector<float> xy((size_t)b.src_packs_pad * 8), gm((size_t)b.src_packs_pad * 4);
    std::vector<float2> pos(N), vel(N);
    std::vector<float> rad(N);
    for (uint32_t i = 0; i < N; i++) {
        pos[i] = make_float2((frand() - 0.5f) * 2e5f, (frand() - 0.5f) * 2e5f);
        vel[i] = make_float2(0.f, 0.f);
        rad[i] = 0.5f + frand() * 9.f;
    }
    for (uint32_t j = 0; j < b.src_packs_pad * 4; j++) {
        const uint32_t o = (j >> 2) * 8 + (j & 3);
        if (j < M) {
            xy[o] = pos[j].x;
            xy[o + 4] = pos[j].y;
            gm[j] = 10.f * (100.f + frand() * 1e4f);
        } else {
            xy[o] = kPadPos;
            xy[o + 4] = kPadPos;
            gm[j] = 0.f;
        }
    }
    NB_CUDA(cudaMalloc(&b.src_xy[0], xy.size() * 4));
    NB_CUDA(cudaMalloc(&b.src_xy[1], xy.size() * 4));
    NB_CUDA(cudaMalloc(&b.src_gm, gm.size() * 4));
    NB_CUDA(cudaMalloc(&b.pos, N * sizeof(float2)));
    NB_CUDA(cudaMalloc(&b.vel, N * sizeof(float2)));
    NB_CUDA(cudaMalloc(&b.acc, N * sizeof(float2)));
    NB_CUDA(cudaMalloc(&b.radius, N * sizeof(float)));
    b.part_elems = (size_t)512 * (N + 4096);
    NB_CUDA(cudaMalloc(&b.part, b.part_elems * sizeof(float2)));
    NB_CUDA(cudaMalloc(&b.cnt, (N / 64 + 16) * sizeof(unsigned)));
    NB_CUDA(cudaMemset(b.cnt, 0, (N / 64 + 16) * sizeof(unsigned)));
    NB_CUDA(cudaMemcpy(b.src_xy[0], xy.data(), xy.size() * 4, cudaMemcpyHostToDevice));
    NB_CUDA(cudaMemcpy(b.src_xy[1], xy.data(), xy.size() * 4, cudaMemcpyHostToDevice));
    NB_CUDA(cudaMemcpy(b.src_gm, gm.data(), gm.size() * 4, cudaMemcpyHostToDevice));
    NB_CUDA(cudaMemcpy(b.pos, pos.data(), N * sizeof(float2), cudaMemcpyHostToDevice));
    NB_CUDA(cudaMemcpy(b.vel, vel.data(), N * sizeof(float2), cudaMemcpyHostToDevice));
    NB_CUDA(cudaMemcpy(b.radius, rad.data(), N * sizeof(float), cudaMemcpyHostToDevice));
    return b;
}

static double g_peak_pairs = 0; /* FP32-pipe ceiling in pairs/s at max clock: SMs*128*clk/9 */
static double g_peak_flops = 0;

template <int THREADS, int T, int TJP, int NS, int MINB, int UNR, int FORM = 0>
static void run_variant(Buffers &b, uint32_t nslab_req, int sms) {
    auto kern = step_kernel<THREADS, T, TJP, NS, MINB, UNR, FORM>;
    const size_t smem = step_smem_bytes<TJP, NS>();
    NB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    NB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem));
    cudaFuncAttributes fa;
    NB_CUDA(cudaFuncGetAttributes(&fa, kern));
    const uint32_t tile = THREADS * T;
    const uint32_t ntiles = (b.N + tile - 1) / tile;
    uint32_t slab_packs = ((b.src_packs + nslab_req - 1) / nslab_req + 3) / 4 * 4;
    uint32_t nslab = (b.src_packs + slab_packs - 1) / slab_packs;
    if ((size_t)nslab * ntiles * tile > b.part_elems) {
        printf("  (skip: partial buffer too small)\n");
        return;
    }
    StepParams p;
    memset(&p, 0, sizeof p);
    p.src_xy = b.src_xy[0];
    p.src_gm = b.src_gm;
    p.dst_xy[0] = (float *)b.src_xy[1];
    p.ndst = 1;
    p.pos = b.pos; p.vel = b.vel; p.acc = b.acc; p.radius = b.radius;
    p.part = b.part; p.tile_cnt = b.cnt;
    p.total_len = b.N; p.mass_len = b.M;
    p.part_stride = ntiles * tile;
    p.src_packs_pad = (b.src_packs + TJP - 1) / TJP * TJP;
    p.slab_packs = slab_packs; p.nslab = nslab;
    p.tile_first = 0;
    p.dt = 0.f; /* dt = 0: positions stay put, every repetition does identical work */
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const uint32_t grid = ntiles * nslab;
    for (int w = 0; w < 2; w++) kern<<<grid, THREADS, smem>>>(p);
    NB_CUDA(cudaDeviceSynchronize());
    float best = 1e30f, sum = 0.f;
    const int reps = 5;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(e0);
        kern<<<grid, THREADS, smem>>>(p);
        cudaEventRecord(e1);
        NB_CUDA(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = std::min(best, ms);
        sum += ms;
    }
    const double pairs = (double)b.N * b.M;
    const double rate = pairs / (best * 1e-3);
    const double waves = (double)grid / (occ * sms);
    printf("  %s t%dx%d tjp%d ns%d unr%d minb%d | regs %3d occ %d | nslab %4u grid %6u waves %7.2f | best %8.3f ms avg %8.3f | "
           "%7.1f Gpairs/s  %5.1f%% of FP32 peak (14 flop/pair)  %5.1f%% of 9-slot ceiling\n",
           FORM == 0 ? "packed" : FORM == 1 ? "tgtpack" : FORM == 2 ? "sclacc" : FORM == 3 ? "scalar" : FORM == 4 ? "ABL-mufu" : FORM == 5 ? "ABL-acc" : FORM == 6 ? "ABL-both" : "ABL-reduce", THREADS, T, TJP, NS, UNR, MINB, fa.numRegs, occ, nslab, grid, waves, best, sum / reps, rate * 1e-9,
           100.0 * rate * 14.0 / g_peak_flops, 100.0 * rate / g_peak_pairs);
    fflush(stdout);
}

/* The GROUPED build of a plan candidate: NSLAB_REQ slabs (whole source tiles) walked by NGROUPS CTAs per tile. */
template <int THREADS, int T, int TJP, int NS, int MINB, int UNR>
static void run_variant_grouped(Buffers &b, uint32_t nslab_req, uint32_t ngroups_req, int sms) {
    auto kern = step_kernel<THREADS, T, TJP, NS, MINB, UNR, kFormPacked, true>;
    const size_t smem = step_smem_bytes_grouped<THREADS, T, TJP, NS>();
    NB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int occ = 0;
    NB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, THREADS, smem));
    cudaFuncAttributes fa;
    NB_CUDA(cudaFuncGetAttributes(&fa, kern));
    const uint32_t tile = THREADS * T;
    const uint32_t ntiles = (b.N + tile - 1) / tile;
    uint32_t slab_packs = std::max<uint32_t>(TJP, ((b.src_packs + nslab_req - 1) / nslab_req) / TJP * TJP);
    const uint32_t nslab = (b.src_packs + slab_packs - 1) / slab_packs;
    const uint32_t gs = (nslab + ngroups_req - 1) / ngroups_req;
    const uint32_t groups = (nslab + gs - 1) / gs;
    StepParams p;
    memset(&p, 0, sizeof p);
    p.src_xy = b.src_xy[0];
    p.src_gm = b.src_gm;
    p.dst_xy[0] = (float *)b.src_xy[1];
    p.ndst = 1;
    p.pos = b.pos; p.vel = b.vel; p.acc = b.acc; p.radius = b.radius;
    p.part = b.part; p.tile_cnt = b.cnt;
    p.total_len = b.N; p.mass_len = b.M;
    p.part_stride = ntiles * tile;
    p.src_packs_pad = (b.src_packs + TJP - 1) / TJP * TJP;
    p.slab_packs = slab_packs * gs; p.nslab = groups;
    p.fold_tiles = slab_packs / TJP;
    p.tile_first = 0;
    p.dt = 0.f;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const uint32_t grid = ntiles * groups;
    for (int w = 0; w < 2; w++) kern<<<grid, THREADS, smem>>>(p);
    NB_CUDA(cudaDeviceSynchronize());
    float best = 1e30f, sum = 0.f;
    const int reps = 5;
    for (int r = 0; r < reps; r++) {
        cudaEventRecord(e0);
        kern<<<grid, THREADS, smem>>>(p);
        cudaEventRecord(e1);
        NB_CUDA(cudaDeviceSynchronize());
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        best = std::min(best, ms);
        sum += ms;
    }
    const double rate = (double)b.N * b.M / (best * 1e-3);
    printf("  grouped t%dx%d tjp%d ns%d unr%d minb%d | regs %3d occ %d | nslab %4u in %3u groups of %3u, grid %6u waves %7.2f | "
           "best %8.3f ms avg %8.3f | %7.1f Gpairs/s  %5.1f%% of FP32 peak (14 flop/pair)\n",
           THREADS, T, TJP, NS, UNR, MINB, fa.numRegs, occ, nslab, groups, gs, grid, (double)grid / (occ * sms), best,
           sum / reps, rate * 1e-9, 100.0 * rate * 14.0 / g_peak_flops);
    fflush(stdout);
}

static void sweep(uint32_t N, uint32_t M, int sms, const char *nslab_list) {
    printf("variant sweep: N = %u targets, M = %u sources, %.3e pairs/step\n", N, M, (double)N * M);
    Buffers b = make_buffers(N, M);
    const uint32_t base = std::max(1u, (uint32_t)(24.0 * sms * 2 / ((N + 1023) / 1024)));
    const char *set = getenv("TUNE_SET");
    const bool full = set && !strcmp(set, "full");
    const bool small = set && !strcmp(set, "small");
    std::vector<uint32_t> list;
    if (nslab_list) {
        for (const char *c = nslab_list; *c;) {
            list.push_back((uint32_t)strtoul(c, (char **)&c, 10));
            if (*c == ',') c++;
        }
    } else {
        list.push_back(base);
    }
    for (uint32_t ns : list) {
        if (set && !strcmp(set, "ship")) { /* the large-N kernel only */
            run_variant<512, 4, 128, 4, 1, 2, 0>(b, ns, sms);
            continue;
        }
        if (set && !strcmp(set, "ship-grouped")) { /* its GROUPED build: TUNE_GROUPS CTAs per tile (default 13) */
            const char *g = getenv("TUNE_GROUPS");
            run_variant_grouped<512, 4, 128, 4, 1, 2>(b, ns, g ? (uint32_t)atoi(g) : 13u, sms);
            continue;
        }
#ifndef TUNE_MINIMAL
        if (set && !strcmp(set, "warps")) { /* 16 .. 24 resident warps per SM at 4 targets per thread */
            run_variant<512, 4, 128, 4, 1, 2>(b, ns, sms);
            run_variant<576, 4, 128, 4, 1, 2>(b, ns, sms);
            run_variant<192, 4, 128, 4, 3, 2>(b, ns, sms);
            run_variant<640, 4, 128, 4, 1, 2>(b, ns, sms);
            run_variant<640, 4, 128, 4, 1, 1>(b, ns, sms);
            run_variant<128, 4, 128, 4, 5, 2>(b, ns, sms);
            run_variant<128, 4, 128, 4, 5, 1>(b, ns, sms);
            run_variant<320, 4, 128, 4, 2, 1>(b, ns, sms);
            run_variant<768, 4, 128, 4, 1, 1>(b, ns, sms);
            continue;
        }
        if (set && !strcmp(set, "ablate")) {
            run_variant<512, 4, 128, 4, 1, 2, 0>(b, ns, sms);
            run_variant<512, 4, 128, 4, 1, 2, 4>(b, ns, sms);
            run_variant<512, 4, 128, 4, 1, 2, 5>(b, ns, sms);
            run_variant<512, 4, 128, 4, 1, 2, 6>(b, ns, sms);
            continue;
        }
        if (set && !strcmp(set, "ablate-reduce")) {
            run_variant<128, 2, 64, 4, 6, 2, 0>(b, ns, sms);
            run_variant<128, 2, 64, 4, 6, 2, 7>(b, ns, sms);
            run_variant<256, 4, 128, 4, 2, 2, 0>(b, ns, sms);
            run_variant<256, 4, 128, 4, 2, 2, 7>(b, ns, sms);
            continue;
        }
        if (small) {
            run_variant<128, 2, 64, 4, 6, 2>(b, ns, sms);
            run_variant<64, 4, 64, 4, 8, 2>(b, ns, sms);
            run_variant<128, 4, 128, 4, 4, 2>(b, ns, sms);
            run_variant<64, 2, 64, 4, 12, 2>(b, ns, sms);
            run_variant<256, 2, 128, 4, 3, 2>(b, ns, sms);
            run_variant<256, 4, 128, 4, 2, 2>(b, ns, sms);
            run_variant<128, 1, 64, 4, 8, 4>(b, ns, sms);
            continue;
        }
        run_variant<256, 4, 128, 4, 2, 2>(b, ns, sms);
        run_variant<512, 4, 128, 4, 1, 2>(b, ns, sms);
        run_variant<256, 4, 128, 4, 3, 1>(b, ns, sms);
        run_variant<256, 4, 128, 4, 3, 2>(b, ns, sms);
        run_variant<256, 4, 128, 4, 2, 2, 2>(b, ns, sms);
        run_variant<256, 4, 128, 4, 2, 1, 2>(b, ns, sms);
        run_variant<512, 4, 128, 4, 1, 2, 2>(b, ns, sms);
        run_variant<256, 4, 128, 4, 3, 1, 2>(b, ns, sms);
        run_variant<256, 4, 128, 4, 2, 2, 3>(b, ns, sms);
        run_variant<256, 4, 128, 4, 2, 1, 3>(b, ns, sms);
        run_variant<512, 4, 128, 4, 1, 2, 3>(b, ns, sms);
        run_variant<256, 4, 128, 4, 3, 1, 3>(b, ns, sms);
        run_variant<256, 2, 128, 4, 4, 2, 3>(b, ns, sms);
        run_variant<256, 8, 128, 4, 1, 1, 3>(b, ns, sms);
        if (!full) continue;
        run_variant<256, 4, 128, 4, 2, 1>(b, ns, sms);
        run_variant<256, 4, 128, 4, 2, 4>(b, ns, sms);
        run_variant<256, 4, 256, 3, 2, 2>(b, ns, sms);
        run_variant<256, 4, 64, 6, 2, 2>(b, ns, sms);
        run_variant<128, 4, 128, 4, 4, 2>(b, ns, sms);
        run_variant<256, 2, 128, 4, 3, 2>(b, ns, sms);
        run_variant<256, 2, 128, 4, 4, 2>(b, ns, sms);
        run_variant<128, 2, 64, 4, 6, 2>(b, ns, sms);
        run_variant<128, 8, 128, 4, 2, 1>(b, ns, sms);
        run_variant<64, 4, 64, 4, 8, 2>(b, ns, sms);
        run_variant<256, 3, 128, 4, 2, 2>(b, ns, sms);
#endif
    }
}

int main(int argc, char **argv) {
    int dev = 0;
    NB_CUDA(cudaSetDevice(dev));
    cudaDeviceProp prop;
    NB_CUDA(cudaGetDeviceProperties(&prop, dev));
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    const int sms = prop.multiProcessorCount;
    g_peak_flops = (double)sms * 128 * 2 * khz * 1e3;
    g_peak_pairs = (double)sms * 128 * khz * 1e3 / 9.0;
    printf("device %s, %d SMs, max SM clock %d MHz, FP32 peak %.2f TFLOP/s, 9-slot ceiling %.1f Gpairs/s\n", prop.name, sms,
           khz / 1000, g_peak_flops * 1e-12, g_peak_pairs * 1e-9);
    const char *cmd = argc > 1 ? argv[1] : "pipes";
#ifndef TUNE_MINIMAL
    if (!strcmp(cmd, "pipes")) pipes(sms);
    else
#endif
    if (!strcmp(cmd, "sweep")) {
        const uint32_t N = argc > 2 ? (uint32_t)atol(argv[2]) : 262144;
        const uint32_t M = argc > 3 ? (uint32_t)atol(argv[3]) : N / 2;
        sweep(N, M, sms, argc > 4 ? argv[4] : nullptr);
    } else {
        fprintf(stderr, "usage: tune pipes | sweep [N] [M]\n");
        return 2;
    }
    return 0;
}
