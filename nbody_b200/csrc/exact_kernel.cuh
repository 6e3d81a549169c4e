/*
 * exact_kernel.cuh -- the reference's CPU arithmetic, operation for operation, on the GPU.
 *
 * The fast kernels (step_kernel.cuh, small_kernel.cuh) compute the same physics with one rsqrt.approx per pair,
 * fused multiply-adds and a hierarchical summation order; they are closer to FP64 than the reference is, but they
 * cannot agree with it to the last bit, and where the reference's own FP32 rounding noise exceeds 1e-5 no other
 * summation order can stay within 1e-5 of it either.  This kernel is the other trade: it
 * reproduces the reference's PackedUpdate (ref src/lib/sim_cpu.c:156-194) EXACTLY --
 *
 *   dx = xj - xi; dy = yj - yi; dist_sq = dx*dx + dy*dy; r2 = dist_sq + radius_i; r1 = sqrt(r2);
 *   gm = mj * G; r3 = r1 * r2; f = gm / r3; ax = ax + dx*f; ay = ay + dy*f
 *
 * every operation separately rounded to nearest (IEEE sqrt.rn / div.rn, no FMA, denormals kept), one accumulator
 * per SIMD lane of the reference build (LANES = 8 AVX, 4 SSE, 1 scalar), source LANES*k + s in lane LANES-1-s
 * because _mm256_set_ps takes its arguments high lane first (ref sim_cpu.c:32-33), the tail pack padded with
 * all-zero particles (ref sim_cpu.c:133-142), lanes added left to right from lane 0 starting at +0
 * (simd_sum, ref sim_cpu.c:146-154), then vel += acc*dt, pos += vel*dt.  The result of UpdateWorld_GPU is then
 * bit-identical to the reference's UpdateWorld_CPU of the same SIMD build (tests/test_gpu_exact.py: golden fixtures
 * produced by the reference itself, 1 and 10 steps; the oracle at larger N; ragged sizes).
 *
 * It is a verification mode (NBODY_EXACT=8|4|1 or NbSimConfig.flags NB_SIM_FLAG_EXACT), not the product's hot path:
 * IEEE sqrt and division cost ~20 instructions per pair where the fast kernel spends one MUFU, so it runs at
 * roughly a quarter of the fast kernel's rate -- still well above the 16-thread CPU it reproduces.
 * One difference remains by construction: 0/0 (a target at the origin with radius 0 against a padding lane)
 * yields a NaN here and on the CPU, but not necessarily the same NaN bit pattern.
 *
 * Layout: one thread per target, 256 targets per CTA, all sources streamed through shared memory in chunks of
 * 512 (x, y, G*m as three float arrays; G*m was rounded once at upload -- the same single rounding as the
 * reference's per-pair simd_mul(pack.m, g)).  Targets read their own old position, sources come from the
 * step's source snapshot: Jacobi, like the reference's PackParticles snapshot (ref world.c:102).
 */
#pragma once

#include "force_kernel.cuh"

namespace nb {

constexpr int kExactThreads = 256;
constexpr int kExactChunk = 512; /* sources per shared-memory chunk: a multiple of 8 (whole reference packs) and of 4 */

template <int LANES>
__global__ void __launch_bounds__(kExactThreads) exact_kernel(const __grid_constant__ StepParams p) {
    static_assert(LANES == 8 || LANES == 4 || LANES == 1, "the reference builds with 8 (AVX), 4 (SSE) or 1 lane");
    __shared__ float s_x[kExactChunk], s_y[kExactChunk], s_g[kExactChunk];

    const uint32_t tid = threadIdx.x;
    const uint32_t i = (p.tile_first + blockIdx.x) * kExactThreads + tid;
    const bool live = i < p.total_len;

    asm volatile("griddepcontrol.wait;" ::: "memory");

    float x = 0.f, y = 0.f, r = 1.f;
    if (live) {
        const float2 q = p.pos[i];
        x = q.x;
        y = q.y;
        r = p.radius[i];
    }
    float ax[LANES], ay[LANES];
#pragma unroll
    for (int l = 0; l < LANES; l++) ax[l] = ay[l] = 0.f;

    /* the reference walks ceil(M / LANES) packs; lanes past mass_len hold all-zero particles */
    const uint32_t padded = (p.mass_len + LANES - 1) / LANES * LANES;
    for (uint32_t base = 0; base < padded; base += kExactChunk) {
        const uint32_t cnt = min((uint32_t)kExactChunk, padded - base);
        __syncthreads();
        for (uint32_t k = tid; k < cnt; k += kExactThreads) {
            const uint32_t j = base + k;
            float sx = 0.f, sy = 0.f, sg = 0.f;
            if (j < p.mass_len) { /* device layout: packs of 4 sources, xs then ys (layout_kernels.cuh) */
                const float *xy = reinterpret_cast<const float *>(p.src_xy) + (size_t)(j >> 2) * 8 + (j & 3u);
                sx = xy[0];
                sy = xy[4];
                sg = reinterpret_cast<const float *>(p.src_gm)[j];
            }
            s_x[k] = sx;
            s_y[k] = sy;
            s_g[k] = sg;
        }
        __syncthreads();
#pragma unroll 1
        for (uint32_t k = 0; k < cnt; k += LANES) {
#pragma unroll
            for (int s = 0; s < LANES; s++) {
                const int l = LANES - 1 - s;
                const float dx = __fsub_rn(s_x[k + s], x);
                const float dy = __fsub_rn(s_y[k + s], y);
                const float dist_sq = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
                const float r2 = __fadd_rn(dist_sq, r);
                const float r1 = __fsqrt_rn(r2);
                const float r3 = __fmul_rn(r1, r2);
                const float f = __fdiv_rn(s_g[k + s], r3);
                ax[l] = __fadd_rn(ax[l], __fmul_rn(dx, f));
                ay[l] = __fadd_rn(ay[l], __fmul_rn(dy, f));
            }
        }
    }
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    float sax = 0.f, say = 0.f;
#pragma unroll
    for (int l = 0; l < LANES; l++) sax = __fadd_rn(sax, ax[l]);
#pragma unroll
    for (int l = 0; l < LANES; l++) say = __fadd_rn(say, ay[l]);
    if (live) integrate_store(p, i, make_float2(sax, say));
}

}  // namespace nb
