/*
 * layout_kernels.cuh -- AoS <-> SoA conversion on the device and FP64 diagnostics.
 *
 * The API surface speaks 32-byte AoS Particles (include/nbody.h); the hot path wants SoA.
 * These kernels replace the host memcpy of ref src/lib/sim_gpu.c:249-256 (Get/SetSimulationData)
 * plus the reference's per-step AoS->SoA PackParticles (ref src/lib/sim_cpu.c:125-143), which here
 * happens once per upload instead of once per step.
 */
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "force_kernel.cuh"
#include "nbody.h"

namespace nb {

/* AoS -> target arrays (pos, vel, acc, radius), one thread per particle. */
__global__ void unpack_targets_kernel(const Particle *__restrict__ aos, uint32_t total_len, float2 *pos,
                                      float2 *vel, float2 *acc, float *radius) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total_len) return;
    const float4 a = reinterpret_cast<const float4 *>(aos)[2 * (size_t)i];     /* pos, vel      */
    const float4 b = reinterpret_cast<const float4 *>(aos)[2 * (size_t)i + 1]; /* acc, mass, r  */
    pos[i] = make_float2(a.x, a.y);
    vel[i] = make_float2(a.z, a.w);
    acc[i] = make_float2(b.x, b.y);
    radius[i] = b.w;
}

/*
 * AoS -> source packs, one thread per pack of 4 sources.  G*m is folded here once
 * (the reference multiplies per pair, ref sim_cpu.c:179 -- same single rounding).  Slots past
 * mass_len are padding: far away and massless, so they add exactly zero.  Both ping-pong
 * position buffers are initialised because the epilogue only ever rewrites slots < mass_len.
 * Packs [pack_first, pack_last) only: a GPU that received just its own particles converts just those.
 */
__global__ void unpack_sources_kernel(const Particle *__restrict__ aos, uint32_t mass_len, uint32_t pack_first,
                                      uint32_t pack_last, float4 *xy0, float4 *xy1, float4 *gm) {
    const uint32_t pk = pack_first + blockIdx.x * blockDim.x + threadIdx.x;
    if (pk >= pack_last) return;
    float x[4], y[4], g[4];
#pragma unroll
    for (int l = 0; l < 4; l++) {
        const uint32_t j = pk * 4 + l;
        if (j < mass_len) {
            const float4 a = reinterpret_cast<const float4 *>(aos)[2 * (size_t)j];
            const float4 b = reinterpret_cast<const float4 *>(aos)[2 * (size_t)j + 1];
            x[l] = a.x;
            y[l] = a.y;
            g[l] = __fmul_rn(b.z, NB_G);
        } else {
            x[l] = kPadPos;
            y[l] = kPadPos;
            g[l] = 0.f;
        }
    }
    const float4 xs = make_float4(x[0], x[1], x[2], x[3]);
    const float4 ys = make_float4(y[0], y[1], y[2], y[3]);
    xy0[2 * (size_t)pk] = xs;
    xy0[2 * (size_t)pk + 1] = ys;
    xy1[2 * (size_t)pk] = xs;
    xy1[2 * (size_t)pk + 1] = ys;
    gm[pk] = make_float4(g[0], g[1], g[2], g[3]);
}

/* Target arrays -> AoS for [first, last): pos, vel, acc are patched, mass and radius were never touched. */
__global__ void pack_targets_kernel(Particle *aos, uint32_t first, uint32_t last, const float2 *__restrict__ pos,
                                    const float2 *__restrict__ vel, const float2 *__restrict__ acc) {
    const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last) return;
    const float2 p = pos[i], v = vel[i], a = acc[i];
    float4 *out = reinterpret_cast<float4 *>(aos) + 2 * (size_t)i;
    out[0] = make_float4(p.x, p.y, v.x, v.y);
    float4 b = out[1];
    b.x = a.x;
    b.y = a.y;
    out[1] = b;
}

/* FP64 accelerations of targets [first, last): the "truth" column of the parity reports. */
__global__ void acc_f64_kernel(const float2 *__restrict__ pos, const float *__restrict__ radius,
                               const float4 *__restrict__ src_xy, const float4 *__restrict__ src_gm,
                               uint32_t mass_len, uint32_t first, uint32_t last, double *out) {
    const uint32_t i = first + blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= last) return;
    const double x = pos[i].x, y = pos[i].y, r = radius[i];
    double ax = 0.0, ay = 0.0;
    const float *sxy = reinterpret_cast<const float *>(src_xy);
    const float *sgm = reinterpret_cast<const float *>(src_gm);
    for (uint32_t j = 0; j < mass_len; j++) {
        const uint32_t o = (j >> 2) * 8u + (j & 3u);
        const double dx = (double)sxy[o] - x;
        const double dy = (double)sxy[o + 4] - y;
        const double r2 = dx * dx + dy * dy + r;
        const double f = (double)sgm[j] / (sqrt(r2) * r2);
        ax += dx * f;
        ay += dy * f;
    }
    out[2 * (size_t)(i - first)] = ax;
    out[2 * (size_t)(i - first) + 1] = ay;
}

/*
 * FP64 diagnostics (drift harness; same definitions as oracle_diagnostics_f64):
 * out[0..1] sum m v, out[2] K, out[3] U over massive pairs with target softening, out[4] sum m,
 * out[5..6] sum m x.  One thread per particle, block reduction, one atomicAdd(double) per block.
 */
__global__ void diagnostics_kernel(const Particle *__restrict__ aos, uint32_t total_len, uint32_t mass_len,
                                   double *out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    double v[7] = {0, 0, 0, 0, 0, 0, 0};
    if (i < total_len) {
        const Particle p = aos[i];
        const double m = p.mass;
        v[0] = m * p.vel.x;
        v[1] = m * p.vel.y;
        v[2] = 0.5 * m * ((double)p.vel.x * p.vel.x + (double)p.vel.y * p.vel.y);
        v[4] = m;
        v[5] = m * p.pos.x;
        v[6] = m * p.pos.y;
        if (i < mass_len) {
            double ui = 0.0;
            for (uint32_t j = 0; j < mass_len; j++) {
                if (j == i) continue;
                const double dx = (double)aos[j].pos.x - p.pos.x;
                const double dy = (double)aos[j].pos.y - p.pos.y;
                ui += (double)aos[j].mass / sqrt(dx * dx + dy * dy + (double)p.radius);
            }
            v[3] = -0.5 * (double)NB_G * m * ui;
        }
    }
    __shared__ double red[7][8];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int k = 0; k < 7; k++) {
        double s = v[k];
        for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (lane == 0) red[k][warp] = s;
    }
    __syncthreads();
    if (threadIdx.x < 7) {
        double s = 0.0;
        for (int w = 0; w < (int)(blockDim.x >> 5); w++) s += red[threadIdx.x][w];
        atomicAdd(&out[threadIdx.x], s);
    }
}

}  // namespace nb
