/*
 * render_kernels.cuh -- headless snapshot of the particle field: a CUDA point splat into a device
 * RGBA buffer, standing in for the reference GUI's DrawParticles (ref src/main.c:237-250) and
 * CreateCamera (ref src/main.c:198-225).  No window, no graphics API: plain compute.
 *
 * Semantics kept from the reference: every particle is a filled disc centred on its position
 * truncated to an integer (ref main.c:244-245), of radius max(radius, min_radius) (ref :246), in a
 * colour chosen by mass class (ref main.c:227-235, colours :20-23); discs are painted in array
 * order, so where discs overlap the particle with the HIGHER index wins.  The camera maps world to
 * screen as  screen = (world - target) * zoom + (width/2, height/2).
 *
 * Parallel painter's algorithm: pass 1 writes, per pixel, the highest particle index whose disc
 * covers the pixel centre (atomicMax on index+1); pass 2 turns the winner's mass class into a colour.
 * Order-independent and deterministic.  One warp per particle; lanes stride over the disc's bounding
 * box, so the few very large discs (galaxy cores) do not serialise on one thread.
 * All coordinate arithmetic is explicitly rounded FP32 (no FMA contraction) so that the CPU
 * restatement in tests/test_render.py reproduces the image bit for bit.
 */
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <string.h>

#include "galaxy.h"
#include "nbody.h"

namespace nb {

struct Camera {
    float target_x, target_y, zoom;
    int32_t width, height;
};

/* FP32 min/max of the positions: 4 values (min x, min y, max x, max y), one atomic per warp.
 * Ordered-int trick: non-negative floats compare like their bit patterns, negative ones reversed. */
__device__ __forceinline__ int float_to_ordered(float f) {
    const int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float ordered_to_float(int i) { return __int_as_float(i >= 0 ? i : i ^ 0x7fffffff); }

static inline int float_to_ordered_host(float f) {
    int i;
    memcpy(&i, &f, 4);
    return i >= 0 ? i : i ^ 0x7fffffff;
}
static inline float ordered_to_float_host(int i) {
    i = i >= 0 ? i : i ^ 0x7fffffff;
    float f;
    memcpy(&f, &i, 4);
    return f;
}

__global__ void bounds_kernel(const Particle *__restrict__ aos, uint32_t n, int *out4) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float x0 = INFINITY, y0 = INFINITY, x1 = -INFINITY, y1 = -INFINITY;
    if (i < n) {
        const float4 a = reinterpret_cast<const float4 *>(aos)[2 * (size_t)i];
        x0 = x1 = a.x;
        y0 = y1 = a.y;
    }
    for (int o = 16; o > 0; o >>= 1) {
        x0 = fminf(x0, __shfl_xor_sync(0xffffffffu, x0, o));
        y0 = fminf(y0, __shfl_xor_sync(0xffffffffu, y0, o));
        x1 = fmaxf(x1, __shfl_xor_sync(0xffffffffu, x1, o));
        y1 = fmaxf(y1, __shfl_xor_sync(0xffffffffu, y1, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(&out4[0], float_to_ordered(x0));
        atomicMin(&out4[1], float_to_ordered(y0));
        atomicMax(&out4[2], float_to_ordered(x1));
        atomicMax(&out4[3], float_to_ordered(y1));
    }
}

/* pass 1: winner[pixel] = max over covering particles of (index + 1) */
__global__ void splat_kernel(const Particle *__restrict__ aos, uint32_t n, Camera cam, float min_radius,
                             unsigned *winner) {
    const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t lane = threadIdx.x & 31;
    if (i >= n) return;
    const float4 a = reinterpret_cast<const float4 *>(aos)[2 * (size_t)i];
    const float4 b = reinterpret_cast<const float4 *>(aos)[2 * (size_t)i + 1];
    /* (int)p.pos.x like the reference, then the camera transform */
    const float wx = truncf(a.x), wy = truncf(a.y);
    const float sx = __fadd_rn(__fmul_rn(__fsub_rn(wx, cam.target_x), cam.zoom), 0.5f * (float)cam.width);
    const float sy = __fadd_rn(__fmul_rn(__fsub_rn(wy, cam.target_y), cam.zoom), 0.5f * (float)cam.height);
    const float r = __fmul_rn(fmaxf(b.w, min_radius), cam.zoom);
    if (!(r >= 0.f) || !isfinite(sx) || !isfinite(sy)) return;
    const float r2 = __fmul_rn(r, r);
    /* bounding box in pixels, clipped to the image */
    const float fx0 = floorf(sx - r - 1.f), fx1 = ceilf(sx + r + 1.f);
    const float fy0 = floorf(sy - r - 1.f), fy1 = ceilf(sy + r + 1.f);
    if (fx1 < 0.f || fy1 < 0.f || fx0 >= (float)cam.width || fy0 >= (float)cam.height) return;
    const int x0 = (int)fmaxf(fx0, 0.f), x1 = (int)fminf(fx1, (float)(cam.width - 1));
    const int y0 = (int)fmaxf(fy0, 0.f), y1 = (int)fminf(fy1, (float)(cam.height - 1));
    const int bw = x1 - x0 + 1, bh = y1 - y0 + 1;
    if (bw <= 0 || bh <= 0) return;
    const long long cells = (long long)bw * bh;
    for (long long c = lane; c < cells; c += 32) {
        const int px = x0 + (int)(c % bw), py = y0 + (int)(c / bw);
        const float ddx = __fsub_rn((float)px + 0.5f, sx), ddy = __fsub_rn((float)py + 0.5f, sy);
        const float d2 = __fadd_rn(__fmul_rn(ddx, ddx), __fmul_rn(ddy, ddy));
        if (d2 <= r2) atomicMax(&winner[(size_t)py * cam.width + px], i + 1u);
    }
}

/* pass 2: colour of the winner's mass class (ref main.c:20-23,227-235); RGBA8, alpha 255 */
__global__ void resolve_kernel(const Particle *__restrict__ aos, const unsigned *__restrict__ winner, uint32_t pixels,
                               uchar4 *rgba) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pixels) return;
    const unsigned w = winner[p];
    uchar4 c = make_uchar4(22, 22, 22, 255); /* background */
    if (w) {
        const float mass = aos[w - 1].mass;
        if (mass <= 0.f) c = make_uchar4(145, 145, 233, 255);            /* massless tracer  */
        else if (mass < MIN_GC_MASS) c = make_uchar4(175, 195, 175, 255); /* ordinary particle */
        else c = make_uchar4(222, 222, 222, 255);                         /* galaxy core       */
    }
    rgba[p] = c;
}

}  // namespace nb
