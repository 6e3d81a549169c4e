/*
 * world.c -- the World API of include/nbody.h on top of the CUDA pipeline (include/nbody_sim.h).
 *
 * Host code stays C, like the reference's src/lib/world.c; it reaches CUDA only through the
 * C-ABI seam (CreateSimPipeline / SetSimulationData / GetSimulationData / PerformSimUpdate).
 * Semantics kept from the reference, with the lines they come from:
 *   - CreateWorld copies the caller's particles and partitions massive ones to the front with the
 *     same unstable two-cursor sweep, so the order seen through GetWorldParticles is identical
 *     (ref world.c:22-46; permutation pinned by ref test/test_particle_sort.c);
 *   - two dirty flags decide lazily which copy (host AoS / device SoA) is current
 *     (ref world.c:59-60,76-89);
 *   - UpdateWorld_GPU(n == 0) does nothing; UpdateWorld_CPU(n == 0) still marks the host copy
 *     as the newer one (ref world.c:99-118).
 * Differences: the device pipeline is created on the first UpdateWorld_GPU (the reference creates
 * its Vulkan pipeline eagerly, ref world.c:52) so that host-only use needs no GPU; DestroyWorld
 * also frees the particle array (the reference leaks it, ref world.c:67-73).
 */
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nbody.h"
#include "nbody_sim.h"

#define WORLD_FAIL(...)                                               \
    do {                                                              \
        fprintf(stderr, "%s:%d [%s] ", __FILE__, __LINE__, __func__); \
        fprintf(stderr, __VA_ARGS__);                                 \
        fprintf(stderr, "\n");                                        \
        abort();                                                      \
    } while (0)

struct World {
    Particle *arr;        /* host AoS copy, World order (massive first)              */
    bool arr_pinned;      /* arr came from NbSimHostAlloc                            */
    SimPipeline *sim;     /* device pipeline, created on first GPU update            */
    uint32_t total_len;
    uint32_t mass_len;
    bool device_stale;    /* host copy changed since the last upload                 */
    bool host_stale;      /* device copy changed since the last download             */
    /* scratch of UpdateWorld_CPU: snapshot of the sources, padded to 8 */
    float *sx, *sy, *sgm;
    uint32_t snap_len;
};

/* Move every particle with mass > 0 in front of the others; returns how many there are.
 * Same swaps, in the same order, as ref world.c:32-46. */
static uint32_t partition_by_mass(Particle *arr, uint32_t size) {
    uint32_t head = 0, tail = size;
    for (;;) {
        while (head < tail && arr[head].mass > 0) head++; /* first massless from the left  */
        while (head < tail) {                             /* last massive from the right    */
            tail--;
            if (!(arr[tail].mass <= 0)) break;
        }
        if (head == tail) return tail;
        const Particle moved = arr[head];
        arr[head] = arr[tail];
        arr[tail] = moved;
    }
}

World *CreateWorld(const Particle *ps, uint32_t size) {
    World *w = calloc(1, sizeof *w);
    if (w == NULL) WORLD_FAIL("Failed to alloc World");

    const size_t bytes = (size_t)size * sizeof(Particle);
    if (NbSimDeviceCount() > 0) {
        w->arr = NbSimHostAlloc(bytes); /* page-locked: uploads/downloads at link speed */
        w->arr_pinned = true;
    } else {
        w->arr = malloc(bytes ? bytes : 1);
    }
    if (w->arr == NULL) WORLD_FAIL("Failed to alloc %u particles", size);
    if (size) memcpy(w->arr, ps, bytes);

    w->total_len = size;
    w->mass_len = partition_by_mass(w->arr, size);
    w->device_stale = true; /* device has nothing yet              (ref world.c:59) */
    w->host_stale = false;  /* host copy is the only, current one  (ref world.c:60) */
    return w;
}

void DestroyWorld(World *w) {
    if (w == NULL) return;
    DestroySimPipeline(w->sim);
    if (w->arr_pinned) NbSimHostFree(w->arr);
    else free(w->arr);
    free(w->sx);
    free(w->sy);
    free(w->sgm);
    free(w);
}

static void pull_from_device(World *w) {
    if (w->host_stale) {
        GetSimulationData(w->sim, w->arr);
        w->host_stale = false;
    }
}

static void push_to_device(World *w) {
    if (w->sim == NULL) {
        const WorldData data = {.total_len = w->total_len, .mass_len = w->mass_len, .dt = 0.f};
        w->sim = CreateSimPipeline(data); /* aborts when there is no usable GPU: no CPU fallback */
    }
    if (w->device_stale) {
        SetSimulationData(w->sim, w->arr);
        w->device_stale = false;
    }
}

const Particle *GetWorldParticles(World *w, uint32_t *size) {
    pull_from_device(w);
    if (size != NULL) *size = w->total_len;
    return w->arr;
}

void UpdateWorld_GPU(World *w, float dt, uint32_t n) {
    if (n == 0) return;
    if (w->total_len == 0) return;
    push_to_device(w);
    PerformSimUpdate(w->sim, n, dt);
    w->host_stale = true;
}

/* Exposed for tools (nbody-bench prints what ran); not part of the reference header. */
SimPipeline *NbWorldPipeline(World *w) { return w->sim; }
/* Create the device pipeline and upload the particles now instead of on the first UpdateWorld_GPU, without
 * stepping: lets a benchmark keep context creation and the first upload out of its timed call. */
void NbWorldPrepareGPU(World *w) {
    if (w->total_len == 0) return;
    push_to_device(w);
}
uint32_t NbWorldMassLen(const World *w) { return w->mass_len; }

/* ---- UpdateWorld_CPU: a host implementation of the same step ------------------------------
 *
 * Present because the header declares it and programs alternate between the two paths on one
 * World (ref src/main.c:112-116,158-162).  It is never used by UpdateWorld_GPU.  The arithmetic is
 * written so that it rounds like the reference's default AVX build: sources in groups of 8, one
 * accumulator per lane, every operation separately rounded (this file is compiled with
 * -ffp-contract=off), lanes added in the reference's order (ref sim_cpu.c:146-154 with the
 * high-lane-first packing of ref sim_cpu.c:32-33), tail group padded with zero particles
 * (ref sim_cpu.c:133-142).
 */
#define CPU_LANES 8

static void cpu_snapshot(World *w) {
    const uint32_t padded = (w->mass_len + CPU_LANES - 1) / CPU_LANES * CPU_LANES;
    if (w->snap_len != padded || w->sx == NULL) {
        free(w->sx);
        free(w->sy);
        free(w->sgm);
        /* C11 aligned_alloc wants a size that is a multiple of the alignment */
        const size_t bytes = ((size_t)(padded ? padded : CPU_LANES) * sizeof(float) + 63) / 64 * 64;
        w->sx = aligned_alloc(64, bytes);
        w->sy = aligned_alloc(64, bytes);
        w->sgm = aligned_alloc(64, bytes);
        if (!w->sx || !w->sy || !w->sgm) WORLD_FAIL("Failed to alloc source snapshot");
        w->snap_len = padded;
    }
    #pragma omp parallel for schedule(static)
    for (uint32_t j = 0; j < padded; j++) {
        const bool real = j < w->mass_len;
        w->sx[j] = real ? w->arr[j].pos.x : 0.f;
        w->sy[j] = real ? w->arr[j].pos.y : 0.f;
        w->sgm[j] = real ? w->arr[j].mass * NB_G : 0.f;
    }
}

typedef float Lanes __attribute__((vector_size(CPU_LANES * sizeof(float)), aligned(32)));

static void cpu_update_one(Particle *p, const float *sx, const float *sy, const float *sgm, uint32_t padded,
                           float dt) {
    /* one 8-lane accumulator per axis, kept in registers (GCC vector extension; with -mavx these are
     * single ymm operations, every one separately rounded: -ffp-contract=off) */
    Lanes ax = {0}, ay = {0};
    const float x = p->pos.x, y = p->pos.y, soft = p->radius;
    const Lanes x8 = {x, x, x, x, x, x, x, x}, y8 = {y, y, y, y, y, y, y, y};
    const Lanes soft8 = {soft, soft, soft, soft, soft, soft, soft, soft};
    for (uint32_t base = 0; base < padded; base += CPU_LANES) {
        const Lanes dx = *(const Lanes *)(sx + base) - x8;
        const Lanes dy = *(const Lanes *)(sy + base) - y8;
        const Lanes r2 = (dx * dx + dy * dy) + soft8;
        Lanes r1;
        for (int s = 0; s < CPU_LANES; s++) r1[s] = sqrtf(r2[s]); /* vsqrtps under -fno-math-errno */
        const Lanes f = *(const Lanes *)(sgm + base) / (r1 * r2);
        ax += dx * f;
        ay += dy * f;
    }
    float sum_x = 0.f, sum_y = 0.f;
    for (int s = CPU_LANES - 1; s >= 0; s--) sum_x += ax[s];
    for (int s = CPU_LANES - 1; s >= 0; s--) sum_y += ay[s];
    p->acc = V2_FROM(sum_x, sum_y);
    p->vel = AddV2(p->vel, ScaleV2(p->acc, dt));
    p->pos = AddV2(p->pos, ScaleV2(p->vel, dt));
}

void UpdateWorld_CPU(World *w, float dt, uint32_t n) {
    pull_from_device(w);
    for (uint32_t it = 0; it < n; it++) {
        cpu_snapshot(w);
        const uint32_t padded = w->snap_len;
        const float *sx = w->sx, *sy = w->sy, *sgm = w->sgm;
        Particle *arr = w->arr;
        #pragma omp parallel for schedule(static, 32)
        for (uint32_t i = 0; i < w->total_len; i++) {
            cpu_update_one(&arr[i], sx, sy, sgm, padded, dt);
        }
    }
    w->device_stale = true;
}
