/*
 * oracle/ref_glue.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Glue that lets the reference's own CPU path be compiled out of /root/reference
 * without its Vulkan half.  It is linked, together with the reference sources
 * (src/lib/sim_cpu.c, src/lib/world.c, src/lib/galaxy.c -- compiled where they
 * lie, never copied), into oracle/_ref/libnbody_ref*.so by oracle/Makefile.
 *
 * Part 1: the five functions of the reference's GPU seam (src/lib/sim_gpu.h:21-36)
 *         that world.c references (world.c:52,69,78,86,115).  The oracle only ever
 *         runs UpdateWorld_CPU, so the seam is inert; PerformSimUpdate aborts.
 * Part 2: refglue_* helpers (our own code) that drive the reference's exported
 *         PackParticles / PackedUpdate (src/lib/sim_cpu.h:21-24) on a sub-range of
 *         targets, with the same OpenMP schedule as world.c:104-107.  They exist so
 *         that the timed CPU baseline can be a bounded sample of a large workload.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <nbody.h>
#include "sim_cpu.h"
#include "sim_gpu.h"

#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- Part 1: inert GPU seam ------------------------------------------------ */

struct SimPipeline {
    WorldData data;
};

SimPipeline *CreateSimPipeline(WorldData data) {
    SimPipeline *s = malloc(sizeof *s);
    if (s == NULL) abort();
    s->data = data;
    return s;
}

void DestroySimPipeline(SimPipeline *sim) { free(sim); }

void GetSimulationData(const SimPipeline *sim, Particle *ps) { (void)sim; (void)ps; }

void SetSimulationData(SimPipeline *sim, const Particle *ps) { (void)sim; (void)ps; }

void PerformSimUpdate(SimPipeline *sim, uint32_t n, float dt) {
    (void)sim; (void)n; (void)dt;
    fprintf(stderr, "oracle/_ref: the reference Vulkan path is unavailable on this box\n");
    abort();
}

/* ---- Part 2: bounded-sample driver over the reference's PackedUpdate ------- */

/* Which SIMD flavour this build of the reference uses (8 = AVX, 4 = SSE, 1 = scalar). */
int refglue_simd_width(void) {
#if defined(USE_AVX)
    return 8;
#elif defined(USE_SSE)
    return 4;
#else
    return 1;
#endif
}

int refglue_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/*
 * One Jacobi step of the reference CPU path restricted to targets [first, last):
 * snapshot the mass_len sources with the reference's PackParticles, then call the
 * reference's PackedUpdate for every target of the range.  ARR is in World order
 * (massive particles first).  With first = 0, last = total_len this is exactly one
 * iteration of UpdateWorld_CPU (world.c:101-108).
 */
void refglue_step_range(Particle *arr, uint32_t mass_len, uint32_t first, uint32_t last, float dt) {
    ParticlePack *pack = NULL;
    uint32_t pack_len = 0;
    AllocPackArray(&pack, &pack_len, mass_len);
    PackParticles(mass_len, arr, pack);

    #pragma omp parallel for schedule(static, 20)
    for (uint32_t i = first; i < last; i++) {
        PackedUpdate(&arr[i], dt, pack_len, pack);
    }
    FreePackArray(pack);
}
