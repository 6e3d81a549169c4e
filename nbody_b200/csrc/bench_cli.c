/*
 * bench_cli.c -- `nbody-bench`: the benchmark program of the World API, in C.
 *
 * Same protocol as the reference's only measuring tool (ref src/bench.c:21-38,42,53): one
 * srand(11037) at start, MakeGalaxies(N, 2) for an ascending list of sizes, dt = 1, 10 warm-up
 * steps in one Update call, then 100 steps in one timed call, wall clock, per-step microseconds.
 * `--cpu` / `--gpu` select one path like the reference's flags.  Extensions (the reference's README
 * lists "parameters through CLI" as a TODO, ref README.md:98):
 *
 *   --n N[,N...]     sizes instead of the reference's list
 *   --steps K        timed steps per call (default 100);  --warmup W (default 10)
 *   --galaxies G     galaxies per world (default 2)
 *   --ngpus G        shard each World over G GPUs of this box (sets NBODY_NGPUS, which the library reads)
 *   --fresh-seed     re-seed before every size (each size then equals a stand-alone run)
 *   --json           one JSON object per size instead of the table
 *
 * The World API has no parameter for the GPU count, so --ngpus travels through env NBODY_NGPUS.
 * Pair count per step = total_len * mass_len (massless particles are targets, never sources).
 * GB/s = algorithmic HBM bytes of a step (56 B per particle: pos, vel, radius read, pos, vel, acc written, the new
 * source position; 12 B per source read once) over the step time -- the number that matters at small N, where the
 * step is latency-bound and neither roofline is near.
 */
#include <stdbool.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "galaxy.h"
#include "nbody.h"
#include "nbody_sim.h"

/* world.c extras (not part of the reference header) */
SimPipeline *NbWorldPipeline(World *w);
uint32_t NbWorldMassLen(const World *w);
void NbWorldPrepareGPU(World *w);

typedef void (*UpdateFn)(World *, float, uint32_t);

static double seconds_now(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* warm-up call, then one timed call of STEPS steps; returns seconds per step.  The device pipeline is created
 * and the particles uploaded before the clock starts even with --warmup 0 (UpdateWorld_GPU(w, dt, 0) is a
 * no-op that creates nothing, ref world.c:113). */
static double time_path(World *w, UpdateFn update, uint32_t warmup, uint32_t steps, float dt) {
    if (update == UpdateWorld_GPU) NbWorldPrepareGPU(w);
    update(w, dt, warmup);
    const double t0 = seconds_now();
    update(w, dt, steps);
    return (seconds_now() - t0) / (double)steps;
}

static const uint32_t kDefaultSizes[] = {250, 500, 800, 1200, 2000, 4000, 10000, 20000, 50000, 100000};

static uint32_t parse_sizes(const char *arg, uint32_t *out, uint32_t cap) {
    uint32_t n = 0;
    while (*arg && n < cap) {
        char *end = NULL;
        const unsigned long v = strtoul(arg, &end, 10);
        if (end == arg) break;
        out[n++] = (uint32_t)v;
        arg = (*end == ',') ? end + 1 : end;
    }
    return n;
}

int main(int argc, char **argv) {
    bool run_cpu = true, run_gpu = true, json = false, fresh_seed = false;
    uint32_t sizes[64], nsizes = 0, steps = 100, warmup = 10, galaxies = 2;
    const float dt = 1.f; /* ref src/bench.c:21 */

    for (int a = 1; a < argc; a++) {
        if (!strcmp(argv[a], "--cpu")) run_gpu = false;
        else if (!strcmp(argv[a], "--gpu")) run_cpu = false;
        else if (!strcmp(argv[a], "--json")) json = true;
        else if (!strcmp(argv[a], "--fresh-seed")) fresh_seed = true;
        else if (!strcmp(argv[a], "--n") && a + 1 < argc) nsizes = parse_sizes(argv[++a], sizes, 64);
        else if (!strcmp(argv[a], "--steps") && a + 1 < argc) steps = (uint32_t)atoi(argv[++a]);
        else if (!strcmp(argv[a], "--warmup") && a + 1 < argc) warmup = (uint32_t)atoi(argv[++a]);
        else if (!strcmp(argv[a], "--galaxies") && a + 1 < argc) galaxies = (uint32_t)atoi(argv[++a]);
        else if (!strcmp(argv[a], "--ngpus") && a + 1 < argc) setenv("NBODY_NGPUS", argv[++a], 1);
        else {
            fprintf(stderr,
                    "usage: nbody-bench [--cpu|--gpu] [--n N,N,...] [--steps K] [--warmup W] [--galaxies G] "
                    "[--ngpus G] [--fresh-seed] [--json]\n");
            return 2;
        }
    }
    if (nsizes == 0) {
        nsizes = sizeof kDefaultSizes / sizeof kDefaultSizes[0];
        memcpy(sizes, kDefaultSizes, sizeof kDefaultSizes);
    }
    if (steps == 0) steps = 1;

    srand(11037); /* ref src/bench.c:42 */
    if (!json) {
        printf("\t      N\t      M");
        if (run_cpu) printf("\t CPU us\tCPU Gp/s");
        if (run_gpu) printf("\t GPU us\tGPU Gp/s\t%%FP32\t  GB/s");
        printf("\n");
    }
    for (uint32_t k = 0; k < nsizes; k++) {
        const uint32_t n = sizes[k];
        if (fresh_seed) srand(11037);
        Particle *ps = MakeGalaxies(n, galaxies);
        double cpu_s = 0, gpu_s = 0, frac = 0;
        uint32_t m = 0;
        int gpus = 0;
        if (run_cpu) {
            World *w = CreateWorld(ps, n);
            m = NbWorldMassLen(w);
            cpu_s = time_path(w, UpdateWorld_CPU, warmup, steps, dt);
            DestroyWorld(w);
        }
        if (run_gpu) {
            World *w = CreateWorld(ps, n);
            m = NbWorldMassLen(w);
            gpu_s = time_path(w, UpdateWorld_GPU, warmup, steps, dt);
            NbSimInfo info;
            NbSimGetInfo(NbWorldPipeline(w), &info);
            gpus = info.local_gpus;
            const double peak = (double)info.sm_count * 128.0 * 2.0 * (double)info.sm_clock_khz * 1e3 * gpus;
            frac = (double)n * m * 14.0 / gpu_s / peak;
            DestroyWorld(w);
        }
        free(ps);
        const double pairs = (double)n * (double)m;
        const double step_bytes = 56.0 * (double)n + 12.0 * (double)m;
        if (json) {
            printf("{\"n\": %u, \"mass_len\": %u, \"steps\": %u, \"warmup\": %u", n, m, steps, warmup);
            if (run_cpu) printf(", \"cpu_us_per_step\": %.3f, \"cpu_gpairs_per_s\": %.3f", cpu_s * 1e6, pairs / cpu_s * 1e-9);
            if (run_gpu)
                printf(", \"gpu_us_per_step\": %.3f, \"gpu_gpairs_per_s\": %.3f, \"gpus\": %d, \"frac_fp32_peak\": %.4f, "
                       "\"algorithmic_gbytes_per_s\": %.2f",
                       gpu_s * 1e6, pairs / gpu_s * 1e-9, gpus, frac, step_bytes / gpu_s * 1e-9);
            printf("}\n");
        } else {
            printf("\t%7u\t%7u", n, m);
            if (run_cpu) printf("\t%7.0f\t%8.2f", cpu_s * 1e6, pairs / cpu_s * 1e-9);
            if (run_gpu)
                printf("\t%7.1f\t%8.1f\t%5.1f\t%6.1f", gpu_s * 1e6, pairs / gpu_s * 1e-9, 100.0 * frac, step_bytes / gpu_s * 1e-9);
            printf("\n");
        }
        fflush(stdout);
    }
    return 0;
}
