/*
 * galaxy.c -- MakeGalaxies of include/galaxy.h: synthetic spiral-galaxy initial conditions.
 *
 * Every benchmark input of this project is defined as "srand(seed); MakeGalaxies(n, g)" of the
 * reference (ref src/lib/galaxy.c:31-221).  This is our own implementation of that generator; it
 * is organised differently (one pass per concern over a small Galaxy record) but consumes libc
 * rand() in exactly the reference's order and evaluates every float expression with the same
 * operand order and precision, so with the same libc and seed the particles are bit-identical
 * (tests/test_host.py::test_make_galaxies_bit_identical_* check this against the compiled reference and the
 * golden fixtures).
 * Compiled with -ffp-contract=off so no FMA sneaks into the float expressions.
 *
 * Draw order per call, one line per rand() (n = particles, g = galaxies):
 *   sizes      g-1 draws   size of each galaxy but the last            (ref :46-54)
 *   cores      g   draws   core radius                                  (ref :69)
 *   placement  3 per try   parent index, squared distance, angle       (ref :89-98)
 *   spirals    2 per gal.  rotation offset, spiral count               (ref :157-160)
 *   particle   7 or 8      angle t, t jitter, r jitter, r sign, t sign, spiral pick, massless
 *                          test, [radius]                              (ref :188-213)
 */
#include <math.h>
#include <stdbool.h>
#include <stdio.h>
#include <stdlib.h>

#include "galaxy.h"

typedef struct Galaxy {
    Particle *first;   /* the core; ordinary particles follow it */
    uint32_t count;    /* particles including the core           */
    float inner, outer; /* min_dist, max_dist of include/galaxy.h */
} Galaxy;

/* uniform float in [lo, hi): double arithmetic, then one rounding (ref galaxy.c:18-20) */
static float draw_float(double lo, double hi) {
    return (float)(lo + (hi - lo) * rand() / RAND_MAX);
}

/* uniform integer in [lo, hi) (ref galaxy.c:23-25) */
static uint32_t draw_uint(uint32_t lo, uint32_t hi) {
    return lo + ((uint32_t)rand() % (hi - lo));
}

static float draw_sign(void) { /* -1 when the low bit of rand() is set (ref galaxy.c:27-29,195-196) */
    return (rand() & 1) ? -1.f : 1.f;
}

static void place_galaxy(Galaxy *all, uint32_t k) {
    Galaxy *self = &all[k];
    for (;;) {
        const uint32_t parent_idx = draw_uint(0, k);
        const Galaxy *parent = &all[parent_idx];
        const float reach = self->outer + parent->outer;
        const float near = MIN_GALAXY_SEPARATION * reach;
        const float far = MAX_GALAXY_SEPARATION * reach;
        const float dist = sqrtf(draw_float(near * near, far * far));
        const float angle = draw_float(0, 2 * PI);
        self->first->pos.x = parent->first->pos.x + dist * cosf(angle);
        self->first->pos.y = parent->first->pos.y + dist * sinf(angle);

        bool clear = true;
        for (uint32_t o = 0; o < k && clear; o++) {
            if (o == parent_idx) continue;
            const float keep = MIN_GALAXY_SEPARATION * (self->outer + all[o].outer);
            const float d2 = SqMagV2(SubV2(self->first->pos, all[o].first->pos));
            if (d2 < keep * keep) clear = false;
        }
        if (clear) return;
    }
}

/* Sideways kick between two cores so that galaxies do not fall straight into each other. */
static void kick_pair(Particle *a, Particle *b) {
    const V2 ab = SubV2(b->pos, a->pos);
    const float dist = MagV2(ab);
    const V2 unit = ScaleV2(ab, 1.f / dist);
    const float speed_a = 0.3f * sqrtf(NB_G * b->mass / dist);
    const float speed_b = 0.3f * sqrtf(NB_G * a->mass / dist);
    a->vel = AddV2(a->vel, ScaleV2(V2_FROM(unit.y, -unit.x), speed_a));
    b->vel = AddV2(b->vel, ScaleV2(V2_FROM(-unit.y, unit.x), speed_b));
}

static void fill_galaxy(const Galaxy *gal) {
    const Particle core = *gal->first;
    const float span = gal->outer - gal->inner;

    /* arms: r(t) = b*t reaches `outer` at t = 2*pi and `inner` at t0 */
    float arm_phase[MAX_SPIRALS];
    const float phase0 = draw_float(0, 2 * PI);
    const uint32_t arms = draw_uint(MIN_SPIRALS, 1 + MAX_SPIRALS);
    const float arm_step = 2 * PI / (float)arms;
    for (uint32_t a = 0; a < arms; a++) arm_phase[a] = phase0 + (float)a * arm_step;

    const float t_end = 2 * PI;
    const float b = gal->outer / t_end;
    const float t_begin = gal->inner / b;

    for (uint32_t k = 1; k < gal->count; k++) {
        Particle *p = &gal->first[k];
        *p = (Particle){0};

        const float t = draw_float(t_begin, t_end);
        const float r = b * t;
        /* squared jitter keeps most particles close to the arm */
        const float jt = draw_float(0, 0.6f * sqrtf(arm_step));
        const float jr = draw_float(0, 0.6f * sqrtf(fminf(b, r - gal->inner)));
        const float dist = r + draw_sign() * (jr * jr);
        const float ang = t + draw_sign() * (jt * jt);

        const float phase = arm_phase[draw_uint(0, arms)];
        const float dx = dist * cosf(ang + phase);
        const float dy = dist * sinf(ang + phase);
        p->pos.x = core.pos.x + dx;
        p->pos.y = core.pos.y + dy;

        /* the outskirts are mostly massless tracer particles */
        if (draw_float(0.f, 1.f) < (dist - gal->inner) / span) {
            p->radius = 0.5f;
            p->mass = 0.f;
        } else {
            p->radius = draw_float(NP_MIN_R, NP_MAX_R);
            p->mass = NP_R_TO_M(p->radius);
        }

        /* circular orbit around the core, riding on the core's own velocity */
        const float speed = sqrtf(NB_G * core.mass / dist);
        p->vel.x = core.vel.x + speed * (dy / dist);
        p->vel.y = core.vel.y + speed * (-dx / dist);
    }
}

Particle *MakeGalaxies(uint32_t particle_count, uint32_t galaxy_count) {
    if (particle_count < galaxy_count * MIN_PARTICLES_PER_GALAXY) {
        fprintf(stderr, "%s:%d [%s] Need at least %u particles to make %u galaxies, called with %u\n", __FILE__,
                __LINE__, __func__, galaxy_count * MIN_PARTICLES_PER_GALAXY, galaxy_count, particle_count);
        abort();
    }
    Particle *all = malloc((size_t)particle_count * sizeof(Particle));
    Galaxy *gal = malloc((size_t)(galaxy_count ? galaxy_count : 1) * sizeof(Galaxy));
    if (all == NULL || gal == NULL) {
        fprintf(stderr, "%s:%d [%s] Failed to alloc %u particles\n", __FILE__, __LINE__, __func__, particle_count);
        abort();
    }

    /* sizes: everyone gets the minimum, the surplus is handed out at random, the last takes the rest */
    uint32_t surplus = particle_count - galaxy_count * MIN_PARTICLES_PER_GALAXY;
    uint32_t offset = 0;
    for (uint32_t k = 0; k < galaxy_count; k++) {
        uint32_t extra = surplus;
        if (k + 1 < galaxy_count) {
            extra = draw_uint(0, 1 + surplus);
            surplus -= extra;
        }
        gal[k].first = all + offset;
        gal[k].count = MIN_PARTICLES_PER_GALAXY + extra;
        offset += gal[k].count;
    }

    /* cores and extents */
    for (uint32_t k = 0; k < galaxy_count; k++) {
        const float core_r = draw_float(GC_MIN_R, GC_MAX_R);
        const float root = sqrtf((float)gal[k].count);
        gal[k].inner = core_r * MIN_PARTICLE_DIST_CR_F;
        gal[k].outer = core_r * MAX_PARTICLE_DIST_CR_F + root * MAX_PARTICLE_DIST_PC_F;
        *gal[k].first = (Particle){.radius = core_r, .mass = GC_R_TO_M(core_r)};
    }

    /* galaxy 0 rests at the origin; the others are placed relative to an earlier one */
    for (uint32_t k = 1; k < galaxy_count; k++) place_galaxy(gal, k);

    for (uint32_t k = 1; k < galaxy_count; k++)
        for (uint32_t o = 0; o < k; o++) kick_pair(gal[k].first, gal[o].first);

    for (uint32_t k = 0; k < galaxy_count; k++) fill_galaxy(&gal[k]);

    free(gal);
    return all;
}
