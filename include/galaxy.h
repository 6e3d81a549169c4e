/*
 * galaxy.h -- physical constants of the synthetic spiral-galaxy scenario and the
 * generator entry point.  Same macro names and values as the reference's
 * include/galaxy.h (lines cited), because every benchmark input is defined by them.
 */
#ifndef NB_GALAXY_H
#define NB_GALAXY_H

#include "nbody.h"

#ifdef __cplusplus
extern "C" {
#endif

#ifndef PI
#define PI 3.1415927f /* ref include/galaxy.h:7 */
#endif

/* spirals per galaxy (ref :10-11) */
#define MIN_SPIRALS 2
#define MAX_SPIRALS 4

/* galaxy cores: radius range and density (ref :13-15) */
#define GC_MIN_R   200.f
#define GC_MAX_R   600.f
#define GC_DENSITY 30.0f

/* ordinary particles: radius range and density (ref :16-18) */
#define NP_MIN_R   1.5f
#define NP_MAX_R   9.5f
#define NP_DENSITY 10.f

/* mass of a sphere of radius R (R evaluated three times) (ref :21) */
#define R_TO_M(R, DENSITY) ((4.f * PI * DENSITY / 3.f) * (R) * (R) * (R))
#define GC_R_TO_M(R)       R_TO_M(R, GC_DENSITY) /* ref :23 */
#define NP_R_TO_M(R)       R_TO_M(R, NP_DENSITY) /* ref :24 */
#define MIN_GC_MASS        GC_R_TO_M(GC_MIN_R)   /* ref :25 */

#define MIN_PARTICLES_PER_GALAXY 100 /* ref :27 */

/*
 * Particle placement around a core of radius R in a galaxy of N particles (ref :29-45):
 *   min_dist = R * MIN_PARTICLE_DIST_CR_F                      (hard lower bound)
 *   max_dist = R * MAX_PARTICLE_DIST_CR_F + sqrt(N) * MAX_PARTICLE_DIST_PC_F
 */
#define MIN_PARTICLE_DIST_CR_F 5.f
#define MAX_PARTICLE_DIST_CR_F 10.f
#define MAX_PARTICLE_DIST_PC_F 300.f

/*
 * Galaxy k > 0 is placed at a random distance in [MIN,MAX]_GALAXY_SEPARATION x
 * (max_dist_k + max_dist_parent) from a random earlier galaxy, retried until no other
 * galaxy is closer than the minimum separation (ref :47-62).
 */
#define MIN_GALAXY_SEPARATION 1.4f
#define MAX_GALAXY_SEPARATION 2.0f

/*
 * Generate PARTICLE_COUNT particles in GALAXY_COUNT galaxies using libc rand(); the
 * caller owns (free()s) the result.  PARTICLE_COUNT must be at least
 * MIN_PARTICLES_PER_GALAXY * GALAXY_COUNT, otherwise the call aborts.  Draws from
 * rand() in the same order as the reference generator (ref src/lib/galaxy.c:31-221),
 * so the same srand() seed yields the same particles.  (ref include/galaxy.h:64)
 */
Particle *MakeGalaxies(uint32_t particle_count, uint32_t galaxy_count);

#ifdef __cplusplus
}
#endif

#endif /* NB_GALAXY_H */
