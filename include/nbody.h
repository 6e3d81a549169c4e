/*
 * nbody.h -- the World API of nbody-b200.
 *
 * This header is the drop-in boundary: it declares the same types, constants and five
 * World functions as the reference's include/nbody.h (reference lines cited per item),
 * so a program written against the reference compiles and links against libnbody.so
 * unchanged.  Behind it the GPU path is hand-written sm_100a CUDA.
 */
#ifndef NB_H
#define NB_H

#include <math.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Gravitational constant: g = NB_G * mass / dist^2.  (ref include/nbody.h:8) */
#define NB_G 10.0f

/* Two floats.  (ref include/nbody.h:11-13) */
typedef struct V2 {
    float x, y;
} V2;

#ifdef __cplusplus
#define V2_FROM(X, Y) (V2{(X), (Y)})
#define V2_ZERO       (V2{0.0f, 0.0f})
#else
#define V2_FROM(X, Y) ((V2){.x = (X), .y = (Y)}) /* ref include/nbody.h:19 */
#define V2_ZERO       ((V2){.x = 0.0f, .y = 0.0f}) /* ref include/nbody.h:16 */
#endif

/* V2 helpers, same names and meaning as ref include/nbody.h:22-44. */
static inline V2 AddV2(V2 a, V2 b) { return V2_FROM(a.x + b.x, a.y + b.y); }
static inline V2 SubV2(V2 a, V2 b) { return V2_FROM(a.x - b.x, a.y - b.y); }
static inline V2 ScaleV2(V2 v, float f) { return V2_FROM(v.x * f, v.y * f); }
static inline float MagV2(V2 v) { return hypotf(v.x, v.y); }
static inline float SqMagV2(V2 v) { return v.x * v.x + v.y * v.y; }

/*
 * One particle, 32 bytes, array-of-structs on the host side of the API.
 * Field order is ABI (ref include/nbody.h:47-50).  On the device the state lives
 * in struct-of-arrays buffers; this layout only exists at the API surface.
 */
typedef struct Particle {
    V2 pos, vel, acc;
    float mass, radius;
} Particle;

#if defined(__cplusplus)
static_assert(sizeof(Particle) == 32, "Particle is 32 bytes in the ABI");
#elif __STDC_VERSION__ >= 201112L
_Static_assert(sizeof(Particle) == 32, "Particle is 32 bytes in the ABI"); /* cf. ref :54 */
#endif

/* Opaque simulated world with a fixed particle count.  (ref include/nbody.h:58) */
typedef struct World World;

/*
 * Copy SIZE particles from PS into a new World; the caller keeps PS.  Particles with
 * mass > 0 are moved to the front by the same unstable partition as the reference
 * (ref src/lib/world.c:32-46), so GetWorldParticles order matches.  (ref :61)
 */
World *CreateWorld(const Particle *ps, uint32_t size);

/* Release a World and its device buffers; NULL is ignored.  (ref :64) */
void DestroyWorld(World *w);

/*
 * Borrow the World-owned particle array (partitioned order); *SIZE receives the count
 * when SIZE is not NULL.  Downloads from the device only if the device copy is newer.
 * The pointer is stable for the World's life.  (ref :67, src/lib/world.c:91-97)
 */
const Particle *GetWorldParticles(World *w, uint32_t *size);

/* N steps of size DT on the host cores.  A separate implementation, never a fallback
 * for the GPU path.  (ref :70, src/lib/world.c:99-110) */
void UpdateWorld_CPU(World *w, float dt, uint32_t n);

/*
 * N steps of size DT on the B200(s), chained on the device with no host round trip
 * between steps; returns when the device has finished (the reference call is
 * synchronous too, ref src/lib/sim_gpu.c:352-353).  N == 0 is a no-op.  (ref :73)
 */
void UpdateWorld_GPU(World *w, float dt, uint32_t n);

#ifdef __cplusplus
}
#endif

#endif /* NB_H */
