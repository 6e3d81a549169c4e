/*
 * tests/c/test_world_partition.c -- the seven cases of the reference's only unit test
 * (ref test/test_particle_sort.c:27-111), run against the PRODUCT's CreateWorld + GetWorldParticles
 * from plain C, the way a C user of include/nbody.h would.  The reference tests a private copy of the
 * partition loop on int arrays; here the same inputs are particle masses and the expected outputs are
 * the orders the reference test pins (e.g. {0,1,2,0,3} -> {3,1,2,0,0}).
 * Exit code 0 = all passed.  No GPU is touched: CreateWorld/GetWorldParticles are host-only.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nbody.h"

uint32_t NbWorldMassLen(const World *w); /* world.c extra */

static int check(const char *name, const int *in, const int *expect, int n, int expect_count) {
    Particle *ps = calloc((size_t)(n ? n : 1), sizeof *ps);
    for (int i = 0; i < n; i++) {
        ps[i].mass = (float)in[i];
        ps[i].radius = 1.f;
        ps[i].pos.x = (float)i;
    }
    World *w = CreateWorld(ps, (uint32_t)n);
    uint32_t size = 12345;
    const Particle *out = GetWorldParticles(w, &size);
    int ok = size == (uint32_t)n && (int)NbWorldMassLen(w) == expect_count;
    for (int i = 0; ok && i < n; i++) ok = out[i].mass == (float)expect[i];
    for (int i = 0; ok && i < n; i++) ok = ps[i].mass == (float)in[i]; /* caller's array untouched */
    printf("%-32s %s\n", name, ok ? "ok" : "FAILED");
    DestroyWorld(w);
    free(ps);
    return ok ? 0 : 1;
}

int main(void) {
    int bad = 0;
    { const int a[] = {1, 2, 3, 4, 5}, e[] = {1, 2, 3, 4, 5}; bad += check("test_on_non_zeros", a, e, 5, 5); }
    { const int a[] = {0, 0, 0, 0, 0}, e[] = {0, 0, 0, 0, 0}; bad += check("test_on_zeros", a, e, 5, 0); }
    { const int a[] = {1, 2, 3, 0, 0}, e[] = {1, 2, 3, 0, 0}; bad += check("test_on_sorted", a, e, 5, 3); }
    { const int a[] = {0, 0, 1, 2, 3}, e[] = {3, 2, 1, 0, 0}; bad += check("test_on_reverse_sorted_odd", a, e, 5, 3); }
    { const int a[] = {0, 0, 0, 1, 2, 3}, e[] = {3, 2, 1, 0, 0, 0}; bad += check("test_on_reverse_sorted_even", a, e, 6, 3); }
    { const int a[] = {0, 1, 2, 0, 3}, e[] = {3, 1, 2, 0, 0}; bad += check("test_on_unsorted_odd", a, e, 5, 3); }
    { const int a[] = {0, 1, 2, 0, 3, 0}, e[] = {3, 1, 2, 0, 0, 0}; bad += check("test_on_unsorted_even", a, e, 6, 3); }
    DestroyWorld(NULL); /* ref world.c:68 */
    return bad ? 1 : 0;
}
