/*
 * Symbol-level parity of the nbody shim (shim/src/nbody/nbody-run.c) against the reference's own
 * object code.  The unmodified reference file src/nbody/nbody-run.c is compiled from where it lies
 * with -Dpair_wise_accel=ref_pair_wise_accel -Daccel=ref_accel -Drun=ref_run (shim/Makefile), so both
 * implementations of every symbol src/nbody/nbody-run.h:23,34,46 declares live in one binary and are
 * compared bit for bit on the same inputs.  TEST INFRASTRUCTURE (built only where the reference
 * tree exists; runs on the GPU box as a prebuilt binary).
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "nbody-run.h"

void ref_pair_wise_accel(Float3D p1, Float3D p2, Float3D *out);
void ref_accel(Float3D *planets, Float3D *buffer, int index, int size);
void ref_run(Float3D *planets, Float3D *buffer, int number_of_planets, int iterations, int number_of_processes);

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static double rnd(void) {   /* xorshift64*, uniform in [0, 1) */
    rng_state ^= rng_state >> 12; rng_state ^= rng_state << 25; rng_state ^= rng_state >> 27;
    return (double) ((rng_state * 0x2545F4914F6CDD1Dull) >> 11) * (1.0 / 9007199254740992.0);
}

static int failures = 0, checks = 0;
static void same(const char *what, int index, const Float3D *a, const Float3D *b) {
    ++checks;
    if (memcmp(a, b, sizeof(Float3D)) != 0) {
        if (failures++ < 10)
            printf("MISMATCH %s[%d]: shim (%a, %a, %a) reference (%a, %a, %a)\n", what, index, a->x, a->y, a->z,
                   b->x, b->y, b->z);
    }
}

int main(void) {
    enum { N = 777 };
    Float3D *pos = (Float3D *) malloc(sizeof(Float3D) * N), *a = (Float3D *) malloc(sizeof(Float3D) * N);
    Float3D *b = (Float3D *) malloc(sizeof(Float3D) * N), *c = (Float3D *) malloc(sizeof(Float3D) * N);
    Float3D *d = (Float3D *) malloc(sizeof(Float3D) * N);
    for (int i = 0; i < N; ++i) { pos[i].x = rnd() * 4 - 2; pos[i].y = rnd() * 4 - 2; pos[i].z = rnd() * 4 - 2; }
    pos[5] = pos[4];                                   /* coincident bodies: d = 0 */
    pos[9].x = 1e155; pos[10].y = -3e-310;             /* overflowing |d|^2 and a denormal coordinate */

    /* pair_wise_accel: accumulates into *out (src/nbody/nbody-run.c:6-15) */
    for (int i = 0; i < 64; ++i) {
        Float3D s = {rnd(), -rnd(), rnd() * 1e-3}, r = s;
        pair_wise_accel(pos[i], pos[(i * 7 + 3) % N], &s);
        ref_pair_wise_accel(pos[i], pos[(i * 7 + 3) % N], &r);
        same("pair_wise_accel", i, &s, &r);
    }
    { Float3D s = {0, 0, 0}, r = s;                    /* the self pair adds +0 */
      pair_wise_accel(pos[3], pos[3], &s); ref_pair_wise_accel(pos[3], pos[3], &r); same("pair_wise_accel self", 3, &s, &r); }

    /* accel: buffer[index] = G * sum over all sources in order (:17-25); other entries untouched */
    memset(a, 0x5a, sizeof(Float3D) * N);
    memcpy(b, a, sizeof(Float3D) * N);
    const int idx[] = {0, 1, 4, 5, 9, 10, 333, N - 1};
    for (unsigned k = 0; k < sizeof idx / sizeof idx[0]; ++k) {
        accel(pos, a, idx[k], N);
        ref_accel(pos, b, idx[k], N);
    }
    for (int i = 0; i < N; ++i) same("accel", i, &a[i], &b[i]);

    /* run: 0..3 iterations incl. the local pointer swap (:27-37) */
    for (int it = 0; it <= 3; ++it) {
        memcpy(a, pos, sizeof(Float3D) * N); memcpy(c, pos, sizeof(Float3D) * N);
        memset(b, 0, sizeof(Float3D) * N); memset(d, 0, sizeof(Float3D) * N);
        a[9].x = 0.25; c[9].x = 0.25;                  /* keep the trajectories finite */
        run(a, b, N, it, 1);
        ref_run(c, d, N, it, 1);
        for (int i = 0; i < N; ++i) { same("run planets", i, &a[i], &c[i]); same("run buffer", i, &b[i], &d[i]); }
    }
    printf("%s: %d checks, %d mismatches\n", failures ? "FAILED" : "PASSED", checks, failures);
    return failures ? 1 : 0;
}
