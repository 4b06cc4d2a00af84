/*
 * Source-level drop-in for the reference's src/nbody/nbody-run.c.
 *
 * Copy (or overlay) this file over src/nbody/nbody-run.c of hg-c-benchmarks, add
 * include/ of this repository to the include path and link with -lhgb200: the
 * reference's main (src/nbody.c) and test suites then drive the B200 kernels
 * through the C ABI of include/hgb200.h.  Valid as C11 and as C++11 (the
 * reference's gtest files #include the .c sources).  There is no CPU fallback:
 * any library error ends the program through the reference's own bail_out().
 */
#include "nbody-run.h"

#include <stdio.h>

#include "hgb200.h"

static void hgb_nbody_or_die(int rc) {
    if (rc != HGB_OK) {
        static char msg[1200];
        snprintf(msg, sizeof msg, "libhgb200 (%d): %s\n", rc, hgb_last_error());
        bail_out(msg);
    }
}

/* Float3D is three packed doubles (src/nbody/nbody-util.h:19-23) */
typedef char hgb_float3d_is_packed[(sizeof(Float3D) == 3 * sizeof(double)) ? 1 : -1];

void pair_wise_accel(Float3D p1, Float3D p2, Float3D *out) {
    hgb_nbody_or_die(hgb_nbody_pairwise(&p1.x, &p2.x, 1, &out->x));
}

void accel(Float3D *planets, Float3D *buffer, int index, int size) {
    hgb_nbody *h = NULL;
    double result[3];
    hgb_nbody_or_die(hgb_nbody_create(&h, size, 1));
    hgb_nbody_or_die(hgb_nbody_upload(h, &planets[0].x));
    hgb_nbody_or_die(hgb_nbody_accel_subset(h, &index, 1, result));
    hgb_nbody_or_die(hgb_nbody_destroy(h));
    buffer[index].x = result[0];
    buffer[index].y = result[1];
    buffer[index].z = result[2];
}

/* number_of_processes (-p) selects how many GPUs the targets are sharded over */
void run(Float3D *planets, Float3D *buffer, int number_of_planets, int iterations, int number_of_processes) {
    hgb_nbody_or_die(hgb_nbody_run(&planets[0].x, &buffer[0].x, number_of_planets, iterations,
                                   number_of_processes));
}
