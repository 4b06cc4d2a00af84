/*
 * hg_oracle.c -- CPU restatement of the hg-c-benchmarks hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA
 * path; nothing in the product (libhgb200.so, the shims, the Python host
 * layer) links, imports or calls it.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may execute it.
 *
 * Parity pinning: PINNED.  oracle/Makefile compiles the unmodified reference
 * sources from /root/reference into oracle/_ref/ and tests/test_oracle_cpu.py
 * checks this restatement bit-for-bit against them (and against the
 * reference's own gtest golden matrices and printed results, committed under
 * tests/golden/).
 *
 * Flat data layout (same one the C ABI in include/hgb200.h uses):
 *   bodies : AoS double[3*n]  == reference `Float3D[n]` (src/nbody/nbody-util.h:19-23)
 *   image  : inner_h rows of pitch = inner_w + 2*padding doubles; row r is the
 *            reference's `image[padding + r]` including its left/right pad
 *            columns (src/convolution/convolution-util.c:375-403).  The aliased
 *            top/bottom pad rows are NOT stored: vertical clamp-to-edge is done
 *            by clamping the row index, which is what the aliasing amounts to.
 *
 * Build with -std=c11 -ffp-contract=off: every + - * below is one IEEE-754
 * binary64 operation, in the reference's association order, never fused.
 */
#include <math.h>
#include <stddef.h>
#include <string.h>

#define HG_G   (9.8)    /* src/nbody/nbody-run.h:12 */
#define HG_EPS (0.005)  /* src/nbody/nbody-run.h:13 */

/* One target against all sources, in index order, self pair included.
 * Follows pair_wise_accel (src/nbody/nbody-run.c:6-15) inlined into
 * accel (src/nbody/nbody-run.c:17-25). */
static void oracle_accel(const double *pos, double *out, int index, int n) {
    const double px = pos[3 * index + 0];
    const double py = pos[3 * index + 1];
    const double pz = pos[3 * index + 2];
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int j = 0; j < n; ++j) {
        double dx = pos[3 * j + 0] - px;
        double dy = pos[3 * j + 1] - py;
        double dz = pos[3 * j + 2] - pz;
        double s = dx * dx + dy * dy + dz * dz + HG_EPS;   /* ((a+b)+c)+EPS */
        double f = 1.0 / sqrt(s * s * s);                  /* (s*s)*s      */
        ax += dx * f;
        ay += dy * f;
        az += dz * f;
    }
    out[3 * index + 0] = HG_G * ax;
    out[3 * index + 1] = HG_G * ay;
    out[3 * index + 2] = HG_G * az;
}

/* One step for targets [t0, t1) -- the unit a target-sharded rank computes. */
void oracle_nbody_step_range(const double *pos, double *out, int n, int t0, int t1,
                             int threads) {
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int i = t0; i < t1; ++i) oracle_accel(pos, out, i, n);
}

/* accel() for an arbitrary list of targets; out_sub is AoS double[3*count]. */
void oracle_nbody_accel_subset(const double *pos, int n, const int *targets, int count,
                               double *out_sub, int threads) {
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int k = 0; k < count; ++k) {
        int index = targets[k];
        const double px = pos[3 * index + 0], py = pos[3 * index + 1], pz = pos[3 * index + 2];
        double ax = 0.0, ay = 0.0, az = 0.0;
        for (int j = 0; j < n; ++j) {
            double dx = pos[3 * j + 0] - px;
            double dy = pos[3 * j + 1] - py;
            double dz = pos[3 * j + 2] - pz;
            double s = dx * dx + dy * dy + dz * dz + HG_EPS;
            double f = 1.0 / sqrt(s * s * s);
            ax += dx * f;
            ay += dy * f;
            az += dz * f;
        }
        out_sub[3 * k + 0] = HG_G * ax;
        out_sub[3 * k + 1] = HG_G * ay;
        out_sub[3 * k + 2] = HG_G * az;
    }
}

/* run() (src/nbody/nbody-run.c:27-37).  The reference swaps its LOCAL copies
 * of the two pointers, so after `iterations` steps state k lives in `planets`
 * when k is even and in `buffer` when k is odd: the caller's `planets` ends up
 * holding state `iterations` (even) or `iterations-1` (odd). */
void oracle_nbody_run(double *planets, double *buffer, int n, int iterations, int threads) {
    double *src = planets, *dst = buffer;
    for (int it = 0; it < iterations; ++it) {
        oracle_nbody_step_range(src, dst, n, 0, n, threads);
        double *t = src; src = dst; dst = t;
    }
}

static inline int clampi(int lo, int v, int hi) { return v < lo ? lo : (hi < v ? hi : v); }

/* apply_kernel_to_padded_image + update_borders for output rows [r0, r1)
 * (src/convolution/convolution-run.c:52-81, src/convolution/convolution-util.c:281-294).
 * Correlation, row-major tap order, val starts at +0.0, separate multiply and add. */
void oracle_conv_step_rows(const double *in, double *out, int w, int h, int p,
                           const double *k, int kw, int kh, int r0, int r1, int threads) {
    const ptrdiff_t pitch = (ptrdiff_t)w + 2 * p;
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(static)
    for (int y = r0; y < r1; ++y) {
        double *orow = out + (ptrdiff_t)y * pitch;
        for (int x = 0; x < w; ++x) {
            double val = 0.0;
            for (int ky = 0; ky < kh; ++ky) {
                const double *irow = in + (ptrdiff_t)clampi(0, y + ky - p, h - 1) * pitch;
                for (int kx = 0; kx < kw; ++kx) {
                    /* padded column index (x + kx - p) + p = x + kx */
                    val += k[ky * kw + kx] * irow[x + kx];
                }
            }
            orow[p + x] = val;
        }
        for (int x = 0; x < p; ++x) {           /* update_borders */
            orow[x] = orow[p];
            orow[pitch - 1 - x] = orow[p + w - 1];
        }
    }
}

/* run_on_padded_image (src/convolution/convolution-run.c:41-49).  The reference
 * swaps the CALLER's pointers every iteration, so on return *padded_img is the
 * newest state; here: `cur` holds state `iterations`, `prev` state iterations-1
 * (prev is untouched input garbage/old data if iterations < 2, exactly as the
 * reference leaves the struct that `*buffer` points at). */
void oracle_conv_run(double *cur, double *prev, int w, int h, int p, const double *k, int kw,
                     int kh, int iterations, int threads) {
    const size_t bytes = (size_t)h * (size_t)(w + 2 * p) * sizeof(double);
    double *a = cur, *b = prev;
    for (int it = 0; it < iterations; ++it) {
        oracle_conv_step_rows(a, b, w, h, p, k, kw, kh, 0, h, threads);
        double *t = a; a = b; b = t;
    }
    if (iterations & 1) {
        /* newest state is in the caller's `prev` memory: exchange contents so the
         * flat API matches what the reference's pointer swap shows its caller */
        double *nu = a, *old = b;   /* a == prev memory (newest), b == cur memory */
        for (size_t i = 0; i < bytes / sizeof(double); ++i) {
            double t = nu[i]; nu[i] = old[i]; old[i] = t;
        }
    }
}

/* get_checksum over the inner pixels, sequential row-major
 * (src/convolution/convolution-util.c:63-71 applied to remove_padding's output). */
double oracle_conv_checksum(const double *img, int w, int h, int p) {
    const ptrdiff_t pitch = (ptrdiff_t)w + 2 * p;
    double val = 0.0;
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) val += img[(ptrdiff_t)y * pitch + p + x];
    return val;
}

/* add_padding (src/convolution/convolution-util.c:258-269): dense w*h image ->
 * pitched rows with clamp-to-edge pad columns. */
void oracle_add_padding(const double *dense, double *flat, int w, int h, int p) {
    const ptrdiff_t pitch = (ptrdiff_t)w + 2 * p;
    for (int y = 0; y < h; ++y) {
        double *row = flat + (ptrdiff_t)y * pitch;
        memcpy(row + p, dense + (ptrdiff_t)y * w, (size_t)w * sizeof(double));
        for (int x = 0; x < p; ++x) {
            row[x] = row[p];
            row[pitch - 1 - x] = row[p + w - 1];
        }
    }
}

/* fill_planet (src/nbody/nbody-util.c:92-96) */
void oracle_fill_planets(double *pos, int n) {
    for (int i = 0; i < n; ++i) {
        pos[3 * i + 0] = i * 1.0;
        pos[3 * i + 1] = i * 0.2;
        pos[3 * i + 2] = i * 30.0;
    }
}

/* Extension: softened potential energy U = -G * sum_{i<j} 1/sqrt(|d|^2+EPS),
 * summed in long double.  No reference counterpart; checker for hgb_nbody_energy. */
double oracle_nbody_energy(const double *pos, int n, int threads) {
    long double total = 0.0L;
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64) reduction(+ : total)
    for (int i = 0; i < n; ++i) {
        long double acc = 0.0L;
        for (int j = i + 1; j < n; ++j) {
            double dx = pos[3 * j + 0] - pos[3 * i + 0];
            double dy = pos[3 * j + 1] - pos[3 * i + 1];
            double dz = pos[3 * j + 2] - pos[3 * i + 2];
            long double s = (long double)dx * dx + (long double)dy * dy + (long double)dz * dz +
                            (long double)HG_EPS;
            acc += 1.0L / sqrtl(s);
        }
        total += acc;
    }
    return (double)(-(long double)HG_G * total);
}

/* ---- integrator extension -------------------------------------------
 * The reference has positions only; this is the CPU checker for hgb_nbody_integrate /
 * hgb_nbody_total_energy (include/hgb200.h).  The pairwise part is pair_wise_accel
 * (src/nbody/nbody-run.c:6-15) with the factor weighted by the source mass; every operator
 * is one IEEE binary64 operation, sources in index order, self pair included. */
static void oracle_accel_mass(const double *pos, const double *mass, int index, int n, double *a) {
    const double px = pos[3 * index + 0], py = pos[3 * index + 1], pz = pos[3 * index + 2];
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (int j = 0; j < n; ++j) {
        double dx = pos[3 * j + 0] - px;
        double dy = pos[3 * j + 1] - py;
        double dz = pos[3 * j + 2] - pz;
        double s = dx * dx + dy * dy + dz * dz + HG_EPS;
        double f = 1.0 / sqrt(s * s * s);
        double w = mass[j] * f;
        ax += dx * w;
        ay += dy * w;
        az += dz * w;
    }
    a[0] = HG_G * ax;
    a[1] = HG_G * ay;
    a[2] = HG_G * az;
}

/* `steps` x { a = accel(pos); vel += dt*a; pos += dt*vel } (symplectic Euler), in place. */
void oracle_nbody_integrate(double *pos, double *vel, const double *mass, int n, double dt, int steps,
                            int threads, double *scratch /* 3n doubles */) {
    if (threads < 1) threads = 1;
    for (int it = 0; it < steps; ++it) {
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int i = 0; i < n; ++i) {
            double a[3];
            oracle_accel_mass(pos, mass, i, n, a);
            for (int c = 0; c < 3; ++c) {
                double t = dt * a[c];
                double v = vel[3 * i + c] + t;
                vel[3 * i + c] = v;
                double u = dt * v;
                scratch[3 * i + c] = pos[3 * i + c] + u;
            }
        }
        memcpy(pos, scratch, (size_t)n * 3 * sizeof(double));
    }
}

/* K = 1/2 sum m v^2 and U = -G sum_{i<j} m_i m_j / sqrt(|d|^2+EPS), both in long double. */
void oracle_nbody_total_energy(const double *pos, const double *vel, const double *mass, int n,
                               int threads, double *kinetic, double *potential) {
    long double k = 0.0L, u = 0.0L;
    if (threads < 1) threads = 1;
#pragma omp parallel for num_threads(threads) schedule(dynamic, 64) reduction(+ : k, u)
    for (int i = 0; i < n; ++i) {
        long double acc = 0.0L;
        for (int j = i + 1; j < n; ++j) {
            double dx = pos[3 * j + 0] - pos[3 * i + 0];
            double dy = pos[3 * j + 1] - pos[3 * i + 1];
            double dz = pos[3 * j + 2] - pos[3 * i + 2];
            long double s = (long double)dx * dx + (long double)dy * dy + (long double)dz * dz +
                            (long double)HG_EPS;
            acc += (long double)mass[j] / sqrtl(s);
        }
        u += (long double)mass[i] * acc;
        k += (long double)mass[i] * ((long double)vel[3 * i] * vel[3 * i] +
                                     (long double)vel[3 * i + 1] * vel[3 * i + 1] +
                                     (long double)vel[3 * i + 2] * vel[3 * i + 2]);
    }
    *kinetic = (double)(0.5L * k);
    *potential = (double)(-(long double)HG_G * u);
}
