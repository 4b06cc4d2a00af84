// nbody_sym.cuh -- pair-symmetric all-pairs FP64 step for sm_100a (EXACT mode, one GPU).
//
// pair_wise_accel (reference src/nbody/nbody-run.c:6-15) is antisymmetric bit for bit:
// d(j->i) = -d(i->j) exactly, s and the factor are the same doubles, so the contribution of
// source i to target j is -(d*f) of the pair evaluated from target i's side.  Computing each
// unordered pair once costs 26 FP64 instructions + one add on each side = 32 per PAIR, i.e. 16
// per interaction instead of 29.  What has to be kept is accel()'s summation order (reference
// src/nbody/nbody-run.c:17-25: every target adds its sources j = 0..n-1 in index order; the
// self pair adds +0 and is skipped, accumulators are never -0).  That makes the step a
// wavefront:
//
//   * worker w (one warp) owns the 32*R targets of block w; lane l owns targets R*l..R*l+R-1
//     with position and accumulator in registers;
//   * the sources j < (block start) stream through the warp as a systolic array: at step s lane
//     l handles source j = s - l, so every lane sees j ascending (its targets' order) and every
//     source meets the targets i ascending (the source's own order, as a target);
//   * the partial sum of source j as a target ("chain" accumulator) travels with it: it sits in
//     a per-warp shared-memory ring next to the source position, each lane subtracts its
//     R pair forces from it in turn (read-modify-write, __syncwarp between steps);
//   * the ring (three chunks of 32 sources) is fed chunk by chunk from global memory, where the
//     previous worker left the chain accumulators as tagged words (payload + tag in one 8-byte
//     store: no fence, no flag; the loads are issued one chunk ahead and the tags checked when the
//     chunk is opened).  Worker w publishes a chunk two chunks (64 steps) after it took it in --
//     the minimum distance between consecutive workers.  Workers take tickets in index order,
//     so worker w only ever waits for a worker that is already running: no deadlock, and the
//     grid is persistent (warps loop over tickets);
//   * the block's own pairs (the diagonal chunks) run through the same skewed schedule with
//     predicates: j < i evaluates the pair, j == i turns the target's finished lower sum into
//     the chain accumulator of source j.
// tests/sym_model.py is a lock-step CPU model of exactly this index logic.
//
// The last worker leaves acc[i] = sum_j pair(i, j) in order for every body; a second tiny kernel
// writes G * acc.  If any operand ever leaves the range of the branch-free IEEE sequences
// (never for sane finite positions) the kernel raises a flag and the ordinary one-target-per-
// thread kernel redoes the step from the untouched source state.
#pragma once
#include <type_traits>

#include "nbody_kernels.cuh"

namespace hgb {

constexpr int kSymCtlInts = 128;   // ctl[0] = ticket counter, ctl[1] = range flag, ctl[8..21] profiling counters, ctl[24..] (PROF) wait cycles by position in the row x worker size class
// Control words of the pair-symmetric kernels (sym_ctl: kSymCtlInts ints, tri_ctl: kTriCtlInts ints).  One owner per
// slot; the host resets kCtlTicket, the range flag(s) and the profiling ranges per step and leaves the abort flag
// and the abandoned-poll record alone until it has read them.
constexpr int kCtlTicket = 0;        // next item / row to hand out
constexpr int kCtlRange = 1;         // sym_ctl: an operand left the range of the packed words -> the step is redone
constexpr int kCtlAbort = 2;         // a poll hit the spin limit: every worker gives up
constexpr int kCtlAbortFirst = 3;    // latch: the first poll to give up fills the record below
constexpr int kCtlRangeParity = 4;   // tri_ctl: range flag of even / odd steps (4, 5)
constexpr int kCtlWantTag = 6;       // abandoned-poll record: tag waited for, tag found (7, 24), chain word (25), row word (26)
constexpr int kCtlSpans = 8;         // 7 x u64 clock64 spans / counts of the profiling variants: ints 8..21
constexpr int kCtlSpanInts = 14;
constexpr int kCtlSymHist = 24;      // sym_ctl, first generation only (no abandoned-poll record there): 40 x u64 histogram
constexpr int kCtlSegPolled = 34;    // segmented kernel, PROF: chunk opens that had to poll (u64: 34, 35)
constexpr int kCtlSegAhead = 40;     // ... chunks the block above has published at an item start (8 x u64: 40..55)
constexpr int kCtlSegPolledQuarter = 56;   // ... polled opens by quarter of the blocks (4 x u64: 56..63)
constexpr int kCtlSegProf0 = 34, kCtlSegProfInts = 30;   // the segmented kernel's PROF range for the per-step reset

// Hand-off words (the "LL" scheme of collective libraries): every 64-bit word carries 32 bits of
// payload and a 32-bit tag, so a naturally aligned 8-byte store publishes data and flag together --
// no fence, no separate flag, and the consumer may issue its loads early and look at the tags later.
// One double = two words; tag = (launch epoch << 21) | (producing worker + 1), never 0.
__device__ __forceinline__ ulonglong2 ll_load(const ulonglong2 *p) {
    ulonglong2 v;
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(v.x), "=l"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void ll_store(ulonglong2 *p, ulonglong2 v) {
    asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(p), "l"(v.x), "l"(v.y) : "memory");
}
__device__ __forceinline__ ulonglong2 ll_pack(double d, unsigned tag) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(d), t = (unsigned long long)tag << 32;
    return make_ulonglong2(t | (b & 0xffffffffull), t | (b >> 32));
}
__device__ __forceinline__ bool ll_ok(ulonglong2 v, unsigned tag) {
    return (unsigned)(v.x >> 32) == tag && (unsigned)(v.y >> 32) == tag;
}
__device__ __forceinline__ double ll_unpack(ulonglong2 v) {
    return __longlong_as_double((long long)((v.y << 32) | (v.x & 0xffffffffull)));
}

// d * f of pair_wise_accel for P independent (target, source) pairs, the dependency chains
// interleaved stage by stage (in-order issue: independent chains must sit next to each other).
template <int P>
__device__ __forceinline__ void pair_forces(const double (&px)[P], const double (&py)[P],
                                            const double (&pz)[P], const double (&xj)[P],
                                            const double (&yj)[P], const double (&zj)[P],
                                            double (&vx)[P], double (&vy)[P], double (&vz)[P],
                                            unsigned &key) {
    double t[P], y[P], g[P], s[P], a[P], e[P], p[P], q[P];
#pragma unroll
    for (int u = 0; u < P; ++u) {
        vx[u] = __dsub_rn(xj[u], px[u]); vy[u] = __dsub_rn(yj[u], py[u]); vz[u] = __dsub_rn(zj[u], pz[u]);
    }
#pragma unroll
    for (int u = 0; u < P; ++u) {
        s[u] = __dmul_rn(vx[u], vx[u]); a[u] = __dmul_rn(vy[u], vy[u]); e[u] = __dmul_rn(vz[u], vz[u]);
    }
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], a[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], e[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], kEps);
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(s[u], s[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(t[u], s[u]);            // (s*s)*s
#pragma unroll
    for (int u = 0; u < P; ++u) { y[u] = mufu_rsqrt(t[u]); key = max(key, range_key(t[u])); }
#pragma unroll
    for (int u = 0; u < P; ++u) a[u] = __dmul_rn(y[u], y[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(t[u], -a[u], 1.0);
#pragma unroll
    for (int u = 0; u < P; ++u) { p[u] = __fma_rn(e[u], 0.375, 0.5); q[u] = __dmul_rn(y[u], e[u]); }
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = __fma_rn(p[u], q[u], y[u]);       // ~1/sqrt(t)
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __dmul_rn(t[u], y[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(-g[u], g[u], t[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __fma_rn(e[u], half_of(y[u]), g[u]);  // RN(sqrt(t))
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = odd_up(y[u]);                     // see inv_sqrt_ieee
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(y[u], -g[u], 1.0);
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = __fma_rn(y[u], e[u], y[u]);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(y[u], -g[u], 1.0);
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = __fma_rn(y[u], e[u], y[u]);       // RN(1/g)
#pragma unroll
    for (int u = 0; u < P; ++u) {
        vx[u] = __dmul_rn(vx[u], y[u]); vy[u] = __dmul_rn(vy[u], y[u]); vz[u] = __dmul_rn(vz[u], y[u]);
    }
}

constexpr int kSymWarps = 4;   // warps (independent workers) per CTA

// R targets per lane, U systolic steps per unrolled group (R*U independent pair chains in
// flight per thread), CTAS resident CTAs per SM (register budget).  src: SoA positions with
// component stride npad (zero beyond n); ll: tagged chain accumulators, [chunk][x,y,z][32 lanes] x 16 B;
// epoch_tag: (launch counter mod 2048) << 21, so words of earlier launches never match.
// PROF: lane 0 of every warp accumulates clock64() spans (flag wait, chunk loads, chunk release, the
// predicate-free loop, predicated steps, whole workers) into ctl[8..] -- a diagnostic instantiation.
template <int R, int U, int CTAS, bool PROF = false>
__global__ void __launch_bounds__(kSymWarps * 32, CTAS)
nbody_sym_kernel(const double *__restrict__ src, int npad, int n, ulonglong2 *ll, unsigned epoch_tag, int *ctl, int hyst) {
    constexpr int BS = 32 * R, CH = 32, RING = 3 * CH;   // block of targets, hand-off chunk, ring entries
    constexpr unsigned FULL = 0xffffffffu;
    static_assert(CH % U == 0, "groups of U steps must not straddle a chunk boundary");
    extern __shared__ __align__(16) double sym_smem[];
    const int lane = threadIdx.x & 31;
    double *rx = sym_smem + (threadIdx.x >> 5) * 6 * RING;
    double *ry = rx + RING, *rz = ry + RING, *rax = rz + RING, *ray = rax + RING, *raz = ray + RING;
    const double *sx = src, *sy = src + npad, *sz = src + 2 * (size_t)npad;
    const int T = (n + BS - 1) / BS;
    unsigned key = 0;
    long long t_wait = 0, t_open = 0, t_close = 0, t_fast = 0, t_slow = 0, t_work = 0, n_chunks = 0;
    auto now = [] { return PROF ? clock64() : 0ll; };

    for (;;) {
        const long long w_begin = now();
        int w = 0;
        if (lane == 0) w = atomicAdd(&ctl[0], 1);   // tickets in index order: worker w-1 is already running
        w = __shfl_sync(FULL, w, 0);
        if (w >= T) break;
        const int i0 = w * BS, jend = min(i0 + BS, n), tgt0 = i0 + R * lane;
        double px[R], py[R], pz[R], ax[R], ay[R], az[R];
        bool valid[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            px[r] = sx[tgt0 + r]; py[r] = sy[tgt0 + r]; pz[r] = sz[tgt0 + r];
            ax[r] = 0.0; ay[r] = 0.0; az[r] = 0.0;
            valid[r] = tgt0 + r < n;
        }
        int e = (RING - lane) % RING;   // ring entry of source j = s - lane
        const int last_chunk_of_row = (jend - 1) / CH;

        // Prefetch registers: one entry (this lane's) of the chunk that will be opened next -- the
        // source position and, for a lower chunk, the tagged chain-accumulator words.  The loads are
        // issued a whole chunk (32 steps) ahead; the tags are only looked at when the chunk is opened.
        int pf_chunk = -1;
        double pf_x = 0.0, pf_y = 0.0, pf_z = 0.0;
        ulonglong2 pf_ax = make_ulonglong2(0, 0), pf_ay = pf_ax, pf_az = pf_ax;
        const unsigned want_tag = epoch_tag | (unsigned)w;         // written by worker w-1
        const unsigned my_tag = epoch_tag | (unsigned)(w + 1);
        auto prefetch_chunk = [&](int c) {
            const int g = c * CH + lane;
            pf_x = sx[g]; pf_y = sy[g]; pf_z = sz[g];
            if (c < w * R) {
                const ulonglong2 *q = ll + (size_t)c * 3 * CH + lane;
                pf_ax = ll_load(q); pf_ay = ll_load(q + CH); pf_az = ll_load(q + 2 * CH);
            }
            pf_chunk = c;
        };
        // chunk c (32 sources) enters the ring: positions always, chain accumulators once worker w-1
        // has published them (chunks of the block itself get theirs from the owning lanes)
        auto open_chunk = [&](int c) {
            const int slot = (c % 3) * CH + lane;
            const bool lower = c < w * R;
            const long long a0 = now();
            if (pf_chunk != c) prefetch_chunk(c);
            if (lower) {
                const ulonglong2 *q = ll + (size_t)c * 3 * CH + lane;
                // every lane polls its own three entries; the warp leaves together
                if (!__all_sync(FULL, ll_ok(pf_ax, want_tag) && ll_ok(pf_ay, want_tag) && ll_ok(pf_az, want_tag))) {
                    // This worker has caught up with the one before it.  hyst > 0: let it fall `hyst` chunks behind
                    // (wait for a later chunk of the predecessor's stream first), so that the next prefetches find
                    // their words instead of paying a round trip at every chunk.
                    const int ahead = min(c + hyst, (w - 1) * R - 1);
                    if (hyst > 0 && ahead > c) {
                        const ulonglong2 *qa = ll + (size_t)ahead * 3 * CH + lane;
                        while (!__all_sync(FULL, ll_ok(ll_load(qa + 2 * CH), want_tag))) __nanosleep(100);
                    }
                    do {
                        pf_ax = ll_load(q); pf_ay = ll_load(q + CH); pf_az = ll_load(q + 2 * CH);
                        if (__all_sync(FULL, ll_ok(pf_ax, want_tag) && ll_ok(pf_ay, want_tag) && ll_ok(pf_az, want_tag))) break;
                        __nanosleep(20);
                    } while (true);
                }
            }
            const long long a1 = now();
            rx[slot] = pf_x; ry[slot] = pf_y; rz[slot] = pf_z;
            if (lower) { rax[slot] = ll_unpack(pf_ax); ray[slot] = ll_unpack(pf_ay); raz[slot] = ll_unpack(pf_az); }
            __syncwarp();
            if (PROF) {
                t_wait += a1 - a0; t_open += now() - a1; ++n_chunks;
                // where the waiting happens: tenth of the row (by chunk) x quarter of the worker range
                if (lane == 0 && (c & 63) == 0 && c + 40 < w * R) {   // every 64th chunk: how many chunks ahead the predecessor has published (0..32+)
                    int ahead = 0;
                    for (int k = 1; k <= 32; ++k) {
                        const ulonglong2 t = ll_load(ll + (size_t)(c + k) * 3 * CH);
                        if ((unsigned)(t.x >> 32) != want_tag) break;
                        ahead = k;
                    }
                    atomicAdd(reinterpret_cast<unsigned long long *>(ctl + 24) + 32 + min(7, ahead >= 32 ? 7 : ahead >= 16 ? 6 : ahead >= 8 ? 5 : ahead >= 4 ? 4 : ahead), 1ull);
                }
                if (lane == 0 && a1 - a0 > 0) {
                    const int tenth = min(9, 10 * c / (last_chunk_of_row + 1)), quarter = min(3, 4 * w / T);
                    atomicAdd(reinterpret_cast<unsigned long long *>(ctl + 24) + (quarter * 8 + min(7, tenth * 8 / 10)), (unsigned long long)(a1 - a0));
                }
            }
        };
        // every lane is through with chunk c: publish its chain accumulators for worker w+1
        auto close_chunk = [&](int c) {
            const int slot = (c % 3) * CH + lane;
            const long long a0 = now();
            __syncwarp();
            ulonglong2 *q = ll + (size_t)c * 3 * CH + lane;
            ll_store(q, ll_pack(rax[slot], my_tag));
            ll_store(q + CH, ll_pack(ray[slot], my_tag));
            ll_store(q + 2 * CH, ll_pack(raz[slot], my_tag));
            if (PROF) { __syncwarp(); t_close += now() - a0; }
        };
        // U predicated steps starting at step s0 (a multiple of U, so the group stays inside one chunk):
        // fill / drain of the systolic skew and the diagonal chunks.  These steps are the critical
        // path of the wavefront (worker w+1 cannot enter its own block before worker w left its), so
        // all R*U pairs are evaluated unconditionally and interleaved like the predicate-free path;
        // the predicates only select what is added (j < i), what starts a chain (j == i) and what is
        // dropped (j > i, padding targets, lanes outside the stream: those read a zero source).
        auto edge_group = [&](int s0) {
            int es[U];
            double qx[R * U], qy[R * U], qz[R * U], jx[R * U], jy[R * U], jz[R * U];
            double vx[R * U], vy[R * U], vz[R * U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                es[u] = e;
                e = (e + 1 == RING) ? 0 : e + 1;
                const int j = s0 + u - lane;
                const bool act = j >= 0 && j < jend;
                const double xj = act ? rx[es[u]] : 0.0, yj = act ? ry[es[u]] : 0.0, zj = act ? rz[es[u]] : 0.0;
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    qx[u * R + r] = px[r]; qy[u * R + r] = py[r]; qz[u * R + r] = pz[r];
                    jx[u * R + r] = xj; jy[u * R + r] = yj; jz[u * R + r] = zj;
                }
            }
            pair_forces<R * U>(qx, qy, qz, jx, jy, jz, vx, vy, vz, key);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int j = s0 + u - lane;
                const bool act = j >= 0 && j < jend;
                bool have_u = act && j < tgt0;   // an earlier lane (or worker) already carries this chain
                double ux = 0.0, uy = 0.0, uz = 0.0;
                if (have_u) { ux = rax[es[u]]; uy = ray[es[u]]; uz = raz[es[u]]; }
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    // branch-free (selects): the diagonal is the serial tail of every worker, and divergent branches
                    // around single DADDs made an edge group cost twice a predicate-free one
                    const int i = tgt0 + r;
                    const bool ok = act && i < n;
                    const bool eq = ok && j == i;      // lower sum of target i is complete: it becomes j's chain
                    const bool lt = ok && j < i;
                    const double nax = __dadd_rn(ax[r], vx[u * R + r]), nay = __dadd_rn(ay[r], vy[u * R + r]),
                                 naz = __dadd_rn(az[r], vz[u * R + r]);
                    const double nux = __dsub_rn(ux, vx[u * R + r]), nuy = __dsub_rn(uy, vy[u * R + r]),
                                 nuz = __dsub_rn(uz, vz[u * R + r]);
                    ux = eq ? ax[r] : (lt ? nux : ux); uy = eq ? ay[r] : (lt ? nuy : uy); uz = eq ? az[r] : (lt ? nuz : uz);
                    ax[r] = lt ? nax : ax[r]; ay[r] = lt ? nay : ay[r]; az[r] = lt ? naz : az[r];
                    have_u = have_u || eq;
                }
                if (have_u) { rax[es[u]] = ux; ray[es[u]] = uy; raz[es[u]] = uz; }
                __syncwarp();
            }
        };

        // U full steps: every lane's source is below the block, no index predicates.  ALLV: the
        // block has no padding targets (always, unless it is the last block of a ragged n).
        auto fast_group = [&](auto allv) {
            constexpr bool ALLV = decltype(allv)::value;
            int es[U];
            double qx[R * U], qy[R * U], qz[R * U], jx[R * U], jy[R * U], jz[R * U];
            double vx[R * U], vy[R * U], vz[R * U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                es[u] = e;
                e = (e + 1 == RING) ? 0 : e + 1;
                const double xj = rx[es[u]], yj = ry[es[u]], zj = rz[es[u]];
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    qx[u * R + r] = px[r]; qy[u * R + r] = py[r]; qz[u * R + r] = pz[r];
                    jx[u * R + r] = xj; jy[u * R + r] = yj; jz[u * R + r] = zj;
                }
            }
            pair_forces<R * U>(qx, qy, qz, jx, jy, jz, vx, vy, vz, key);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                double ux = rax[es[u]], uy = ray[es[u]], uz = raz[es[u]];
#pragma unroll
                for (int r = 0; r < R; ++r) {   // targets ascending: the chain's order
                    ax[r] = __dadd_rn(ax[r], vx[u * R + r]);
                    ay[r] = __dadd_rn(ay[r], vy[u * R + r]);
                    az[r] = __dadd_rn(az[r], vz[u * R + r]);
                    if (ALLV || valid[r]) {
                        ux = __dsub_rn(ux, vx[u * R + r]);
                        uy = __dsub_rn(uy, vy[u * R + r]);
                        uz = __dsub_rn(uz, vz[u * R + r]);
                    }
                }
                rax[es[u]] = ux; ray[es[u]] = uy; raz[es[u]] = uz;
                __syncwarp();   // lane l+1 reads this entry in the next step
            }
        };
        const bool full = i0 + BS <= n;

        int s = 0;
        bool early_closed = false;
        const int s_end = jend - 1 + 31, last_chunk = (jend - 1) / CH;
#pragma unroll 1
        while (s <= s_end) {
            if (s % CH == 0 && s / CH <= last_chunk) {
                const int c = s / CH;
                if (c >= 2) close_chunk(c - 2);   // lane 31 left it 2 steps ago at the latest
                open_chunk(c);
                if (c + 1 <= last_chunk) prefetch_chunk(c + 1);
            } else if (s == (last_chunk + 1) * CH && last_chunk >= 1 && !early_closed) {
                // The same rule one chunk further: every lane has left chunk last_chunk-1 (the first chunk of the block
                // itself).  Worker w+1 can stream it while this worker finishes its diagonal: the serial tail per
                // worker (stream the predecessor's two chunks, then the own diagonal) loses one chunk time, and with a
                // wavefront whose leaders advance 4 chunks per tail that tail -- not the FP64 pipe -- sets the pace.
                close_chunk(last_chunk - 1);
                early_closed = true;
            }
            const int lim = min(i0, (s / CH + 1) * CH);   // end of the predicate-free run inside this chunk
            const long long b0 = now();
            if (s >= 31 && s + U <= lim) {
                if (full) {
#pragma unroll 1
                    do { fast_group(std::true_type{}); s += U; } while (s + U <= lim);
                } else {
#pragma unroll 1
                    do { fast_group(std::false_type{}); s += U; } while (s + U <= lim);
                }
                if (PROF) t_fast += now() - b0;
            } else {
                edge_group(s);   // may run up to U-1 steps past s_end: every lane is inactive there
                s += U;
                if (PROF) t_slow += now() - b0;
            }
        }
        if (last_chunk >= 1 && !early_closed) close_chunk(last_chunk - 1);
        close_chunk(last_chunk);   // the block's own chunks, tagged w+1, let worker w+1 in
        if (PROF) t_work += now() - w_begin;
    }
    if (PROF && lane == 0) {
        unsigned long long *cnt = reinterpret_cast<unsigned long long *>(ctl + 8);
        atomicAdd(cnt + 0, (unsigned long long)t_wait); atomicAdd(cnt + 1, (unsigned long long)t_open);
        atomicAdd(cnt + 2, (unsigned long long)t_close); atomicAdd(cnt + 3, (unsigned long long)t_fast);
        atomicAdd(cnt + 4, (unsigned long long)t_slow); atomicAdd(cnt + 5, (unsigned long long)t_work);
        atomicAdd(cnt + 6, (unsigned long long)n_chunks);
    }
    if (key >= kRangeLimit) ctl[1] = 1;
}

// buffer[i] = G * acc[i] (reference src/nbody/nbody-run.c:22-24); acc = what the last worker published
__global__ void nbody_sym_finish_kernel(const ulonglong2 *__restrict__ ll, int npad, int n, PeerOut out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const ulonglong2 *q = ll + (size_t)(i / 32) * 3 * 32 + (i % 32);
    const double gx = __dmul_rn(kG, ll_unpack(q[0])), gy = __dmul_rn(kG, ll_unpack(q[32])),
                 gz = __dmul_rn(kG, ll_unpack(q[64]));
    for (int p = 0; p < out.count; ++p) {
        double *o = out.base[p];
        o[i] = gx; o[(size_t)npad + i] = gy; o[2 * (size_t)npad + i] = gz;
    }
}

}  // namespace hgb
