// nbody_seg.cuh -- pair-symmetric all-pairs FP64 step for sm_100a: the FIRST generation's systolic warp
// (nbody_sym.cuh: 16 FP64 instructions per interaction, no transposition through shared memory, 94 % of the
// FP64 pipe while it computes) on the SECOND generation's 2D schedule (nbody_tri.cuh: rows cut into segments,
// items handed out in anti-diagonal key order).
//
// Why: in nbody_sym_kernel a worker is a whole row, tickets go in row order, and the 1184 concurrent workers
// form ONE tandem line in which everybody follows its predecessor at the minimum distance -- 80 % of the chunk
// opens find nothing published beyond the chunk itself, every delay upstream is inherited downstream, 20 % of
// the worker time waits (profiles/r3t_sym_wait_histogram.txt).  Here an ITEM is a block of 32 R targets
// against one SEGMENT of `seg` sources; item (w, k) streams the chain words of (w-1, k) and continues the row
// sums of (w, k-1), which it gets through a tagged buffer.  Consecutive blocks of one segment are handed out a
// whole anti-diagonal of tickets apart, so a worker finds its predecessor's words published long before it
// needs them, and a delay somewhere is absorbed by that slack instead of travelling down a line.
//
// Arithmetic and order are those of nbody_sym_kernel (EXACT mode, bit-identical to accel(), reference
// src/nbody/nbody-run.c:17-25): lane l owns targets R l .. R l + R - 1 of the block, at step s it handles source
// j = s - l, the chain accumulator of source j travels with it through the per-warp ring.  What an item adds:
//   * the stream starts at j_begin = k seg and ends at j_stop = (k+1) seg (or, in the row's last item, at the end
//     of the block itself, diagonal included); the skew fills and drains once per item (31 + 31 predicated steps
//     for >= 8192 predicate-free ones);
//   * the block's 32 R x 3 row sums come from item (w, k-1) and go to (w, k+1) as tagged words;
//   * every poll has a spin limit: a hand-off that never arrives raises ctl[2] and the step ends with garbage
//     and an error instead of hanging the GPU.
#pragma once
#include "nbody_sym.cuh"

namespace hgb {

// pair_forces (nbody_sym.cuh) with a hook after every stage: hook(std::integral_constant<int, stage>) lets the caller
// weave other work (the ordered read-modify-write steps of the PREVIOUS group of sources) between the stages of the
// eight independent force chains, in one basic block.
template <int P, class Hook>
__device__ __forceinline__ void pair_forces_hooked(const double (&px)[P], const double (&py)[P], const double (&pz)[P],
                                                   const double (&xj)[P], const double (&yj)[P], const double (&zj)[P],
                                                   double (&vx)[P], double (&vy)[P], double (&vz)[P], unsigned &key, Hook hook) {
    double t[P], y[P], g[P], s[P], a[P], e[P], p[P], q[P];
#define HGB_STAGE(N_, BODY_)                      \
    _Pragma("unroll") for (int u = 0; u < P; ++u) { BODY_ } \
    hook(std::integral_constant<int, N_>{});
    HGB_STAGE(0, vx[u] = __dsub_rn(xj[u], px[u]); vy[u] = __dsub_rn(yj[u], py[u]); vz[u] = __dsub_rn(zj[u], pz[u]);)
    HGB_STAGE(1, s[u] = __dmul_rn(vx[u], vx[u]); a[u] = __dmul_rn(vy[u], vy[u]); e[u] = __dmul_rn(vz[u], vz[u]);)
    HGB_STAGE(2, s[u] = __dadd_rn(s[u], a[u]);)
    HGB_STAGE(3, s[u] = __dadd_rn(s[u], e[u]);)
    HGB_STAGE(4, s[u] = __dadd_rn(s[u], kEps);)
    HGB_STAGE(5, t[u] = __dmul_rn(s[u], s[u]);)
    HGB_STAGE(6, t[u] = __dmul_rn(t[u], s[u]);)
    HGB_STAGE(7, y[u] = mufu_rsqrt(t[u]); key = max(key, range_key(t[u]));)
    HGB_STAGE(8, a[u] = __dmul_rn(y[u], y[u]);)
    HGB_STAGE(9, e[u] = __fma_rn(t[u], -a[u], 1.0);)
    HGB_STAGE(10, p[u] = __fma_rn(e[u], 0.375, 0.5); q[u] = __dmul_rn(y[u], e[u]);)
    HGB_STAGE(11, y[u] = __fma_rn(p[u], q[u], y[u]);)
    HGB_STAGE(12, g[u] = __dmul_rn(t[u], y[u]);)
    HGB_STAGE(13, e[u] = __fma_rn(-g[u], g[u], t[u]);)
    HGB_STAGE(14, g[u] = __fma_rn(e[u], half_of(y[u]), g[u]);)
    HGB_STAGE(15, y[u] = odd_up(y[u]);)
    HGB_STAGE(16, e[u] = __fma_rn(y[u], -g[u], 1.0);)
    HGB_STAGE(17, y[u] = __fma_rn(y[u], e[u], y[u]);)
    HGB_STAGE(18, e[u] = __fma_rn(y[u], -g[u], 1.0);)
    HGB_STAGE(19, y[u] = __fma_rn(y[u], e[u], y[u]);)
    HGB_STAGE(20, vx[u] = __dmul_rn(vx[u], y[u]); vy[u] = __dmul_rn(vy[u], y[u]); vz[u] = __dmul_rn(vz[u], y[u]);)
#undef HGB_STAGE
}

struct SegArgs {
    const double *src;        // SoA positions, component stride npad, zero beyond n
    int npad, n;
    int T;                    // blocks of 32 R targets
    int seg, tr;              // sources per segment (multiple of 32), blocks per tile (consecutive tickets)
    const int2 *tiles;        // tiles (b, k) in ticket order
    int ntiles;
    ulonglong2 *ll;           // chain words [chunk of 32 sources][x, y, z][32] x 16 B, tag epoch | producing block + 1
    ulonglong2 *racc;         // row sums between the items of a block: [block][R x 3][32] x 16 B, tag epoch | next item
    unsigned epoch_tag;
    int *ctl;                 // ctl[0] ticket, ctl[1] range flag, ctl[2] abort flag, ctl[8..] span counters (PROF)
    long long spin_limit;
    int hyst;                 // chunks a block falls behind its predecessor after catching up with it (0: none)
    // several GPUs (blocks dealt in turns of `deal`, as in nbody_tri.cuh): where the last block of a turn publishes, who to tell
    ulonglong2 *ll_down;      // chain-word buffer of the GPU that runs the next turn (== ll on one GPU)
    int deal;                 // blocks per turn of a GPU
    int npeers;
    int *ctl_all[kMaxPeers];  // every GPU's ctl: range and abort flags are raised everywhere
    int range_slot;           // ctl index of this step's range flag
};

// raise flag ctl[which] on every GPU (range flag: before the words computed from the operand are published)
__device__ __noinline__ void seg_raise(const SegArgs &a, int which) {
    for (int p = 0; p < a.npeers; ++p) *reinterpret_cast<volatile int *>(a.ctl_all[p] + which) = 1;
    __threadfence_system();
}

// poll one tagged word (all lanes together); gives up after the spin limit and raises the abort flag
__device__ __noinline__ ulonglong2 seg_poll(const SegArgs &a, const ulonglong2 *q, unsigned tag, int &dead) {
    const long long t0 = clock64();
    ulonglong2 w = make_ulonglong2(0, 0);
    while (!dead) {
        w = ll_load(q);
        if (__all_sync(0xffffffffu, ll_ok(w, tag))) break;
        bool g = *reinterpret_cast<volatile int *>(a.ctl + 2) != 0;
        if (!g && clock64() - t0 > a.spin_limit) {
            if ((threadIdx.x & 31) == 0 && atomicCAS(a.ctl + 3, 0, 1) == 0) {
                a.ctl[6] = (int)tag; a.ctl[7] = (int)(w.x >> 32);
                a.ctl[25] = (int)(q - a.ll); a.ctl[26] = (int)(q - a.racc);
            }
            seg_raise(a, 2);
            g = true;
        }
        dead = __any_sync(0xffffffffu, g);
        __nanosleep(40);
    }
    return w;
}

// R targets per lane, U systolic steps per unrolled group, CTAS resident CTAs per SM.  PROF: lane 0 of every
// warp accumulates clock64 spans in ctl[8..]: waits for chain words, waits for row sums, the predicate-free loop,
// predicated groups, whole items, item count, chunk count.
// PIPE: software pipeline inside the predicate-free run of a chunk -- the forces of group g+1 are evaluated while the
// ordered read-modify-write steps of group g are applied (one basic block per group, the steps woven between the
// stages of the force chains), instead of "32-instruction chains, then four latency-bound steps" per group.
template <int R, int U, int CTAS, bool PROF = false, bool PIPE = false>
__global__ void __launch_bounds__(kSymWarps * 32, CTAS)
nbody_seg_kernel(const __grid_constant__ SegArgs a) {
    static_assert(!PIPE || U == 4 || U == 2, "the pipelined group weaves four (or two) steps into the force stages");
    constexpr int BS = 32 * R, CH = 32, RING = 3 * CH;
    constexpr unsigned FULL = 0xffffffffu;
    static_assert(CH % U == 0, "groups of U steps must not straddle a chunk boundary");
    extern __shared__ __align__(16) double seg_smem[];
    const int lane = threadIdx.x & 31;
    double *rx = seg_smem + (threadIdx.x >> 5) * 6 * RING;
    double *ry = rx + RING, *rz = ry + RING, *rax = rz + RING, *ray = rax + RING, *raz = ray + RING;
    const double *sx = a.src, *sy = a.src + a.npad, *sz = a.src + 2 * (size_t)a.npad;
    const int n = a.n;
    ulonglong2 *const ll = a.ll;
    unsigned key = 0;
    int dead = 0;
    long long t_wait = 0, t_racc = 0, t_fast = 0, t_slow = 0, t_item = 0, n_items = 0, n_chunks = 0, n_polled = 0;
    auto now = [] { return PROF ? clock64() : 0ll; };

    for (;;) {
        int ticket = 0;
        if (lane == 0) ticket = atomicAdd(&a.ctl[0], 1);
        ticket = __shfl_sync(FULL, ticket, 0);
        const int ti = ticket / a.tr, rr = ticket - ti * a.tr;
        if (ti >= a.ntiles) break;
        const int2 bk = a.tiles[ti];
        const int w = bk.x * a.tr + rr, k = bk.y;
        if (w >= a.T) continue;
        const int i0 = w * BS, K = i0 / a.seg;             // K = the block's last item
        if (k > K) continue;                               // this block's row ends before segment k
        const long long it0 = now();
        const bool last_item = k == K;
        const int jend = min(i0 + BS, n), tgt0 = i0 + R * lane;
        const int j_begin = k * a.seg, j_stop = last_item ? jend : (k + 1) * a.seg;   // sources of this item
        double px[R], py[R], pz[R], ax[R], ay[R], az[R];
        bool valid[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            px[r] = sx[tgt0 + r]; py[r] = sy[tgt0 + r]; pz[r] = sz[tgt0 + r];
            ax[r] = 0.0; ay[r] = 0.0; az[r] = 0.0;
            valid[r] = tgt0 + r < n;
        }
        if (k > 0) {   // the block's row sums so far, left by item (w, k-1)
            const long long r0 = now();
            const ulonglong2 *q = a.racc + (size_t)w * (3 * R * 32) + lane;
            const unsigned tag = a.epoch_tag | (unsigned)k;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                ax[r] = ll_unpack(seg_poll(a, q + (3 * r + 0) * 32, tag, dead));
                ay[r] = ll_unpack(seg_poll(a, q + (3 * r + 1) * 32, tag, dead));
                az[r] = ll_unpack(seg_poll(a, q + (3 * r + 2) * 32, tag, dead));
            }
            if (PROF) t_racc += now() - r0;
        }
        int e = (((j_begin - lane) % RING) + RING) % RING;   // ring entry of source j = s - lane at s = j_begin
        if (PROF && lane == 0 && w > 0 && j_begin / CH + 40 < min((w - 1) * R, (j_stop - 1) / CH)) {
            // diagnostic: how many chunks of this segment block w-1 has already published when this item starts
            int ahead = 0;
            for (int kk = 0; kk < 32; ++kk) {
                const ulonglong2 t = ll_load(a.ll + (size_t)(j_begin / CH + kk) * 3 * CH);
                if ((unsigned)(t.x >> 32) != (a.epoch_tag | (unsigned)w)) break;
                ahead = kk + 1;
            }
            atomicAdd(reinterpret_cast<unsigned long long *>(a.ctl + kCtlSegAhead) + (ahead >= 32 ? 7 : ahead >= 16 ? 6 : ahead >= 8 ? 5 : ahead >= 4 ? 4 : ahead), 1ull);
        }
        const int last_chunk_of_item = (j_stop - 1) / CH;

        int pf_chunk = -1;
        double pf_x = 0.0, pf_y = 0.0, pf_z = 0.0;
        ulonglong2 pf_ax = make_ulonglong2(0, 0), pf_ay = pf_ax, pf_az = pf_ax;
        const unsigned want_tag = a.epoch_tag | (unsigned)w;         // written by block w-1
        const unsigned my_tag = a.epoch_tag | (unsigned)(w + 1);
        auto prefetch_chunk = [&](int c) {
            const int g = c * CH + lane;
            pf_x = sx[g]; pf_y = sy[g]; pf_z = sz[g];
            if (c < w * R) {
                const ulonglong2 *q = ll + (size_t)c * 3 * CH + lane;
                pf_ax = ll_load(q); pf_ay = ll_load(q + CH); pf_az = ll_load(q + 2 * CH);
            }
            pf_chunk = c;
        };
        auto open_chunk = [&](int c) {
            const int slot = (c % 3) * CH + lane;
            const bool lower = c < w * R;
            const long long a0 = now();
            if (pf_chunk != c) prefetch_chunk(c);
            if (lower && !__all_sync(FULL, ll_ok(pf_ax, want_tag) && ll_ok(pf_ay, want_tag) && ll_ok(pf_az, want_tag))) {
                if (PROF) { ++n_polled; if (lane == 0) atomicAdd(reinterpret_cast<unsigned long long *>(a.ctl + kCtlSegPolledQuarter) + min(3, 4 * w / a.T), 1ull); }
                // caught up with block w-1: fall a.hyst chunks behind first (a later chunk of the same segment that block
                // w-1 streams, not one of its own), so that the next prefetches find their words
                const int ahead = min(min(c + a.hyst, last_chunk_of_item), (w - 1) * R - 1);
                if (a.hyst > 0 && ahead > c) seg_poll(a, ll + (size_t)ahead * 3 * CH + 2 * CH + lane, want_tag, dead);
                // the three words of this lane's entry together, inline: ONE round trip per attempt (a block that has caught up
                // with the one above pays this at every chunk -- three polls one after the other cost three round trips)
                const ulonglong2 *q = ll + (size_t)c * 3 * CH + lane;
                const long long t0 = clock64();
                while (!dead) {
                    pf_ax = ll_load(q); pf_ay = ll_load(q + CH); pf_az = ll_load(q + 2 * CH);
                    if (__all_sync(FULL, ll_ok(pf_ax, want_tag) && ll_ok(pf_ay, want_tag) && ll_ok(pf_az, want_tag))) break;
                    if (clock64() - t0 > a.spin_limit || *reinterpret_cast<volatile int *>(a.ctl + 2) != 0)
                        pf_ax = seg_poll(a, q, want_tag, dead);   // out of line: gives up, records what never came, raises the abort flag
                }
            }
            if (PROF) { t_wait += now() - a0; ++n_chunks; }   // the tag check (it also waits for the prefetch loads) and the polls
            rx[slot] = pf_x; ry[slot] = pf_y; rz[slot] = pf_z;
            if (lower) { rax[slot] = ll_unpack(pf_ax); ray[slot] = ll_unpack(pf_ay); raz[slot] = ll_unpack(pf_az); }
            __syncwarp();
        };
        // block w+1 reads what this block publishes: here, or on the GPU of the next turn; the last block's words are the sums
        ulonglong2 *const pub = (w % a.deal == a.deal - 1 && w != a.T - 1) ? a.ll_down : ll;
        auto check_range = [&]() {   // raised everywhere BEFORE the words computed from such an operand are published
            if (__any_sync(FULL, key >= kRangeLimit)) { seg_raise(a, a.range_slot); key = 0; }
        };
        auto close_chunk = [&](int c) {
            const int slot = (c % 3) * CH + lane;
            __syncwarp();
            check_range();
            ulonglong2 *q = pub + (size_t)c * 3 * CH + lane;
            ll_store(q, ll_pack(rax[slot], my_tag));
            ll_store(q + CH, ll_pack(ray[slot], my_tag));
            ll_store(q + 2 * CH, ll_pack(raz[slot], my_tag));
        };
        // U predicated steps starting at step s0: the skew's fill and drain at both ends of the item and the
        // block's own diagonal.  Branch-free: everything is evaluated, the predicates select what is kept.
        auto edge_group = [&](int s0) {
            int es[U];
            double qx[R * U], qy[R * U], qz[R * U], jx[R * U], jy[R * U], jz[R * U];
            double vx[R * U], vy[R * U], vz[R * U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                es[u] = e;
                e = (e + 1 == RING) ? 0 : e + 1;
                const int j = s0 + u - lane;
                const bool act = j >= j_begin && j < j_stop;
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
                const bool act = j >= j_begin && j < j_stop;
                bool have_u = act && j < tgt0;   // an earlier lane (or block) already carries this chain
                double ux = 0.0, uy = 0.0, uz = 0.0;
                if (have_u) { ux = rax[es[u]]; uy = ray[es[u]]; uz = raz[es[u]]; }
#pragma unroll
                for (int r = 0; r < R; ++r) {
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
        // U full steps: every lane's source lies inside the item and below the block, no index predicates
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
        // ---- the same work as fast_group, split so that it can be pipelined: gather + forces of one group, and its four steps
        auto gather = [&](int (&es)[U], double (&jx)[R * U], double (&jy)[R * U], double (&jz)[R * U]) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                es[u] = e;
                e = (e + 1 == RING) ? 0 : e + 1;
                const double xj = rx[es[u]], yj = ry[es[u]], zj = rz[es[u]];
#pragma unroll
                for (int r = 0; r < R; ++r) { jx[u * R + r] = xj; jy[u * R + r] = yj; jz[u * R + r] = zj; }
            }
        };
        auto rmw_apply = [&](auto allv, const int (&es)[U], int u, const double (&vx)[R * U], const double (&vy)[R * U],
                             const double (&vz)[R * U], double ux, double uy, double uz) {
            constexpr bool ALLV = decltype(allv)::value;
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
        };
        auto forces_only = [&](int (&es)[U], double (&vx)[R * U], double (&vy)[R * U], double (&vz)[R * U]) {
            double qx[R * U], qy[R * U], qz[R * U], jx[R * U], jy[R * U], jz[R * U];
            gather(es, jx, jy, jz);
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int r = 0; r < R; ++r) { qx[u * R + r] = px[r]; qy[u * R + r] = py[r]; qz[u * R + r] = pz[r]; }
            pair_forces<R * U>(qx, qy, qz, jx, jy, jz, vx, vy, vz, key);
        };
        auto rmw_all = [&](auto allv, const int (&es)[U], const double (&vx)[R * U], const double (&vy)[R * U], const double (&vz)[R * U]) {
#pragma unroll
            for (int u = 0; u < U; ++u) rmw_apply(allv, es, u, vx, vy, vz, rax[es[u]], ray[es[u]], raz[es[u]]);
        };
        // forces of the NEW group (esn, vn) with the four steps of the PREVIOUS one (esp, vp) woven in
        auto fused = [&](auto allv, int (&esn)[U], double (&vnx)[R * U], double (&vny)[R * U], double (&vnz)[R * U],
                         const int (&esp)[U], const double (&vpx)[R * U], const double (&vpy)[R * U], const double (&vpz)[R * U]) {
            double qx[R * U], qy[R * U], qz[R * U], jx[R * U], jy[R * U], jz[R * U];
            gather(esn, jx, jy, jz);
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int r = 0; r < R; ++r) { qx[u * R + r] = px[r]; qy[u * R + r] = py[r]; qz[u * R + r] = pz[r]; }
            double ux = rax[esp[0]], uy = ray[esp[0]], uz = raz[esp[0]];
            pair_forces_hooked<R * U>(qx, qy, qz, jx, jy, jz, vnx, vny, vnz, key, [&](auto st) {
                constexpr int S = decltype(st)::value;
                if constexpr (U == 2) {
                    if constexpr (S == 4) { rmw_apply(allv, esp, 0, vpx, vpy, vpz, ux, uy, uz); ux = rax[esp[1]]; uy = ray[esp[1]]; uz = raz[esp[1]]; }
                    if constexpr (S == 13) { rmw_apply(allv, esp, 1, vpx, vpy, vpz, ux, uy, uz); }
                    return;
                }
                if constexpr (S == 1) { rmw_apply(allv, esp, 0, vpx, vpy, vpz, ux, uy, uz); ux = rax[esp[1]]; uy = ray[esp[1]]; uz = raz[esp[1]]; }
                if constexpr (S == 6) { rmw_apply(allv, esp, 1, vpx, vpy, vpz, ux, uy, uz); ux = rax[esp[2]]; uy = ray[esp[2]]; uz = raz[esp[2]]; }
                if constexpr (S == 11) { rmw_apply(allv, esp, 2, vpx, vpy, vpz, ux, uy, uz); ux = rax[esp[3]]; uy = ray[esp[3]]; uz = raz[esp[3]]; }
                if constexpr (S == 16) { rmw_apply(allv, esp, 3, vpx, vpy, vpz, ux, uy, uz); }
            });
        };
        // one predicate-free run inside a chunk, pipelined: forces(0) | forces(g+1) + steps(g) ... | steps(last)
        auto piped_run = [&](auto allv, int lim, int &s_) {
            auto s_ref = [&]() -> int & { return s_; };
            int esA[U], esB[U];
            double vAx[R * U], vAy[R * U], vAz[R * U], vBx[R * U], vBy[R * U], vBz[R * U];
            forces_only(esA, vAx, vAy, vAz);
            s_ref() += U;
#pragma unroll 1
            while (s_ref() + 2 * U <= lim) {
                fused(allv, esB, vBx, vBy, vBz, esA, vAx, vAy, vAz);
                fused(allv, esA, vAx, vAy, vAz, esB, vBx, vBy, vBz);
                s_ref() += 2 * U;
            }
            if (s_ref() + U <= lim) {
                fused(allv, esB, vBx, vBy, vBz, esA, vAx, vAy, vAz);
                s_ref() += U;
                rmw_all(allv, esB, vBx, vBy, vBz);
            } else {
                rmw_all(allv, esA, vAx, vAy, vAz);
            }
        };
        const bool full = i0 + BS <= n;

        int s = j_begin;
        bool early_closed = false;
        const int s_end = j_stop - 1 + 31, first_chunk = j_begin / CH, last_chunk = (j_stop - 1) / CH;
        const int fast_from = j_begin + 31, fast_to = min(i0, j_stop);   // steps whose 32 lanes all hold a lower source of the item
#pragma unroll 1
        while (s <= s_end) {
            if (s % CH == 0 && s / CH <= last_chunk) {
                const int c = s / CH;
                if (c - 2 >= first_chunk) close_chunk(c - 2);   // lane 31 left it 2 steps ago at the latest
                open_chunk(c);
                if (c + 1 <= last_chunk) prefetch_chunk(c + 1);
            } else if (s == (last_chunk + 1) * CH && last_chunk - 1 >= first_chunk && !early_closed) {
                close_chunk(last_chunk - 1);                    // the same rule one chunk further (the drain / the diagonal follow)
                early_closed = true;
            }
            const int lim = min(fast_to, (s / CH + 1) * CH);    // end of the predicate-free run inside this chunk
            const long long b0 = now();
            if (s >= fast_from && s + U <= lim) {
                if constexpr (PIPE) {
                    if (full) piped_run(std::true_type{}, lim, s);
                    else piped_run(std::false_type{}, lim, s);
                } else if (full) {
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
        if (last_chunk - 1 >= first_chunk && !early_closed) close_chunk(last_chunk - 1);
        close_chunk(last_chunk);
        if (!last_item) {   // hand the block's row sums to item (w, k+1)
            check_range();
            ulonglong2 *q = a.racc + (size_t)w * (3 * R * 32) + lane;
            const unsigned tag = a.epoch_tag | (unsigned)(k + 1);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                ll_store(q + (3 * r + 0) * 32, ll_pack(ax[r], tag));
                ll_store(q + (3 * r + 1) * 32, ll_pack(ay[r], tag));
                ll_store(q + (3 * r + 2) * 32, ll_pack(az[r], tag));
            }
        }
        if (PROF) { t_item += now() - it0; ++n_items; }
    }
    if (PROF && lane == 0) {
        unsigned long long *cnt = reinterpret_cast<unsigned long long *>(a.ctl + kCtlSpans);
        atomicAdd(cnt + 0, (unsigned long long)t_wait); atomicAdd(cnt + 1, (unsigned long long)t_racc);
        atomicAdd(reinterpret_cast<unsigned long long *>(a.ctl + kCtlSegPolled), (unsigned long long)n_polled);
        atomicAdd(cnt + 2, (unsigned long long)t_item); atomicAdd(cnt + 3, (unsigned long long)t_fast);
        atomicAdd(cnt + 4, (unsigned long long)t_slow); atomicAdd(cnt + 5, (unsigned long long)n_items);
        atomicAdd(cnt + 6, (unsigned long long)n_chunks);
    }
    if (__any_sync(FULL, key >= kRangeLimit)) seg_raise(a, a.range_slot);
}

}  // namespace hgb
