// nbody_tri.cuh -- pair-symmetric all-pairs FP64 step for sm_100a, second generation: the force of
// every unordered pair is evaluated once (EXACT mode, bit-identical to accel(), reference
// src/nbody/nbody-run.c:17-25) and the ORDERED accumulation is decoupled from the force evaluation,
// so the wavefront's critical path carries 3 adds per pair instead of the whole 32-instruction
// chain, rows can be cut into segments (no long tail), and the work can be dealt over several GPUs.
//
// Pair (i, j), j < i, contributes v = d * f (pair_wise_accel, reference src/nbody/nbody-run.c:6-15,
// evaluated from target i's side) to acc_i at position j of i's sum and -v to acc_j at position i of
// j's sum -- bit for bit what the reference adds there (d(j->i) = -d(i->j) exactly, same s, t, f).
//
//   * ROW I = the 32 targets 32I .. 32I+31, one per lane of a warp, position and accumulator in
//     registers.  Its lower sources j < 32I are cut into SEGMENTS of `seg` sources; ITEM (I, k) =
//     row I against segment k (the last item of a row also does the row's own 32 x 32 triangle).
//     An item is the unit of scheduling: a warp takes a ticket, runs one item, takes the next ticket.
//   * MINITILE = 32 targets x 8 consecutive sources.  Each lane evaluates its 8 pairs (8 interleaved
//     dependency chains, 208 FP64 instructions per warp), adds them to its own accumulator in source
//     order (ROW pass: 24 instructions), and writes the 8 x 3 values into a shared-memory transposed
//     buffer.  COLUMN pass: lane L < 24 owns (component L / 8, source L % 8) of the minitile, takes
//     that source's running sum ("chain": the accumulator of body j as a TARGET, which has already
//     absorbed every target below this row), subtracts the 32 targets' values in target order (32
//     dependent instructions with 24 of 32 lanes busy: 264 instead of the ideal 256 per minitile)
//     and publishes the result for the next row.  The column pass of minitile m is issued INSIDE the
//     force evaluation of minitile m+1 (one basic block, independent instruction streams), so its
//     latency hides under the throughput-bound part.  (Measured alternative: the chain packed into the
//     first stages and published mid-block, chain words prefetched one minitile ahead instead of two --
//     a shorter stagger between rows, but 8-12 % less throughput on one GPU and no gain on 4 / 8 GPUs,
//     profiles/r2m_tri_multi_early_publish_g8_g4.txt.)
//   * Hand-off between rows: tagged words (payload + tag in every 8-byte half: one 16-byte store
//     publishes value and flag, no fence; the "LL" scheme of collective libraries).  Row I expects
//     tag (epoch | I), written by row I-1, and publishes with (epoch | I+1) -- into the local buffer,
//     or into the buffer of the GPU that runs row I+1 (peer-mapped pointer: the store crosses NVLink).
//     The words of the last row are the finished accumulators: every GPU's finish kernel reads them from
//     the buffer of the GPU that ran the last row (peer-mapped loads), so no position exchange follows.
//   * Hand-off between the items of one row: the row's 32 x 3 accumulators through a tagged buffer.
//   * Order.  Items are grouped in TILES (b, k) = the `tr` = seg/32 rows of super-row b against segment
//     k, k <= b.  Item (I, k) needs (I, k-1) finished (row accumulators) and streams the chain words of
//     (I-1, k) minitile by minitile; the host lists the tiles of each GPU in anti-diagonal order
//     (b + k ascending), a linear extension of those dependencies, and warps take tickets in that
//     order, so a warp only ever waits for an item with an earlier ticket, which a resident warp is
//     already running (persistent grid, no deadlock); super-row b lives on GPU b mod g.
// tests/tri_model.py is a CPU model of exactly this index logic.
#pragma once
#include "nbody_sym.cuh"

namespace hgb {

constexpr int kTriWarps = 4;          // warps (independent workers) per CTA
constexpr int kTriCtlInts = 64;       // ctl[0] ticket, ctl[2] abort flag, ctl[4], ctl[5] range flag of even / odd steps, ctl[8..19] profiling, ctl[3], [6], [7], [24..26] what an abandoned poll waited for
constexpr int kTbPitch = 34;          // doubles per row of the transposed buffer (16-byte rows, conflict-free LDS.128)
constexpr int kTriWarpDoubles = 2 * 3 * 32 + 2 * 24 * kTbPitch + 32;   // ring + 2 transposed buffers + chain seeds
constexpr size_t kTriSmem = (size_t)kTriWarps * kTriWarpDoubles * sizeof(double);

struct TriArgs {
    const double *src;        // SoA positions, component stride npad, zero beyond n
    int npad, n;
    int T;                    // rows = ceil(n / 32)
    int seg, tr;              // sources per segment (multiple of 32), rows per tile (handed out as consecutive tickets)
    int deal;                 // rows per turn of a GPU (a multiple of tr): rows I with I / deal = t (mod GPUs) run on GPU t
    const int2 *tiles;        // this GPU's tiles (b, k) in ticket order
    int ntiles;
    ulonglong2 *ll;           // chain words this GPU reads: [chunk of 32 sources][x, y, z][32] x 16 B
    ulonglong2 *ll_down;      // where the LAST row of a super-row publishes (buffer of the next super-row's GPU)
    int npeers;
    ulonglong2 *racc;         // row accumulators between the items of a row: [row][x, y, z][32] x 16 B
    unsigned epoch_tag;       // (launch epoch) << 21
    int *ctl;
    int *ctl_all[kMaxPeers];  // peers' ctl (range / abort flags are raised everywhere)
    long long spin_limit;     // clock64 ticks a poll may last before the step is abandoned
    int range_slot;           // ctl index of this step's range flag (4 + step parity)
};

// d * f of pair_wise_accel for one target against 8 sources, fused with the COLUMN pass of the
// previous minitile: `col` points at this lane's row of that minitile's transposed buffer (32 values,
// target order) and `chain` is the running sum they are subtracted from.  One basic block: ptxas
// interleaves the dependent chain (LDS -> DSUB -> DSUB ...) with the eight independent force chains.
template <bool ALLV>
__device__ __forceinline__ void forces_and_column(double px, double py, double pz, const double (&xj)[8],
                                                  const double (&yj)[8], const double (&zj)[8],
                                                  double (&vx)[8], double (&vy)[8], double (&vz)[8],
                                                  unsigned &key, const double *col, double &chain, int cnt, bool act) {
    constexpr int P = 8;
    double t[P], y[P], g[P], s[P], a[P], e[P], p[P], q[P];
    int ct = 0;   // column pairs consumed so far (compile-time after unrolling)
    auto column = [&](int pairs) {
#pragma unroll
        for (int w = 0; w < pairs; ++w, ++ct) {
            double2 v = make_double2(0.0, 0.0);
            if (act) v = *reinterpret_cast<const double2 *>(col + 2 * ct);   // lanes 24..31 own no chain: no load, no shared-memory wavefront
            if (ALLV || 2 * ct < cnt) chain = __dsub_rn(chain, v.x);
            if (ALLV || 2 * ct + 1 < cnt) chain = __dsub_rn(chain, v.y);
        }
    };
#pragma unroll
    for (int u = 0; u < P; ++u) { vx[u] = __dsub_rn(xj[u], px); vy[u] = __dsub_rn(yj[u], py); vz[u] = __dsub_rn(zj[u], pz); }
    column(2);
#pragma unroll
    for (int u = 0; u < P; ++u) { s[u] = __dmul_rn(vx[u], vx[u]); a[u] = __dmul_rn(vy[u], vy[u]); e[u] = __dmul_rn(vz[u], vz[u]); }
    column(2);
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], a[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], e[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], kEps);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(s[u], s[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(t[u], s[u]);            // (s*s)*s
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) { y[u] = mufu_rsqrt(t[u]); key = max(key, range_key(t[u])); }
#pragma unroll
    for (int u = 0; u < P; ++u) a[u] = __dmul_rn(y[u], y[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(t[u], -a[u], 1.0);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) { p[u] = __fma_rn(e[u], 0.375, 0.5); q[u] = __dmul_rn(y[u], e[u]); }
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = __fma_rn(p[u], q[u], y[u]);       // ~1/sqrt(t)
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __dmul_rn(t[u], y[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(-g[u], g[u], t[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __fma_rn(e[u], half_of(y[u]), g[u]);  // RN(sqrt(t))
    column(1);
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
    for (int u = 0; u < P; ++u) { vx[u] = __dmul_rn(vx[u], y[u]); vy[u] = __dmul_rn(vy[u], y[u]); vz[u] = __dmul_rn(vz[u], y[u]); }
    static_assert(2 + 2 + 12 == 16, "16 column pairs = 32 targets");
}

// The same block with the ROW pass of the previous minitile fused in as well ("row pipelining"): `pvx/pvy/pvz`
// are that minitile's 8 x 3 values, added to the row's sums in source order (reference src/nbody/nbody-run.c:19-21),
// one source per stage of the force evaluation, so the 8-deep dependent add chains cost no issue time of their
// own; `after_row()` runs right after the last add (the cooperative-row kernel hands the sums on there).
template <bool ALLV, class AfterRow>
__device__ __forceinline__ void forces_column_row(double px, double py, double pz, const double (&xj)[8],
                                                  const double (&yj)[8], const double (&zj)[8],
                                                  double (&vx)[8], double (&vy)[8], double (&vz)[8],
                                                  unsigned &key, const double *col, double &chain, int cnt, bool act,
                                                  const double (&pvx)[8], const double (&pvy)[8], const double (&pvz)[8],
                                                  double &ax, double &ay, double &az, AfterRow after_row) {
    constexpr int P = 8;
    double t[P], y[P], g[P], s[P], a[P], e[P], p[P], q[P];
    int ct = 0, ru = 0;   // column pairs / row sources consumed so far (compile-time after unrolling)
    auto column = [&](int pairs) {
#pragma unroll
        for (int w = 0; w < pairs; ++w, ++ct) {
            double2 v = make_double2(0.0, 0.0);
            if (act) v = *reinterpret_cast<const double2 *>(col + 2 * ct);
            if (ALLV || 2 * ct < cnt) chain = __dsub_rn(chain, v.x);
            if (ALLV || 2 * ct + 1 < cnt) chain = __dsub_rn(chain, v.y);
        }
    };
    auto row = [&]() {
        ax = __dadd_rn(ax, pvx[ru]); ay = __dadd_rn(ay, pvy[ru]); az = __dadd_rn(az, pvz[ru]);
        ++ru;
    };
#pragma unroll
    for (int u = 0; u < P; ++u) { vx[u] = __dsub_rn(xj[u], px); vy[u] = __dsub_rn(yj[u], py); vz[u] = __dsub_rn(zj[u], pz); }
    column(2); row();
#pragma unroll
    for (int u = 0; u < P; ++u) { s[u] = __dmul_rn(vx[u], vx[u]); a[u] = __dmul_rn(vy[u], vy[u]); e[u] = __dmul_rn(vz[u], vz[u]); }
    column(2); row();
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], a[u]);
    column(1); row();
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], e[u]);
    column(1); row();
#pragma unroll
    for (int u = 0; u < P; ++u) s[u] = __dadd_rn(s[u], kEps);
    column(1); row();
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(s[u], s[u]);
    column(1); row();
#pragma unroll
    for (int u = 0; u < P; ++u) t[u] = __dmul_rn(t[u], s[u]);            // (s*s)*s
    column(1); row();
#pragma unroll
    for (int u = 0; u < P; ++u) { y[u] = mufu_rsqrt(t[u]); key = max(key, range_key(t[u])); }
#pragma unroll
    for (int u = 0; u < P; ++u) a[u] = __dmul_rn(y[u], y[u]);
    column(1); row();
    after_row();
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(t[u], -a[u], 1.0);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) { p[u] = __fma_rn(e[u], 0.375, 0.5); q[u] = __dmul_rn(y[u], e[u]); }
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) y[u] = __fma_rn(p[u], q[u], y[u]);       // ~1/sqrt(t)
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __dmul_rn(t[u], y[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) e[u] = __fma_rn(-g[u], g[u], t[u]);
    column(1);
#pragma unroll
    for (int u = 0; u < P; ++u) g[u] = __fma_rn(e[u], half_of(y[u]), g[u]);  // RN(sqrt(t))
    column(1);
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
    for (int u = 0; u < P; ++u) { vx[u] = __dmul_rn(vx[u], y[u]); vy[u] = __dmul_rn(vy[u], y[u]); vz[u] = __dmul_rn(vz[u], y[u]); }
}

// ---- cold paths, out of line: the hot loop must stay small (a 2^17-body profile of the first version showed
// one "no instruction" stall per issued instruction: the unrolled cold code had pushed the loop out of the
// instruction cache) ------------------------------------------------------------------------------------
// raise flag ctl[which] (1: operand out of range, 2: step abandoned) on every GPU
__device__ __noinline__ void tri_raise(const TriArgs &a, int which) {
    for (int p = 0; p < a.npeers; ++p) *reinterpret_cast<volatile int *>(a.ctl_all[p] + which) = 1;
    __threadfence_system();
}
// poll one tagged word until its tag shows (every lane of the warp calls this together and all leave
// together); a poll that outlasts the spin limit raises the abort flag everywhere, after which nobody waits
// any more (results are garbage, the host reports the failure).  Returns the clock ticks spent.
__device__ __noinline__ ulonglong2 tri_poll(const TriArgs &a, const ulonglong2 *q, unsigned tag, int &dead,
                                            unsigned ns, long long &spent) {
    const long long t0 = clock64();
    ulonglong2 w = make_ulonglong2(0, 0);
    while (!dead) {
        w = ll_load(q);
        if (__all_sync(0xffffffffu, ll_ok(w, tag))) break;
        bool g = *reinterpret_cast<volatile int *>(a.ctl + 2) != 0;
        if (!g && clock64() - t0 > a.spin_limit) {
            if ((threadIdx.x & 31) == 0 && atomicCAS(a.ctl + 3, 0, 1) == 0) {   // first to give up: say what never came
                a.ctl[6] = (int)tag;
                a.ctl[7] = (int)(unsigned)w.x; a.ctl[24] = (int)(w.x >> 32);
                const long long off = q - a.ll, roff = q - a.racc;
                a.ctl[25] = (off >= 0 && off < 3ll * a.npad) ? (int)off : -1;
                a.ctl[26] = (roff >= 0 && roff < 3ll * a.npad) ? (int)roff : -1;
            }
            tri_raise(a, 2); g = true;
        }
        dead = __any_sync(0xffffffffu, g);
        __nanosleep(ns);
    }
    spent += clock64() - t0;
    return w;
}

// CTAS resident CTAs per SM (register budget).  PROF: lane 0 of every warp accumulates clock64 spans
// (waits for chain words, waits for row accumulators, whole items, the diagonal triangles) in ctl[8..].
// ROWPIPE: the row pass of a minitile is fused into the force block of the NEXT one as well (forces_column_row).
template <int CTAS, bool PROF = false, bool ROWPIPE = false>
__global__ void __launch_bounds__(kTriWarps * 32, CTAS)
nbody_tri_kernel(const __grid_constant__ TriArgs a) {
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) double tri_smem[];
    const int lane = threadIdx.x & 31;
    double *ring = tri_smem + (threadIdx.x >> 5) * kTriWarpDoubles;   // [slot 2][comp 3][32]
    double *tb = ring + 2 * 3 * 32;                                   // [buffer 2][24][kTbPitch]
    double *seed = tb + 2 * 24 * kTbPitch;                            // [24] chain seeds of a diagonal minitile
    const double *sx = a.src, *sy = a.src + a.npad, *sz = a.src + 2 * (size_t)a.npad;
    const int col_lane = lane < 24 ? lane : 23;                       // lanes 24..31 shadow lane 23 and never publish
    const int cc = col_lane >> 3, cu = col_lane & 7;                  // component, source within the minitile
    unsigned key = 0;
    // only the out-of-line poll ever takes the address of these (they may live in local memory; the hot
    // loop never touches them)
    int dead = 0;                                                     // the step was abandoned (spin limit)
    long long t_chain = 0, t_racc = 0;
    long long t_item = 0, t_diag = 0, n_items = 0;
    auto now = [] { return PROF ? clock64() : 0ll; };

    for (;;) {
        int ticket = 0;
        if (lane == 0) ticket = atomicAdd(&a.ctl[0], 1);
        ticket = __shfl_sync(FULL, ticket, 0);
        const int ti = ticket / a.tr, r = ticket - ti * a.tr;
        if (ti >= a.ntiles) break;
        const int2 bk = a.tiles[ti];
        const int I = bk.x * a.tr + r, k = bk.y;
        if (I >= a.T) continue;
        const int i0 = 32 * I, K = i0 / a.seg;            // K = the row's last item
        if (k > K) continue;                              // this row ends before segment k
        const long long it0 = now();
        const bool last_item = k == K;
        const int j_begin = k * a.seg, j_end = last_item ? i0 : (k + 1) * a.seg;
        const int cnt = min(32, a.n - i0);                // valid targets of the row (32 except in the last row)
        const bool final_row = I == a.T - 1;
        const unsigned want_tag = a.epoch_tag | (unsigned)I;          // written by row I-1
        const unsigned my_tag = a.epoch_tag | (unsigned)(I + 1);
        // row I+1 is the next tile-row's first row; the last row of the step keeps its words (the finished sums) here
        ulonglong2 *pub = (I % a.deal == a.deal - 1 && !final_row) ? a.ll_down : a.ll;

        const double px = sx[i0 + lane], py = sy[i0 + lane], pz = sz[i0 + lane];
        double ax = 0.0, ay = 0.0, az = 0.0;
        if (k > 0) {   // the row's sums so far, left by item (I, k-1)
            const ulonglong2 *q = a.racc + (size_t)I * 96 + lane;
            const unsigned tag = a.epoch_tag | (unsigned)k;
            ax = ll_unpack(tri_poll(a, q, tag, dead, 100, t_racc));
            ay = ll_unpack(tri_poll(a, q + 32, tag, dead, 100, t_racc));
            az = ll_unpack(tri_poll(a, q + 64, tag, dead, 100, t_racc));
        }

        // An operand left the range of the branch-free IEEE sequences (never for sane finite positions): the
        // flag is raised on every GPU BEFORE the last words computed from it are published, so whoever has
        // seen all final words of the step also sees the flag, and the ordinary kernel redoes the step.
        auto check_range = [&]() {
            if (__any_sync(FULL, key >= kRangeLimit)) { tri_raise(a, a.range_slot); key = 0; }
        };
        auto publish = [&](ulonglong2 *to, double chain) {
            if (lane < 24) ll_store(to, ll_pack(chain, my_tag));
        };
        auto wait_word = [&](ulonglong2 w, const ulonglong2 *from) -> double {
            if (!__all_sync(FULL, ll_ok(w, want_tag))) w = tri_poll(a, from, want_tag, dead, 40, t_chain);
            return ll_unpack(w);
        };

        auto lower = [&](auto allv) {
            constexpr bool ALLV = decltype(allv)::value;
            const int nchunks = (j_end - j_begin) >> 5, M = 4 * nchunks;
            if (nchunks <= 0) return;
            // Chain words of this item for this column lane: word of minitile m at wp0 + (m / 4) * 96 + (m % 4) * 8.
            // The loop below walks them with one pointer and a step that alternates 16 / 80.
            const ulonglong2 *wp = a.ll + (size_t)(j_begin >> 5) * 96 + cc * 32 + cu;   // word of minitile 0
            const ptrdiff_t to_pub = pub - a.ll;
            // positions: chunk 0 straight into ring slot 0, chunk ch+1 into the other slot at the start of
            // chunk ch (from registers loaded one chunk earlier), so that the sources of the NEXT minitile can
            // always be fetched from shared memory before the current one is finished
            ring[lane] = sx[j_begin + lane]; ring[64 + lane] = sy[j_begin + lane]; ring[128 + lane] = sz[j_begin + lane];
            double fx = 0.0, fy = 0.0, fz = 0.0;                       // positions of chunk ch+1
            if (nchunks > 1) { const int g = j_begin + 32 + lane; fx = sx[g]; fy = sy[g]; fz = sz[g]; }
            // chain words in flight, two minitiles ahead of their use: `we` of the even minitiles, `wo` of the
            // odd ones (two registers that are never copied: a copy would wait for the load)
            ulonglong2 we = ll_load(wp), wo = ll_load(wp + 8);
            __syncwarp();
            double xj[8], yj[8], zj[8];
            auto fetch_sources = [&](int m) {   // the 8 sources of minitile m: ring[x, y, z][slot (m / 4) % 2][32]
                const double *from = ring + (m & 7) * 8;
#pragma unroll
                for (int u = 0; u < 8; u += 2) {
                    const double2 x2 = *reinterpret_cast<const double2 *>(from + u);
                    const double2 y2 = *reinterpret_cast<const double2 *>(from + 64 + u);
                    const double2 z2 = *reinterpret_cast<const double2 *>(from + 128 + u);
                    xj[u] = x2.x; xj[u + 1] = x2.y; yj[u] = y2.x; yj[u + 1] = y2.y; zj[u] = z2.x; zj[u + 1] = z2.y;
                }
            };
            auto next_chunk = [&](int ch) {   // start of chunk ch: chunk ch+1 into the other ring slot (last read during chunk ch-1)
                if (ch + 1 < nchunks) {
                    double *other = ring + ((ch + 1) & 1) * 32 + lane;
                    other[0] = fx; other[64] = fy; other[128] = fz;
                    if (ch + 2 < nchunks) { const int g = j_begin + 32 * (ch + 2) + lane; fx = sx[g]; fy = sy[g]; fz = sz[g]; }
                }
            };
            auto row_pass = [&](double *mine, const double (&vx)[8], const double (&vy)[8], const double (&vz)[8]) {
#pragma unroll
                for (int u = 0; u < 8; ++u) {   // sources ascending, the target's order
                    ax = __dadd_rn(ax, vx[u]); ay = __dadd_rn(ay, vy[u]); az = __dadd_rn(az, vz[u]);
                    mine[u * kTbPitch] = vx[u]; mine[(8 + u) * kTbPitch] = vy[u]; mine[(16 + u) * kTbPitch] = vz[u];
                }
            };
            double *const tb0 = tb + lane, *const tb1 = tb + 24 * kTbPitch + lane;              // where this lane writes
            const double *const col0 = tb + col_lane * kTbPitch, *const col1 = col0 + 24 * kTbPitch;   // ... and reads
            // minitile m (parity known at compile time): forces + row pass, fused with the column pass of minitile
            // m-1, whose word (at `at`) is in `w`; the register is refilled with the word of minitile m+1 (at
            // `refill`, or null) and the sources of minitile m+1 are fetched under the row pass (FETCH)
            auto minitile = [&](int m, ulonglong2 &w, const ulonglong2 *at, const ulonglong2 *refill,
                                const double *col_prev, double *mine, auto fetch) {
                double chain = wait_word(w, at);
                if (refill) w = ll_load(refill);
                double vx[8], vy[8], vz[8];
                forces_and_column<ALLV>(px, py, pz, xj, yj, zj, vx, vy, vz, key, col_prev, chain, cnt, lane < 24);
                if (decltype(fetch)::value) fetch_sources(m + 1);
                publish(const_cast<ulonglong2 *>(at) + to_pub, chain);   // before the row pass: the next row gets it that much earlier
                row_pass(mine, vx, vy, vz);
                __syncwarp();   // column lanes read these values during the next minitile; ring slot hand-over
            };
            if constexpr (ROWPIPE) {
                // two sets of 8 x 3 values, alternating: a minitile's values are added to the row's sums during the next one
                double vxa[8], vya[8], vza[8], vxb[8], vyb[8], vzb[8];
                auto store_tb = [&](double *mine, const double (&vx)[8], const double (&vy)[8], const double (&vz)[8]) {
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        mine[u * kTbPitch] = vx[u]; mine[(8 + u) * kTbPitch] = vy[u]; mine[(16 + u) * kTbPitch] = vz[u];
                    }
                };
                auto minitile_p = [&](int m, ulonglong2 &w, const ulonglong2 *at, const ulonglong2 *refill, const double *col_prev,
                                      double *mine, auto fetch, const double (&ox)[8], const double (&oy)[8], const double (&oz)[8],
                                      double (&nx)[8], double (&ny)[8], double (&nz)[8]) {
                    double chain = wait_word(w, at);
                    if (refill) w = ll_load(refill);
                    forces_column_row<ALLV>(px, py, pz, xj, yj, zj, nx, ny, nz, key, col_prev, chain, cnt, lane < 24,
                                            ox, oy, oz, ax, ay, az, [] {});
                    if (decltype(fetch)::value) fetch_sources(m + 1);
                    store_tb(mine, nx, ny, nz);
                    publish(const_cast<ulonglong2 *>(at) + to_pub, chain);
                    __syncwarp();
                };
                {   // minitile 0
                    fetch_sources(0);
                    next_chunk(0);
                    double qx[8], qy[8], qz[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u) { qx[u] = px; qy[u] = py; qz[u] = pz; }
                    pair_forces<8>(qx, qy, qz, xj, yj, zj, vxa, vya, vza, key);
                    fetch_sources(1);
                    store_tb(tb0, vxa, vya, vza);
                    __syncwarp();
                }
                int step = 16;
#pragma unroll 1
                for (int m = 1; m < M - 2; m += 2) {
                    check_range();
                    minitile_p(m, we, wp, wp + step, col0, tb1, std::true_type{}, vxa, vya, vza, vxb, vyb, vzb);
                    if (step == 80) next_chunk((m + 1) >> 2);
                    minitile_p(m + 1, wo, wp + 8, wp + step + 8, col1, tb0, std::true_type{}, vxb, vyb, vzb, vxa, vya, vza);
                    wp += step;
                    step = 96 - step;
                }
                check_range();
                minitile_p(M - 1, we, wp, nullptr, col0, tb1, std::false_type{}, vxa, vya, vza, vxb, vyb, vzb);
                double chain = wait_word(wo, wp + 8);
#pragma unroll
                for (int tcol = 0; tcol < 16; ++tcol) {   // row pass and column pass of the segment's last minitile
                    const double2 v = *reinterpret_cast<const double2 *>(col1 + 2 * tcol);
                    if (ALLV || 2 * tcol < cnt) chain = __dsub_rn(chain, v.x);
                    if (ALLV || 2 * tcol + 1 < cnt) chain = __dsub_rn(chain, v.y);
                    if (tcol < 8) { ax = __dadd_rn(ax, vxb[tcol]); ay = __dadd_rn(ay, vyb[tcol]); az = __dadd_rn(az, vzb[tcol]); }
                }
                check_range();
                publish(const_cast<ulonglong2 *>(wp) + 8 + to_pub, chain);
                __syncwarp();
                return;
            }
            {   // minitile 0: nothing before it to do a column pass for
                fetch_sources(0);
                next_chunk(0);
                double qx[8], qy[8], qz[8], vx[8], vy[8], vz[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) { qx[u] = px; qy[u] = py; qz[u] = pz; }
                pair_forces<8>(qx, qy, qz, xj, yj, zj, vx, vy, vz, key);
                fetch_sources(1);
                row_pass(tb0, vx, vy, vz);
                __syncwarp();
            }
            int step = 16;   // words from minitile m-1 (even) to m+1: 16 inside a chunk, 80 across the chunk boundary
#pragma unroll 1
            for (int m = 1; m < M - 2; m += 2) {   // pairs (odd, even) = minitiles 1 .. M-2; wp = word of minitile m-1
                check_range();
                minitile(m, we, wp, wp + step, col0, tb1, std::true_type{});
                if (step == 80) next_chunk((m + 1) >> 2);
                minitile(m + 1, wo, wp + 8, wp + step + 8, col1, tb0, std::true_type{});
                wp += step;
                step = 96 - step;
            }
            check_range();
            minitile(M - 1, we, wp, nullptr, col0, tb1, std::false_type{});
            // column pass of the segment's last minitile (M-1 is odd: buffer 1, word in `wo`)
            double chain = wait_word(wo, wp + 8);
#pragma unroll
            for (int tcol = 0; tcol < 16; ++tcol) {
                const double2 v = *reinterpret_cast<const double2 *>(col1 + 2 * tcol);
                if (ALLV || 2 * tcol < cnt) chain = __dsub_rn(chain, v.x);
                if (ALLV || 2 * tcol + 1 < cnt) chain = __dsub_rn(chain, v.y);
            }
            check_range();
            publish(const_cast<ulonglong2 *>(wp) + 8 + to_pub, chain);
            __syncwarp();
        };
        if (cnt == 32) lower(std::true_type{});
        else lower(std::false_type{});

        if (!last_item) {   // hand the row's sums to item (I, k+1)
            ulonglong2 *q = a.racc + (size_t)I * 96 + lane;
            const unsigned tag = a.epoch_tag | (unsigned)(k + 1);
            ll_store(q, ll_pack(ax, tag)); ll_store(q + 32, ll_pack(ay, tag)); ll_store(q + 64, ll_pack(az, tag));
        } else {
            // the row's own 32 x 32 triangle: sources j = i0 + 8 mt + u against targets i > j.  The sources are
            // the warp's own targets (positions by shuffle).  After the row pass of minitile mt the lanes
            // 8 mt .. 8 mt + 7 hold their complete lower sums: those seed the chains of the minitile's sources.
            const long long d0 = now();
#pragma unroll 1
            for (int mt = 0; mt < 4; ++mt) {
                double qx[8], qy[8], qz[8], jx[8], jy[8], jz[8], vx[8], vy[8], vz[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    jx[u] = __shfl_sync(FULL, px, mt * 8 + u); jy[u] = __shfl_sync(FULL, py, mt * 8 + u);
                    jz[u] = __shfl_sync(FULL, pz, mt * 8 + u);
                    qx[u] = px; qy[u] = py; qz[u] = pz;
                }
                pair_forces<8>(qx, qy, qz, jx, jy, jz, vx, vy, vz, key);
                double *mine = tb + lane;
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    if (mt * 8 + u < lane) {   // j < i
                        ax = __dadd_rn(ax, vx[u]); ay = __dadd_rn(ay, vy[u]); az = __dadd_rn(az, vz[u]);
                    }
                    mine[u * kTbPitch] = vx[u]; mine[(8 + u) * kTbPitch] = vy[u]; mine[(16 + u) * kTbPitch] = vz[u];
                }
                if ((lane >> 3) == mt) { seed[lane & 7] = ax; seed[8 + (lane & 7)] = ay; seed[16 + (lane & 7)] = az; }
                __syncwarp();
                double chain = seed[col_lane];
                const double *col = tb + col_lane * kTbPitch;
                for (int i = mt * 8 + cu + 1; i < cnt; ++i) chain = __dsub_rn(chain, col[i]);
                check_range();
                publish(pub + (size_t)I * 96 + cc * 32 + mt * 8 + cu, chain);
                __syncwarp();
            }
            if (PROF) t_diag += now() - d0;
        }
        if (PROF) { t_item += now() - it0; ++n_items; }
    }
    if (PROF && lane == 0) {
        unsigned long long *c = reinterpret_cast<unsigned long long *>(a.ctl + 8);
        atomicAdd(c + 0, (unsigned long long)t_chain); atomicAdd(c + 1, (unsigned long long)t_racc);
        atomicAdd(c + 2, (unsigned long long)t_item); atomicAdd(c + 3, (unsigned long long)t_diag);
        atomicAdd(c + 4, (unsigned long long)n_items);
    }
    if (__any_sync(FULL, key >= kRangeLimit)) tri_raise(a, a.range_slot);
}

// buffer[i] = G * acc[i] (reference src/nbody/nbody-run.c:22-24).  acc = what the LAST row published (tag
// epoch | T) into this GPU's buffer -- possibly from another GPU, whose kernel may still be running, so
// the words are polled; a poll that outlasts the spin limit raises the abort flag.
__global__ void __launch_bounds__(256)
nbody_tri_finish_kernel(const ulonglong2 *__restrict__ ll, int npad, int n, unsigned final_tag, double *out,
                        int *ctl, long long spin_limit) {
    // a small persistent grid (grid-stride): while it polls it must not crowd out a wavefront kernel
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const ulonglong2 *q = ll + (size_t)(i >> 5) * 96 + (i & 31);
        ulonglong2 wx = ll_load(q), wy = ll_load(q + 32), wz = ll_load(q + 64);
        const long long t0 = clock64();
        while (!(ll_ok(wx, final_tag) && ll_ok(wy, final_tag) && ll_ok(wz, final_tag))) {
            if (*reinterpret_cast<volatile int *>(ctl + 2) != 0) break;
            if (clock64() - t0 > spin_limit) { *reinterpret_cast<volatile int *>(ctl + 2) = 1; break; }
            __nanosleep(200);
            wx = ll_load(q); wy = ll_load(q + 32); wz = ll_load(q + 64);
        }
        out[i] = __dmul_rn(kG, ll_unpack(wx));
        out[(size_t)npad + i] = __dmul_rn(kG, ll_unpack(wy));
        out[2 * (size_t)npad + i] = __dmul_rn(kG, ll_unpack(wz));
    }
}

}  // namespace hgb
