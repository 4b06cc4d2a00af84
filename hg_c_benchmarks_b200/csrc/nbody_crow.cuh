// nbody_crow.cuh -- pair-symmetric all-pairs FP64 step for sm_100a, third generation: COOPERATIVE ROWS.
//
// Same arithmetic, same hand-off buffers, same ticket plan and same multi-GPU dealing as the second
// generation (nbody_tri.cuh: every unordered pair evaluated once, EXACT mode, bit-identical to accel(),
// reference src/nbody/nbody-run.c:17-25) -- what changes is WHO runs an item.  There one warp walked the
// N/8 minitiles of a row one after another, so the last row alone took N/8 minitile times (111 ms at
// N = 2^20, more than half an 8-GPU step) and consecutive rows trailed each other by 2-3 of those
// minitile times: two serial chains that no number of GPUs shortens.
//
// Here an item (row I x segment k) belongs to a whole CTA of W warps, one per SM sub-partition:
//   * minitile m (32 targets x 8 sources) of the item goes to warp m mod W.  Every warp holds the SAME 32
//     targets and evaluates the forces of its own minitiles -- the 208 order-free instructions -- in
//     parallel with the other warps;
//   * the ORDERED part of the row, 8 x 3 adds per minitile (reference src/nbody/nbody-run.c:19-21, sources
//     ascending), ripples through the warps: warp q takes the row's sums after minitile m-1 from warp q-1
//     (a "token": 3 x 32 tagged 16-byte words in shared memory, payload + tag in each 8-byte half, written
//     and polled lane to lane -- no barrier, no fence), adds its 8 sources in order and hands the sums to
//     warp q+1.  The ripple costs 8 dependent DADDs + one shared-memory round trip per minitile (~150
//     cycles) where the force block costs >= 528 FP64-pipe cycles, so W = 4 warps keep a row moving at W
//     times the speed of one warp and stay throughput-bound;
//   * the column pass (every source's own running sum minus the 32 targets' values, target order, chain
//     words between rows) is unchanged: each warp does it for its own minitiles, fused into the force block
//     of its next one; rows still follow each other through the tagged chain words in L2 / over NVLink, but
//     a "minitile time" of a row is now 1/W of a warp's, so the stagger between rows shrinks with it.
// Slot reuse is safe without extra synchronisation for W >= 2: warp q overwrites its token (minitile m)
// when it finishes minitile m+W, whose sums went through m+1 -- i.e. through the consumer of token m.
// Items are taken by CTA (one ticket, two __syncthreads per item); the row's own 32 x 32 triangle and the
// hand-off of the row's sums to the next item are done by the warp next in rotation after the last minitile.
#pragma once
#include "nbody_tri.cuh"

namespace hgb {

template <int W>
constexpr size_t crow_smem_bytes() {
    return (size_t)W * kTriWarpDoubles * sizeof(double) + (size_t)W * 96 * sizeof(double) + (size_t)W * 8 + 16;
}

// One attempt to see phase `parity` of an mbarrier complete; the warp is suspended in hardware for up to `ns`
// nanoseconds instead of re-issuing loads (a first version polled tagged words in shared memory: three spinning
// warps per CTA saturated the shared-memory pipe and every hand-off took ~1200 cycles).
__device__ __forceinline__ bool mbar_try(unsigned bar, unsigned parity, unsigned ns) {
    unsigned ok;
    asm volatile(
        "{\n.reg .pred P1;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2, %3;\n"
        "selp.u32 %0, 1, 0, P1;\n}"
        : "=r"(ok) : "r"(bar), "r"(parity), "r"(ns) : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_arrive_at(unsigned bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// wait for a token (all lanes of the warp together); the global abort flag and the spin limit are looked at
// every 64 attempts, so a lost hand-off ends the step instead of hanging the GPU
__device__ __noinline__ void crow_wait(const TriArgs &a, unsigned bar, unsigned parity, int &dead, long long &spent) {
    const long long t0 = clock64();
    unsigned spins = 0;
    while (!dead) {
        if (__all_sync(0xffffffffu, mbar_try(bar, parity, 1000))) break;
        if ((++spins & 63u) == 0) {
            bool g = *reinterpret_cast<volatile int *>(a.ctl + 2) != 0;
            if (!g && clock64() - t0 > a.spin_limit) {
                if ((threadIdx.x & 31) == 0 && atomicCAS(a.ctl + 3, 0, 1) == 0) {
                    a.ctl[6] = (int)parity; a.ctl[7] = 0; a.ctl[24] = 0;
                    a.ctl[25] = -2; a.ctl[26] = (int)(threadIdx.x >> 5);   // -2: a token between the warps of a CTA
                }
                tri_raise(a, 2); g = true;
            }
            dead = __any_sync(0xffffffffu, g);
        }
    }
    spent += clock64() - t0;
}

// W warps per row (a multiple of 4: a warp's minitiles then sit at a fixed place inside the 32-source chunks of
// the chain-word layout), CTAS resident CTAs per SM.  PROF: lane 0 of every warp accumulates clock64 spans in
// ctl[8..] (chain-word waits, row-accumulator waits, items, triangles, item count, token waits).
// ROWPIPE: the ordered adds of a minitile are fused into the force block of the warp's NEXT minitile
// (forces_column_row): the token is taken at the start of that block and handed on after its first stages.
template <int W, int CTAS, bool PROF = false, bool ROWPIPE = false>
__global__ void __launch_bounds__(W * 32, CTAS)
nbody_crow_kernel(const __grid_constant__ TriArgs a) {
    static_assert(W % 4 == 0 && W >= 4, "warps per row");
    constexpr unsigned FULL = 0xffffffffu;
    constexpr int WS = 24 * W;        // chain words (16 B units) between consecutive minitiles of one warp
    extern __shared__ __align__(16) double crow_smem[];
    const int lane = threadIdx.x & 31, q = threadIdx.x >> 5;
    double *ring = crow_smem + q * kTriWarpDoubles;                   // [comp 3][slot 2][32]
    double *tb = ring + 2 * 3 * 32;                                   // [buffer 2][24][kTbPitch]
    double *seed = tb + 2 * 24 * kTbPitch;                            // [24] chain seeds of a diagonal minitile
    double *tokens = crow_smem + W * kTriWarpDoubles;                 // [warp][x, y, z][32]: the row's sums a warp hands on
    uint64_t *bars = reinterpret_cast<uint64_t *>(tokens + W * 96);   // one mbarrier per giving warp, one arrival (lane 0) per phase
    volatile int *s_ticket = reinterpret_cast<volatile int *>(bars + W);
    double *tok_mine = tokens + q * 96 + lane;
    const double *tok_prev = tokens + ((q + W - 1) % W) * 96 + lane;
    const unsigned bar_mine = smem_u32(bars + q), bar_prev = smem_u32(bars + (q + W - 1) % W);
    unsigned taken = 0;               // tokens taken from the previous warp so far: the phase parity to wait for
    const double *sx = a.src, *sy = a.src + a.npad, *sz = a.src + 2 * (size_t)a.npad;
    const int col_lane = lane < 24 ? lane : 23;                       // lanes 24..31 shadow lane 23 and never publish
    const int cc = col_lane >> 3, cu = col_lane & 7;                  // component, source within the minitile
    unsigned key = 0;
    int dead = 0;
    long long t_chain = 0, t_racc = 0, t_tok = 0;
    long long t_item = 0, t_diag = 0, n_items = 0;
    auto now = [] { return PROF ? clock64() : 0ll; };

    if (threadIdx.x == 0) {
        for (int w = 0; w < W; ++w) mbar_init(bars + w, 1);
        mbar_init_fence();
    }
    for (;;) {
        __syncthreads();                                              // every warp has left the previous item
        if (threadIdx.x == 0) *s_ticket = atomicAdd(&a.ctl[0], 1);
        __syncthreads();
        const int ticket = *s_ticket;
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
        ulonglong2 *pub = (I % a.deal == a.deal - 1 && !final_row) ? a.ll_down : a.ll;
        const int M = (j_end - j_begin) >> 3;             // minitiles of the item (a multiple of 4)
        const int C = M > q ? (M - q + W - 1) / W : 0;    // ... of which this warp runs m = q + W i, i < C

        const double px = sx[i0 + lane], py = sy[i0 + lane], pz = sz[i0 + lane];
        double ax = 0.0, ay = 0.0, az = 0.0;

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
        // the row's sums after `tok` minitiles of this item: from the previous item (tok == 0: global tagged
        // buffer, or zero for the row's first item), else from the warp that ran minitile tok-1
        auto take_token = [&](int tok) {
            if (tok == 0) {
                ax = 0.0; ay = 0.0; az = 0.0;
                if (k > 0) {
                    const ulonglong2 *g = a.racc + (size_t)I * 96 + lane;
                    const unsigned tag = a.epoch_tag | (unsigned)k;
                    ax = ll_unpack(tri_poll(a, g, tag, dead, 100, t_racc));
                    ay = ll_unpack(tri_poll(a, g + 32, tag, dead, 100, t_racc));
                    az = ll_unpack(tri_poll(a, g + 64, tag, dead, 100, t_racc));
                }
                return;
            }
            // Every token given on an edge is taken before the next one is given (the giver's next minitile
            // depends, through the ring of warps, on what the taker does with this one), so phases alternate.
            if (!__all_sync(FULL, mbar_try(bar_prev, taken & 1, 0))) crow_wait(a, bar_prev, taken & 1, dead, t_tok);
            ++taken;
            ax = tok_prev[0]; ay = tok_prev[32]; az = tok_prev[64];
        };
        auto give_token = [&]() {
            tok_mine[0] = ax; tok_mine[32] = ay; tok_mine[64] = az;
            __syncwarp();                             // orders the 32 lanes' stores before lane 0's release
            if (lane == 0) mbar_arrive_at(bar_mine);  // one arrival per phase (32 would serialise on the barrier word)
        };

        auto lower = [&](auto allv) {
            constexpr bool ALLV = decltype(allv)::value;
            if (C <= 0) return;
            const int nch = (C + 3) >> 2;                 // this warp's chunks of 4 minitiles = 32 sources
            // chain word of this warp's minitile i for this column lane
            const ulonglong2 *wp = a.ll + (size_t)((j_begin >> 5) + (q >> 2)) * 96 + cc * 32 + (q & 3) * 8 + cu;
            const ptrdiff_t to_pub = pub - a.ll;
            // lane l of chunk ch: source l % 8 of this warp's minitile 4 ch + l / 8 (clamped inside the item)
            auto src_of = [&](int ch) { return min(j_begin + 8 * (q + W * (4 * ch + (lane >> 3))) + (lane & 7), j_end - 1); };
            { const int g = src_of(0); ring[lane] = sx[g]; ring[64 + lane] = sy[g]; ring[128 + lane] = sz[g]; }
            double fx = 0.0, fy = 0.0, fz = 0.0;                       // positions of chunk ch+1
            if (nch > 1) { const int g = src_of(1); fx = sx[g]; fy = sy[g]; fz = sz[g]; }
            // chain words in flight, two of this warp's minitiles ahead of their use: `we` of the even ones,
            // `wo` of the odd ones (two registers that are never copied: a copy would wait for the load)
            ulonglong2 we = ll_load(wp), wo = make_ulonglong2(0, 0);
            if (C > 1) wo = ll_load(wp + WS);
            __syncwarp();
            double xj[8], yj[8], zj[8];
            auto fetch_sources = [&](int i) {   // the 8 sources of minitile i: ring[x, y, z][slot (i / 4) % 2][32]
                const double *from = ring + (i & 7) * 8;
#pragma unroll
                for (int u = 0; u < 8; u += 2) {
                    const double2 x2 = *reinterpret_cast<const double2 *>(from + u);
                    const double2 y2 = *reinterpret_cast<const double2 *>(from + 64 + u);
                    const double2 z2 = *reinterpret_cast<const double2 *>(from + 128 + u);
                    xj[u] = x2.x; xj[u + 1] = x2.y; yj[u] = y2.x; yj[u + 1] = y2.y; zj[u] = z2.x; zj[u + 1] = z2.y;
                }
            };
            auto next_chunk = [&](int ch) {   // start of chunk ch: chunk ch+1 into the other ring slot (last read during chunk ch-1)
                if (ch + 1 < nch) {
                    double *other = ring + ((ch + 1) & 1) * 32 + lane;
                    other[0] = fx; other[64] = fy; other[128] = fz;
                    if (ch + 2 < nch) { const int g = src_of(ch + 2); fx = sx[g]; fy = sy[g]; fz = sz[g]; }
                }
            };
            // ordered part of minitile i: the row's sums arrive, absorb the 8 sources in order, leave at once;
            // only then the values go to the transposed buffer for the column pass
            auto row_pass = [&](int i, double *mine, const double (&vx)[8], const double (&vy)[8], const double (&vz)[8]) {
                const int m = q + W * i;
                // The 24 values must EXIST before the token is looked at: without this the compiler sinks the whole
                // force block below the poll (fewer live registers across it) and the 208-instruction block becomes
                // part of every hand-off (measured: 1200 cycles per hop, 3x slower than one warp per row).
                asm volatile("" ::"d"(vx[0]), "d"(vx[1]), "d"(vx[2]), "d"(vx[3]), "d"(vx[4]), "d"(vx[5]), "d"(vx[6]), "d"(vx[7]),
                             "d"(vy[0]), "d"(vy[1]), "d"(vy[2]), "d"(vy[3]), "d"(vy[4]), "d"(vy[5]), "d"(vy[6]), "d"(vy[7]),
                             "d"(vz[0]), "d"(vz[1]), "d"(vz[2]), "d"(vz[3]), "d"(vz[4]), "d"(vz[5]), "d"(vz[6]), "d"(vz[7]));
                take_token(m);
#pragma unroll
                for (int u = 0; u < 8; ++u) {   // sources ascending, the target's order; the stores issue under the adds' latency
                    ax = __dadd_rn(ax, vx[u]); ay = __dadd_rn(ay, vy[u]); az = __dadd_rn(az, vz[u]);
                    mine[u * kTbPitch] = vx[u]; mine[(8 + u) * kTbPitch] = vy[u]; mine[(16 + u) * kTbPitch] = vz[u];
                }
                give_token();
            };
            double *const tb0 = tb + lane, *const tb1 = tb + 24 * kTbPitch + lane;              // where this lane writes
            const double *const col0 = tb + col_lane * kTbPitch, *const col1 = col0 + 24 * kTbPitch;   // ... and reads
            // this warp's minitile i >= 1: forces, fused with the column pass of its minitile i-1 (word at `at`, in
            // `w`); `more`: minitile i+1 exists (its word refills the register, its sources are fetched)
            auto block = [&](int i, ulonglong2 &w, const ulonglong2 *at, const double *col_prev, double *mine, bool more) {
                double chain = wait_word(w, at);
                if (more) w = ll_load(at + 2 * WS);
                double vx[8], vy[8], vz[8];
                forces_and_column<ALLV>(px, py, pz, xj, yj, zj, vx, vy, vz, key, col_prev, chain, cnt, lane < 24);
                if (more) fetch_sources(i + 1);
                publish(const_cast<ulonglong2 *>(at) + to_pub, chain);   // before the row pass: the next row gets it that much earlier
                row_pass(i, mine, vx, vy, vz);
                __syncwarp();   // column lanes read these values during the next minitile; ring slot hand-over
            };
            if constexpr (ROWPIPE) {
                double vxa[8], vya[8], vza[8], vxb[8], vyb[8], vzb[8];   // values of this warp's minitiles, alternating
                auto store_tb = [&](double *mine, const double (&vx)[8], const double (&vy)[8], const double (&vz)[8]) {
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        mine[u * kTbPitch] = vx[u]; mine[(8 + u) * kTbPitch] = vy[u]; mine[(16 + u) * kTbPitch] = vz[u];
                    }
                };
                // this warp's minitile i >= 1: its forces, the column pass AND the ordered adds of its minitile i-1
                auto block_p = [&](int i, ulonglong2 &w, const ulonglong2 *at, const double *col_prev, double *mine, bool more,
                                   const double (&ox)[8], const double (&oy)[8], const double (&oz)[8],
                                   double (&nx)[8], double (&ny)[8], double (&nz)[8]) {
                    double chain = wait_word(w, at);
                    if (more) w = ll_load(at + 2 * WS);
                    take_token(q + W * (i - 1));
                    forces_column_row<ALLV>(px, py, pz, xj, yj, zj, nx, ny, nz, key, col_prev, chain, cnt, lane < 24,
                                            ox, oy, oz, ax, ay, az, [&] { give_token(); });
                    if (more) fetch_sources(i + 1);
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
                    if (C > 1) fetch_sources(1);
                    store_tb(tb0, vxa, vya, vza);
                    __syncwarp();
                }
                int i = 1;
#pragma unroll 1
                for (; i + 1 < C; i += 2) {
                    check_range();
                    block_p(i, we, wp, col0, tb1, true, vxa, vya, vza, vxb, vyb, vzb);
                    if ((i & 3) == 3) next_chunk((i + 1) >> 2);
                    block_p(i + 1, wo, wp + WS, col1, tb0, i + 2 < C, vxb, vyb, vzb, vxa, vya, vza);
                    wp += 2 * WS;
                }
                check_range();
                const double *col_last = col0;
                if (i < C) {   // one odd minitile left; its values move to the first set (once per item)
                    block_p(i, we, wp, col0, tb1, false, vxa, vya, vza, vxb, vyb, vzb);
                    wp += WS;
                    col_last = col1;
#pragma unroll
                    for (int u = 0; u < 8; ++u) { vxa[u] = vxb[u]; vya[u] = vyb[u]; vza[u] = vzb[u]; }
                }
                // ordered adds and column pass of this warp's last minitile C-1
                take_token(q + W * (C - 1));
#pragma unroll
                for (int u = 0; u < 8; ++u) { ax = __dadd_rn(ax, vxa[u]); ay = __dadd_rn(ay, vya[u]); az = __dadd_rn(az, vza[u]); }
                give_token();
                double chain = wait_word(col_last == col0 ? we : wo, wp);
#pragma unroll
                for (int tcol = 0; tcol < 16; ++tcol) {
                    double2 v = make_double2(0.0, 0.0);
                    if (lane < 24) v = *reinterpret_cast<const double2 *>(col_last + 2 * tcol);
                    if (ALLV || 2 * tcol < cnt) chain = __dsub_rn(chain, v.x);
                    if (ALLV || 2 * tcol + 1 < cnt) chain = __dsub_rn(chain, v.y);
                }
                check_range();
                publish(const_cast<ulonglong2 *>(wp) + to_pub, chain);
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
                if (C > 1) fetch_sources(1);
                row_pass(0, tb0, vx, vy, vz);
                __syncwarp();
            }
            int i = 1;
#pragma unroll 1
            for (; i + 1 < C; i += 2) {   // pairs (odd, even); wp = word of minitile i-1
                check_range();
                block(i, we, wp, col0, tb1, true);
                if ((i & 3) == 3) next_chunk((i + 1) >> 2);
                block(i + 1, wo, wp + WS, col1, tb0, i + 2 < C);
                wp += 2 * WS;
            }
            check_range();
            const double *col_last = col0;
            if (i < C) {   // one odd minitile left
                block(i, we, wp, col0, tb1, false);
                wp += WS;
                col_last = col1;
            }
            // column pass of this warp's last minitile C-1 (word in `we` if C-1 is even, else `wo`); wp = its word
            double chain = wait_word(col_last == col0 ? we : wo, wp);
#pragma unroll
            for (int tcol = 0; tcol < 16; ++tcol) {
                const double2 v = *reinterpret_cast<const double2 *>(col_last + 2 * tcol);
                if (ALLV || 2 * tcol < cnt) chain = __dsub_rn(chain, v.x);
                if (ALLV || 2 * tcol + 1 < cnt) chain = __dsub_rn(chain, v.y);
            }
            check_range();
            publish(const_cast<ulonglong2 *>(wp) + to_pub, chain);
            __syncwarp();
        };
        if (cnt == 32) lower(std::true_type{});
        else lower(std::false_type{});

        if (q == M % W) {   // next in rotation after the last minitile: the row's sums end here
            take_token(M);
            if (!last_item) {   // hand the row's sums to item (I, k+1)
                ulonglong2 *g = a.racc + (size_t)I * 96 + lane;
                const unsigned tag = a.epoch_tag | (unsigned)(k + 1);
                ll_store(g, ll_pack(ax, tag)); ll_store(g + 32, ll_pack(ay, tag)); ll_store(g + 64, ll_pack(az, tag));
            } else {
                // the row's own 32 x 32 triangle (as in nbody_tri_kernel): sources j = i0 + 8 mt + u against targets
                // i > j; after the row pass of minitile mt the lanes 8 mt .. 8 mt + 7 hold their complete lower sums,
                // which seed the chains of the minitile's sources
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
        }
        if (PROF) { t_item += now() - it0; ++n_items; }
    }
    if (PROF && lane == 0) {
        unsigned long long *c = reinterpret_cast<unsigned long long *>(a.ctl + 8);
        atomicAdd(c + 0, (unsigned long long)t_chain); atomicAdd(c + 1, (unsigned long long)t_racc);
        atomicAdd(c + 2, (unsigned long long)t_item); atomicAdd(c + 3, (unsigned long long)t_diag);
        atomicAdd(c + 4, (unsigned long long)n_items); atomicAdd(c + 5, (unsigned long long)t_tok);
    }
    if (__any_sync(FULL, key >= kRangeLimit)) tri_raise(a, a.range_slot);
}

}  // namespace hgb
