// conv_kernels.cuh -- FP64 K x K clamp-to-edge stencil for sm_100a.
//
// Replaces apply_kernel_to_padded_point / apply_kernel_to_padded_image
// (reference src/convolution/convolution-run.c:52-81) fused with update_borders
// (src/convolution/convolution-util.c:281-294).
//
// Device image layout ("band"): rows of `dp` doubles; row of GLOBAL image row gr
// sits at local row gr - row0 + halo; inside a row the inner pixel x is at
// kLead + x (128-byte aligned), the reference's left pad columns at
// [kLead-p, kLead), the right ones at [kLead+W, kLead+W+p).  `halo` rows above
// and below the band hold the neighbouring ranks' edge rows (multi-GPU); at the
// image's top/bottom the kernels clamp the row index instead, which is exactly
// what the reference's aliased pad rows do (convolution-util.c:394-400).
//
// Exactness: val starts at +0.0 and every tap is one __dmul_rn followed by one
// __dadd_rn in row-major tap order -- bit-identical to the reference (ISO C, no
// contraction) for any finite data.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <type_traits>

#include "async_copy.cuh"

namespace hgb {

constexpr int kLead = 16;      // doubles in front of inner pixel 0 of every device row
constexpr int kMaxTaps = 49;   // specialised kernels: 3x3, 5x5, 7x7

struct Taps { double k[kMaxTaps]; };   // kernel parameter => constant bank operands

struct BandGeom {
    int W, H;        // inner image size (global)
    int p;           // reference padding
    int row0, rows;  // this band: global rows [row0, row0+rows) -- defines the addressing
    int crow0, crows;  // rows this LAUNCH computes (a sub-range of the band: boundary / interior split)
    int halo;        // halo rows stored above/below the band
    int dp;          // device row pitch in doubles
};

// One tap.  EXACT: val + k*v as the reference computes it under ISO C (no contraction): one rounding for
// the product, one for the sum (reference src/convolution/convolution-run.c:71-77).  FAST (opt-in, floating-
// point images only, never the parity path): one fused multiply-add -- half the FP64 instructions, a
// single rounding per tap, results within ~1e-15 relative per iteration of the exact ones.
template <bool FAST>
__device__ __forceinline__ double conv_tap(double val, double k, double v) {
    return FAST ? __fma_rn(k, v, val) : __dadd_rn(val, __dmul_rn(k, v));
}

__device__ __forceinline__ size_t band_row_off(const BandGeom &g, int gr) {
    return (size_t)(gr - g.row0 + g.halo) * g.dp;
}

// ---- generic fallback: any kw, kh <= 2p+1 --------------------------------
template <bool FAST = false>
__global__ void __launch_bounds__(256)
conv_generic_kernel(const double *__restrict__ in, double *__restrict__ out, BandGeom g,
                    const double *__restrict__ taps, int kw, int kh) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = g.crow0 + blockIdx.y;
    if (x >= g.W) return;
    double val = 0.0;
    for (int ky = 0; ky < kh; ++ky) {
        const int gr = min(max(y + ky - g.p, 0), g.H - 1);
        const double *row = in + band_row_off(g, gr) + kLead - g.p + x;   // padded column x + kx
        for (int kx = 0; kx < kw; ++kx) val = conv_tap<FAST>(val, taps[ky * kw + kx], row[kx]);
    }
    double *orow = out + band_row_off(g, y) + kLead;
    orow[x] = val;
    if (x == 0)
        for (int i = 1; i <= g.p; ++i) orow[-i] = val;            // update_borders, left
    if (x == g.W - 1)
        for (int i = 0; i < g.p; ++i) orow[g.W + i] = val;        // update_borders, right
}

// ---- specialised K x K (K = 2p+1) row-streaming kernel ---------------------
// A CTA owns a strip of TW columns and a chunk of `chunk` output rows.  It
// streams the chunk's input rows (chunk + 2p of them, row index clamped at the
// image edge) through a shared-memory ring, K rows per stage, filled by bulk
// copies issued by one thread and tracked with one mbarrier per stage.  Each
// thread owns C adjacent columns and keeps K x C running sums in registers: input
// row i feeds tap row ky of output row i - ky, so every output pixel receives
// its taps in the reference's row-major order while each input value is read
// from shared memory once per thread.
template <int K, int TW, int C, int NST>
struct ConvCfg {
    static constexpr int P = K / 2;
    static constexpr int LP = (P + 1) & ~1;          // halo columns staged per side (16 B granules)
    static constexpr int ROWW = TW + 2 * LP;         // doubles per staged row
    static constexpr int THREADS = TW / C;
    static constexpr int STAGE_ROWS = K;
    static constexpr size_t SMEM = (size_t)NST * STAGE_ROWS * ROWW * sizeof(double) + NST * 8;
};

template <int K, int TW, int C, int NST, bool FAST = false>
__global__ void __launch_bounds__(TW / C)
conv_stream_kernel(const double *__restrict__ in, double *__restrict__ out, BandGeom g, int chunk,
                   const __grid_constant__ Taps taps) {
    using Cfg = ConvCfg<K, TW, C, NST>;
    constexpr int P = Cfg::P, LP = Cfg::LP, ROWW = Cfg::ROWW;
    constexpr int NIN = C + 2 * P;                   // input values per thread per row
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *ring = reinterpret_cast<double *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(ring + (size_t)NST * K * ROWW);

    const int x0 = blockIdx.x * TW;
    const int y0 = g.crow0 + blockIdx.y * chunk;                 // first output row (global)
    const int nrows = min(chunk, g.crow0 + g.crows - y0);        // output rows of this CTA
    const int nin = nrows + 2 * P;                               // input rows streamed
    const int nstages = (nin + K - 1) / K;
    const int tid = threadIdx.x;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < NST; ++s) mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int st) {   // thread 0 only: fill ring slot st % NST with stage st
        const int slot = st % NST;
        const int i0 = st * K;
        const int cnt = min(K, nin - i0);
        mbar_expect_tx(&full[slot], (uint32_t)(cnt * ROWW * sizeof(double)));
        for (int r = 0; r < cnt; ++r) {
            const int gr = min(max(y0 - P + i0 + r, 0), g.H - 1);   // aliased pad rows == clamp
            const double *src = in + band_row_off(g, gr) + kLead + x0 - LP;
            bulk_g2s(ring + (size_t)(slot * K + r) * ROWW, src, ROWW * sizeof(double), &full[slot]);
        }
    };
    if (tid == 0)
        for (int st = 0; st < NST - 1 && st < nstages; ++st) issue(st);

    double acc[K][C];
#pragma unroll
    for (int a = 0; a < K; ++a)
#pragma unroll
        for (int c = 0; c < C; ++c) acc[a][c] = 0.0;

    const int xt = x0 + C * tid;                       // first column of this thread
    const bool any_col = xt < g.W;

    for (int st = 0; st < nstages; ++st) {
        const int slot = st % NST;
        // slot (st-1)%NST was drained by every thread before the barrier at the end of
        // the previous iteration, so it can be refilled with stage st+NST-1 now
        if (tid == 0 && st + NST - 1 < nstages) issue(st + NST - 1);
        mbar_wait(&full[slot], (uint32_t)((st / NST) & 1));
#pragma unroll
        for (int ph = 0; ph < K; ++ph) {
            const int i = st * K + ph;                 // input row index within the chunk
            if (i < nin) {
                const double *srow = ring + (size_t)(slot * K + ph) * ROWW + C * tid + (LP - P);
                double v[NIN];
#pragma unroll
                for (int q = 0; q < NIN; ++q) v[q] = srow[q];
                // steady-state rows feed all K output rows; the first/last K-1 rows of a chunk only
                // some of them (no output above/below the chunk): one warp-uniform test per row
                auto row_body = [&](auto full_tag) {
                    constexpr bool FULL = decltype(full_tag)::value;
                    constexpr bool EDGE = K <= 5;
#pragma unroll
                    for (int ky = 0; ky < K; ++ky) {
                        const int a = (ph - ky + K) % K;   // compile-time accumulator slot of row i-ky
                        if (!FULL && (i - ky < 0 || i - ky >= nrows)) continue;
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            double val = (ky == 0) ? 0.0 : acc[a][c];
#pragma unroll
                            for (int kx = 0; kx < K; ++kx)
                                val = conv_tap<FAST>(val, taps.k[ky * K + kx], v[c + kx]);
                            acc[a][c] = val;
                        }
                        if (ky == K - 1 && any_col && (EDGE || i >= K - 1)) {   // output row i-(K-1) is complete
                            const int gy = y0 + i - (K - 1);
                            double *orow = out + band_row_off(g, gy) + kLead;
                            if (C == 2 && xt + 1 < g.W) {
                                *reinterpret_cast<double2 *>(orow + xt) = make_double2(acc[a][0], acc[a][1]);
                            } else {
#pragma unroll
                                for (int c = 0; c < C; ++c)
                                    if (xt + c < g.W) orow[xt + c] = acc[a][c];
                            }
                            if (xt == 0) {                                // update_borders, left
#pragma unroll
                                for (int q = 1; q <= P; ++q) orow[-q] = acc[a][0];
                            }
#pragma unroll
                            for (int c = 0; c < C; ++c)
                                if (xt + c == g.W - 1) {                  // update_borders, right
#pragma unroll
                                    for (int q = 0; q < P; ++q) orow[g.W + q] = acc[a][c];
                                }
                        }
                    }
                };
                // (7x7: the duplicated body would not fit the instruction cache -- measured slower --
                //  so it computes the few useless edge taps instead)
                constexpr bool EDGE = K <= 5;
                if (!EDGE || (i >= K - 1 && i < nrows)) row_body(std::true_type{});
                else row_body(std::false_type{});
            }
        }
        __syncthreads();   // everyone is done reading this slot
    }
}

// ---- 3x3: two iterations per pass (temporal blocking) -----------------------
// The 3x3 stencil is HBM-bound (16 B and 18 FP64 ops per pixel), so a pass that applies the
// kernel TWICE halves the DRAM traffic per iteration.  A CTA streams input rows y0-2 .. y0+n+1 of
// a strip through a shared-memory ring (bulk copies, one mbarrier per row); stage 1 turns them
// into rows of the intermediate state (iteration t+1), which live only in a two-row shared-memory
// buffer; stage 2 consumes each intermediate row at once and emits rows of iteration t+2.  Both
// stages keep 3x2 running sums per thread in registers and apply the taps in the reference's
// row-major order with one __dmul_rn + one __dadd_rn each, so every value -- intermediate and
// final -- is bit-identical to two separate passes.  update_borders between the two iterations
// (pad columns = edge pixels) and the aliased pad rows (row index clamp) are applied to the
// intermediate state on the fly:
//   * columns: a strip computes 256 intermediate columns xs .. xs+255 and outputs the 254 inside
//     them; an intermediate column left of 0 / right of W-1 is replaced by column 0 / W-1;
//   * rows: intermediate row -1 / H would alias row 0 / H-1, so output row 0 takes its first tap
//     row from intermediate row 0 as well, and output row H-1 its last one from row H-1.
constexpr int kF2TW = 256;              // intermediate columns per CTA
constexpr int kF2Out = kF2TW - 2;       // output columns per CTA
constexpr int kF2Pitch = kF2TW + 4;     // shared-memory row pitch (doubles): 16-byte multiple, room for the reads past TW
constexpr int kF2Ring = 6;              // input rows in flight: two halves of three (the loop is unrolled by 3)
constexpr size_t kF2Smem = (size_t)(kF2Ring + 2) * kF2Pitch * sizeof(double) + kF2Ring * sizeof(uint64_t);

// C = columns per thread (2 or 4); kF2TW / C threads per CTA.  Measured equal on B200 (a pass moves
// 4.39 GB, so its HBM floor is ~650 us; both variants take ~800 us for two iterations with the FP64
// pipe ~69 % busy, ncu: profiles/r1c_conv3_fused_ncu_summary.txt); C = 2 is the default.
template <int C, bool FAST = false>
__global__ void __launch_bounds__(kF2TW / C)
conv3_fused2_kernel(const double *__restrict__ in, double *__restrict__ out, BandGeom g, int chunk,
                    const __grid_constant__ Taps taps) {
    constexpr int T = kF2TW / C, NV = C + 2;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *inring = reinterpret_cast<double *>(smem_raw);
    double *mid = inring + kF2Ring * kF2Pitch;
    uint64_t *full = reinterpret_cast<uint64_t *>(mid + 2 * kF2Pitch);

    const int xs = blockIdx.x * kF2Out - 1;                      // first intermediate column
    const int y0 = g.crow0 + blockIdx.y * chunk;                 // first output row (global)
    const int nrows = min(chunk, g.crow0 + g.crows - y0);
    const int nin = nrows + 4;                                   // input rows y0-2 .. y0+nrows+1
    const int tid = threadIdx.x;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kF2Ring; ++s) mbar_init(&full[s], 1);
        mbar_init_fence();
    }
    __syncthreads();

    // thread 0: input row i of the chunk into ring slot `slot` (= i % 6)
    const double *in_strip = in + kLead + (xs - 1);              // even column: 16-byte aligned
    auto issue = [&](int i, int slot) {
        const int gr = min(max(y0 - 2 + i, 0), g.H - 1);         // aliased pad rows == clamp
        mbar_expect_tx(&full[slot], (uint32_t)((kF2TW + 2) * sizeof(double)));
        bulk_g2s(inring + slot * kF2Pitch, in_strip + band_row_off(g, gr), (kF2TW + 2) * sizeof(double),
                 &full[slot]);
    };
    if (tid == 0)
        for (int i = 0; i < kF2Ring && i < nin; ++i) issue(i, i);

    const int a = xs + C * tid;            // global column of this thread's first intermediate value
    const int c0 = a + 1;                  // first of its C output columns
    // output columns of the strip end at xs + TW - 2 (the last thread owns fewer) and at W - 1
    const int nvalid = max(0, min(C, min(g.W, xs + kF2TW - 1) - c0));
    const bool pad_left = nvalid > 0 && c0 == 0;                 // update_borders of the output
    const int pad_right = g.W - 1 - c0;                          // this thread's output that is column W-1, if in [0, nvalid)
    const bool edge_strip = blockIdx.x == 0 || xs + kF2TW + 2 > g.W;
    const int last = g.W - 1 - a;          // position of column W-1 among this thread's intermediate values
    // chunks that hold image row 0 or H-1 take the variant of stage 2 with the row-alias rules
    const bool edge_rows = y0 == 0 || y0 + nrows >= g.H;
    double *const out_col = out + band_row_off(g, y0) + kLead + c0;

    double a1[3][C], a2[3][C];
#pragma unroll
    for (int q = 0; q < 3; ++q)
#pragma unroll
        for (int c = 0; c < C; ++c) { a1[q][c] = 0.0; a2[q][c] = 0.0; }

    auto store_row = [&](int o, const double (&val)[C]) {
        double *orow = out_col + (size_t)o * g.dp;
#pragma unroll
        for (int c = 0; c < C; c += 2) {
            if (c + 1 < nvalid) *reinterpret_cast<double2 *>(orow + c) = make_double2(val[c], val[c + 1]);
            else if (c < nvalid) orow[c] = val[c];
        }
        if (pad_left) orow[-1] = val[0];
        if (pad_right >= 0 && pad_right < nvalid) {
            double v = val[0];
#pragma unroll
            for (int c = 1; c < C; ++c) v = pad_right == c ? val[c] : v;
            orow[pad_right + 1] = v;
        }
    };
    auto load_row = [&](const double *row, double (&v)[NV]) {    // NV consecutive doubles, 16-byte aligned
#pragma unroll
        for (int q = 0; q < NV; q += 2) {
            const double2 t = *reinterpret_cast<const double2 *>(row + q);
            v[q] = t.x; v[q + 1] = t.y;
        }
    };

    for (int k = 0, i0 = 0; i0 < nin; ++k, i0 += 3) {
        const int half = k & 1;                                  // ring slot of phase ph is 3*half + ph
        const uint32_t parity = (uint32_t)(k >> 1) & 1u;
        const double *in_half = inring + half * 3 * kF2Pitch + C * tid;
        double *mid_a = mid + half * kF2Pitch + C * tid;         // intermediate row j lives in slot j & 1 == (k + ph) & 1
        double *mid_b = mid + (half ^ 1) * kF2Pitch + C * tid;
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            const int i = i0 + ph;
            if (i >= nin) break;
            mbar_wait(&full[3 * half + ph], parity);
            {   // stage 1: input row i feeds the intermediate rows it is the top / middle / bottom row of
                double v[NV];                                    // input columns a-1 .. a+C
                load_row(in_half + ph * kF2Pitch, v);
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const int sl = (ph - ky + 3) % 3;
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        double val = (ky == 0) ? 0.0 : a1[sl][c];
#pragma unroll
                        for (int kx = 0; kx < 3; ++kx) val = conv_tap<FAST>(val, taps.k[ky * 3 + kx], v[c + kx]);
                        a1[sl][c] = val;
                    }
                }
            }
            // intermediate row j = i-2 (global row y0-1+j) is complete in accumulator slot (ph+1)%3
            const int j = i - 2;
            double *mrow = (ph == 1) ? mid_b : mid_a;
            if (j >= 0) {
#pragma unroll
                for (int c = 0; c < C; c += 2)
                    *reinterpret_cast<double2 *>(mrow + c) = make_double2(a1[(ph + 1) % 3][c], a1[(ph + 1) % 3][c + 1]);
            }
            __syncthreads();               // intermediate row visible; input slot drained by everyone
            if (tid == 0 && i + kF2Ring < nin) issue(i + kF2Ring, 3 * half + ph);
            if (j < 0) continue;
            // stage 2: intermediate row j feeds output rows j, j-1, j-2 (relative to y0)
            double w[NV];                                        // intermediate columns a .. a+C+1
            load_row(mrow, w);
            if (edge_strip) {              // update_borders of the intermediate state
                if (a < 0) w[0] = w[1];
                if (last >= 0 && last < NV - 1) {
                    double wl = w[0];
#pragma unroll
                    for (int q = 1; q < NV - 1; ++q) wl = last == q ? w[q] : wl;
#pragma unroll
                    for (int q = 1; q < NV; ++q) w[q] = q > last ? wl : w[q];
                }
            }
            auto stage2 = [&](auto edge_tag) {
                constexpr bool EDGE = decltype(edge_tag)::value;
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const int sl = (ph + 1 - ky + 3) % 3;
                    const int o = j - ky;                        // output row relative to y0
                    const bool top = EDGE && ky == 1 && y0 + o == 0;            // warp-uniform
                    const bool bottom = EDGE && ky == 1 && y0 + o == g.H - 1;
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        double val = (ky == 0) ? 0.0 : a2[sl][c];
                        if (EDGE && top) { // intermediate row -1 aliases row 0: tap row 0 comes from this row too
                            val = 0.0;
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx) val = conv_tap<FAST>(val, taps.k[kx], w[c + kx]);
                        }
#pragma unroll
                        for (int kx = 0; kx < 3; ++kx) val = conv_tap<FAST>(val, taps.k[ky * 3 + kx], w[c + kx]);
                        if (EDGE && bottom) {   // intermediate row H aliases row H-1: tap row 2 from this row as well
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx) val = conv_tap<FAST>(val, taps.k[6 + kx], w[c + kx]);
                        }
                        a2[sl][c] = val;
                    }
                    if (EDGE) {
                        const bool done = (ky == 2 && y0 + o != g.H - 1) || bottom;
                        if (done && o >= 0 && o < nrows) store_row(o, a2[sl]);
                    } else {
                        if (ky == 2 && o >= 0) store_row(o, a2[sl]);   // o < nrows holds: i < nin
                    }
                }
            };
            if (edge_rows) stage2(std::true_type{});
            else stage2(std::false_type{});
        }
    }
}

// ---- 3x3, two iterations per pass, warp-independent variant ------------------------------
// Same arithmetic and the same row/column alias rules as conv3_fused2_kernel, but the intermediate
// state never touches shared memory: a WARP owns 64 intermediate columns (2 per lane) and outputs
// the 62 inside them; the two intermediate values a lane needs from its right neighbour come by
// warp shuffle, so there is no CTA-wide barrier per row.  The input ring is still shared by the
// CTA's warps: `full` mbarriers announce a landed row, `empty` mbarriers (one arrival per warp)
// hand a slot back to the producer lane, which refills it five rows ahead.
constexpr int kF2wWarps = 4;
constexpr int kF2wOut = 62 * kF2wWarps;      // output columns per CTA (248)
constexpr int kF2wRow = 256;                 // staged doubles per input row: columns xs-1 .. xs+254
constexpr size_t kF2wSmem = (size_t)kF2Ring * kF2wRow * sizeof(double) + 2 * kF2Ring * sizeof(uint64_t);

__global__ void __launch_bounds__(kF2wWarps * 32)
conv3_fused2w_kernel(const double *__restrict__ in, double *__restrict__ out, BandGeom g, int chunk,
                     const __grid_constant__ Taps taps) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *inring = reinterpret_cast<double *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(inring + kF2Ring * kF2wRow);
    uint64_t *empty = full + kF2Ring;

    const int xs = blockIdx.x * kF2wOut - 1;                     // first intermediate column of the CTA
    const int y0 = g.crow0 + blockIdx.y * chunk;                 // first output row (global)
    const int nrows = min(chunk, g.crow0 + g.crows - y0);
    const int nin = nrows + 4;                                   // input rows y0-2 .. y0+nrows+1
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kF2Ring; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], kF2wWarps); }
        mbar_init_fence();
    }
    __syncthreads();

    const double *in_strip = in + kLead + (xs - 1);              // even column: 16-byte aligned
    auto issue = [&](int i, int slot) {                          // producer lane: input row i into `slot`
        const int gr = min(max(y0 - 2 + i, 0), g.H - 1);         // aliased pad rows == clamp
        mbar_expect_tx(&full[slot], (uint32_t)(kF2wRow * sizeof(double)));
        bulk_g2s(inring + slot * kF2wRow, in_strip + band_row_off(g, gr), kF2wRow * sizeof(double), &full[slot]);
    };
    if (tid == 0)
        for (int i = 0; i < kF2Ring && i < nin; ++i) issue(i, i);

    const int a = xs + 62 * warp + 2 * lane;   // global column of this lane's first intermediate value
    const int c0 = a + 1;                      // first of its two output columns
    const int nvalid = lane == 31 ? 0 : max(0, min(2, g.W - c0));
    const bool pad_left = nvalid > 0 && c0 == 0;
    const int pad_right = g.W - 1 - c0;        // which of the two outputs is column W-1 (if 0 or 1)
    const bool edge_strip = blockIdx.x == 0 || xs + kF2wRow > g.W;
    const int last = g.W - 1 - a;
    const bool edge_rows = y0 == 0 || y0 + nrows >= g.H;
    double *const out_col = out + band_row_off(g, y0) + kLead + c0;

    double a1[3][2], a2[3][2];
#pragma unroll
    for (int q = 0; q < 3; ++q) { a1[q][0] = a1[q][1] = 0.0; a2[q][0] = a2[q][1] = 0.0; }

    auto store_row = [&](int o, double v0, double v1) {
        double *orow = out_col + (size_t)o * g.dp;
        if (nvalid == 2) *reinterpret_cast<double2 *>(orow) = make_double2(v0, v1);
        else if (nvalid == 1) orow[0] = v0;
        if (pad_left) orow[-1] = v0;
        if (pad_right >= 0 && pad_right < nvalid) orow[pad_right + 1] = pad_right == 0 ? v0 : v1;
    };
    for (int k = 0, i0 = 0; i0 < nin; ++k, i0 += 3) {
        const int half = k & 1;                                  // ring slot of phase ph is 3*half + ph
        const uint32_t parity = (uint32_t)(k >> 1) & 1u;
        const double *in_half = inring + half * 3 * kF2wRow + 62 * warp + 2 * lane;
#pragma unroll
        for (int ph = 0; ph < 3; ++ph) {
            const int i = i0 + ph;
            if (i >= nin) break;
            const int slot = 3 * half + ph;
            mbar_wait(&full[slot], parity);
            const double *srow = in_half + ph * kF2wRow;
            const double2 p0 = *reinterpret_cast<const double2 *>(srow);
            const double2 p1 = *reinterpret_cast<const double2 *>(srow + 2);
            const double v[4] = {p0.x, p0.y, p1.x, p1.y};        // input columns a-1 .. a+2
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[slot]);            // this warp has the row in registers
            if (tid == 0 && i >= 1 && i + kF2Ring - 1 < nin) {
                // row i+5 goes where row i-1 was, once every warp has handed that slot back
                const int pslot = ph == 0 ? 3 * (half ^ 1) + 2 : slot - 1;
                mbar_wait(&empty[pslot], (uint32_t)(((i - 1) / kF2Ring) & 1));
                issue(i + kF2Ring - 1, pslot);
            }
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {                     // stage 1
                const int sl = (ph - ky + 3) % 3;
#pragma unroll
                for (int c = 0; c < 2; ++c) {
                    double val = (ky == 0) ? 0.0 : a1[sl][c];
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) val = __dadd_rn(val, __dmul_rn(taps.k[ky * 3 + kx], v[c + kx]));
                    a1[sl][c] = val;
                }
            }
            const int j = i - 2;           // intermediate row (global y0-1+j) complete in slot (ph+1)%3
            if (j < 0) continue;
            double w[4];                   // intermediate columns a .. a+3: own pair + the right neighbour's
            w[0] = a1[(ph + 1) % 3][0];
            w[1] = a1[(ph + 1) % 3][1];
            w[2] = __shfl_down_sync(0xffffffffu, w[0], 1);
            w[3] = __shfl_down_sync(0xffffffffu, w[1], 1);
            if (edge_strip) {              // update_borders of the intermediate state
                if (a < 0) w[0] = w[1];
                if (last >= 0 && last < 3) {
                    const double wl = last == 0 ? w[0] : last == 1 ? w[1] : w[2];
                    if (last < 1) w[1] = wl;
                    if (last < 2) w[2] = wl;
                    w[3] = wl;
                }
            }
            auto stage2 = [&](auto edge_tag) {
                constexpr bool EDGE = decltype(edge_tag)::value;
#pragma unroll
                for (int ky = 0; ky < 3; ++ky) {
                    const int sl = (ph + 1 - ky + 3) % 3;
                    const int o = j - ky;                        // output row relative to y0
                    const bool top = EDGE && ky == 1 && y0 + o == 0;
                    const bool bottom = EDGE && ky == 1 && y0 + o == g.H - 1;
#pragma unroll
                    for (int c = 0; c < 2; ++c) {
                        double val = (ky == 0) ? 0.0 : a2[sl][c];
                        if (EDGE && top) {
                            val = 0.0;
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx) val = __dadd_rn(val, __dmul_rn(taps.k[kx], w[c + kx]));
                        }
#pragma unroll
                        for (int kx = 0; kx < 3; ++kx) val = __dadd_rn(val, __dmul_rn(taps.k[ky * 3 + kx], w[c + kx]));
                        if (EDGE && bottom) {
#pragma unroll
                            for (int kx = 0; kx < 3; ++kx) val = __dadd_rn(val, __dmul_rn(taps.k[6 + kx], w[c + kx]));
                        }
                        a2[sl][c] = val;
                    }
                    if (EDGE) {
                        const bool done = (ky == 2 && y0 + o != g.H - 1) || bottom;
                        if (done && o >= 0 && o < nrows) store_row(o, a2[sl][0], a2[sl][1]);
                    } else {
                        if (ky == 2 && o >= 0) store_row(o, a2[sl][0], a2[sl][1]);
                    }
                }
            };
            if (edge_rows) stage2(std::true_type{});
            else stage2(std::false_type{});
        }
    }
}

// add_padding's pad columns (reference src/convolution/convolution-util.c:258-269 via
// init_padded_image/update_borders): for stored rows [r0, r1) of the band (halo rows included)
// copy inner pixel 0 into the p columns left of it and pixel W-1 into the p columns right of it.
__global__ void __launch_bounds__(128)
conv_pad_columns_kernel(double *__restrict__ img, BandGeom g, int r0, int r1) {
    const int y = r0 + blockIdx.x * (blockDim.x / 8) + threadIdx.x / 8;
    const int i = threadIdx.x % 8;          // p <= 7 for the specialised kernels; loop covers more
    if (y >= r1) return;
    double *row = img + band_row_off(g, y) + kLead;
    const double left = row[0], right = row[g.W - 1];
    for (int q = i; q < g.p; q += 8) { row[-1 - q] = left; row[g.W + q] = right; }
}

// FAST mode guard: is every inner pixel of rows [r0, r1) an integer-valued double?  (north_star: integer
// images must stay bit-exact, so they keep the exact kernels even when FAST is requested.)  flag[0] is
// set to 1 by any thread that finds a non-integer (or non-finite) pixel.
// Only the W inner pixels of each row are looked at.  The caller's pad columns (read by the first iteration,
// rewritten by every iteration) are left out on purpose: missing a non-integer value there can only keep the
// EXACT kernels where FAST would have been allowed -- the error is on the bit-identical side.
__global__ void __launch_bounds__(256)
conv_not_integer_kernel(const double *__restrict__ img, BandGeom g, int r0, int r1, int *flag) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= g.W) return;
    bool bad = false;
    for (int y = r0 + blockIdx.y; y < r1; y += gridDim.y) {
        const double v = img[band_row_off(g, y) + kLead + x];
        bad |= !(v == rint(v)) || !(fabs(v) <= 9007199254740992.0);
    }
    if (bad) flag[0] = 1;
}

// sequential row-major checksum (reference get_checksum, convolution-util.c:63-71):
// one warp stages 32 doubles at a time, lane 0 adds them in order.  Exact but
// serial; used by tests and the CLI path only.
__global__ void __launch_bounds__(32)
conv_checksum_seq_kernel(const double *__restrict__ img, BandGeom g, double carry_in,
                         double *__restrict__ result) {
    __shared__ double buf[2][1024];
    double val = carry_in;
    const int lane = threadIdx.x;
    for (int y = g.row0; y < g.row0 + g.rows; ++y) {
        const double *row = img + band_row_off(g, y) + kLead;
        for (int x0 = 0; x0 < g.W; x0 += 1024) {
            const int cnt = min(1024, g.W - x0);
            double *b = buf[(x0 >> 10) & 1];
            for (int q = lane; q < cnt; q += 32) b[q] = row[x0 + q];
            __syncwarp();
            if (lane == 0)
                for (int q = 0; q < cnt; ++q) val = __dadd_rn(val, b[q]);
            __syncwarp();
        }
    }
    if (lane == 0) result[0] = val;
}

// rows of `cols` doubles from pitch `sp` to pitch `dp_` (both in doubles): the streamed drop-in moves row
// chunks between its staging buffers (host pitch) and the image (device pitch) with SMs, so that the
// copy engines stay free for the PCIe transfers running beside it
__global__ void __launch_bounds__(256)
conv_copy_rows_kernel(double *__restrict__ dst, size_t dp_, const double *__restrict__ src, size_t sp, int cols, int rows) {
    for (int r = blockIdx.y; r < rows; r += gridDim.y) {
        const double *s = src + (size_t)r * sp;
        double *d = dst + (size_t)r * dp_;
        for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < cols; c += gridDim.x * blockDim.x) d[c] = s[c];
    }
}

}  // namespace hgb
