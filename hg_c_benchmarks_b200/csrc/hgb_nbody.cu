// hgb_nbody.cu -- host side of the nbody path: device-resident handle, target
// sharding, position exchange (ncclAllGather or fused peer stores) and the
// drop-in for the reference's run() (src/nbody/nbody-run.c:27-37).
#include <algorithm>

#include "hgb_common.cuh"
#include "nbody_kernels.cuh"
#include "nbody_sym.cuh"
#include "nbody_tri.cuh"
#include "nbody_crow.cuh"
#include "nbody_seg.cuh"

using namespace hgb;

namespace {
constexpr int kAlign = 1024;  // >= largest TILE (256 threads x 4 loads)

struct NbShard {
    DevCtx *dev = nullptr;
    int index = 0;               // which target slice this device owns
    double *state[2] = {nullptr, nullptr};  // ping-pong SoA, 3*npad doubles each
    double *stage = nullptr;     // AoS staging, 3*npad doubles
    double *partials = nullptr;  // energy block sums
    double *vel = nullptr;       // integrator extension: SoA velocities, 3*npad (own slice is live)
    double *mass = nullptr;      // integrator extension: npad masses, zero beyond n
    int *tlist = nullptr;
    int tlist_cap = 0;
    int *sym_ctl = nullptr;      // pair-symmetric kernel: ticket, range flag, profiling counters (nbody_sym.cuh)
    ulonglong2 *sym_ll = nullptr; // ... tagged chain accumulators, 48 B per body
    unsigned sym_epoch = 0;
    // second-generation pair-symmetric kernel (nbody_tri.cuh)
    int *tri_ctl = nullptr;                       // ticket, range flag, abort flag, profiling counters
    ulonglong2 *tri_ll[2] = {nullptr, nullptr};   // chain words, double-buffered by step parity, 48 B per body each
    ulonglong2 *tri_racc = nullptr;               // row accumulators between the items of a row, 48 B per body
    int2 *tri_tiles = nullptr;                    // this shard's tiles in ticket order
    int tri_ntiles = 0, tri_plan_seg = 0, tri_plan_rows = 0, tri_plan_weight = 0, tri_plan_deal = 0, tri_plan_rs = 0;
    // first-generation warp on the segmented schedule (nbody_seg.cuh): row sums between items, tiles in ticket order
    ulonglong2 *seg_racc = nullptr;
    int2 *seg_tiles = nullptr;
    int seg_ntiles = 0, seg_plan_seg = 0, seg_plan_rows = 0, seg_plan_weight = 0;
};
}  // namespace

struct hgb_nbody {
    int n = 0, npad = 0, slice = 0, world = 1;
    bool rank_mode = false;
    int mode = HGB_NBODY_EXACT, exchange = HGB_EXCHANGE_NCCL;
    int block = 0, unroll = 0, warps = 0;
    int cur = 0;  // state[cur] holds the newest positions
    bool uploaded = false;
    bool dynamics = false;  // hgb_nbody_set_dynamics was called
    int sym = 0;            // pair-symmetric kernel variant R*100 + U*10 + CTAs/SM, or 2000 + CTAs/SM (0 = off)
    int sym_auto = 0;       // the library picks the variant per GPU count (sym_variant_in_effect)
    int sym_min_n = 0;      // ... used when n >= sym_min_n, EXACT mode
    int tri_seg = 0;        // second-generation kernel (variants 2000+): sources per segment,
    int tri_rows = 0;       // ... rows per tile,
    int tri_weight = 0;     // ... weight of the segment index in the ticket order (0 = defaults)
    int tri_deal = 0;       // ... rows per turn of a GPU (0 = default)
    unsigned tri_step = 0;  // steps run with it so far (epoch and buffer parity)
    int tri_status = 0;     // abort flag read back after the last advance
    int seg_hyst = 0;       // HGB_NBODY_SEG_HYST, read once when the handle is made
    int seg_inject_abort = -1;   // HGB_NBODY_SEG_INJECT_ABORT=k (test aid): raise the abort flag before one-GPU segmented step k
    unsigned seg_step = 0;  // one-GPU segmented steps launched on this handle so far
    // chain-word buffers and control words of every shard as THIS process addresses them (own allocations,
    // peer allocations of the same process, or cudaIpcOpenMemHandle mappings of the other ranks')
    ulonglong2 *tri_peer_ll[2][kMaxPeers] = {};
    int *tri_peer_ctl[kMaxPeers] = {};
    int tri_connected = 0;  // 0 not tried, 1 tables filled, -1 the other ranks' buffers cannot be mapped (standard kernel is used)
    double dt = 0.0;
    std::vector<NbShard> sh;
};

static int g_default_mode = HGB_NBODY_EXACT;
static int g_default_exchange = HGB_EXCHANGE_NCCL;

template <int BLOCK, int UNROLL, int WARPS>
static void launch_variant(bool fast, int grid, cudaStream_t st, const double *src, int npad, int n,
                           int t0, int count, const int *tlist, const PeerOut &out, double *out_list,
                           const int *run_if) {
    const Dynamics none{nullptr, nullptr, 0.0};
    if (fast)
        nbody_step_kernel<BLOCK, UNROLL, WARPS, true><<<grid, BLOCK, 0, st>>>(src, npad, n, t0, count,
                                                                            tlist, out, out_list, none, run_if);
    else
        nbody_step_kernel<BLOCK, UNROLL, WARPS, false><<<grid, BLOCK, 0, st>>>(src, npad, n, t0, count,
                                                                             tlist, out, out_list, none, run_if);
}

// CTA shape for `count` targets when the caller did not pin one (hgb_nbody_tune): 128-thread CTAs, 6 per
// SM, two double2 source loads per group (measured best at N = 2^20, profiles/r1_nbody_sweep.txt).  Below
// one wave (N = 65536: 3.46 CTAs per SM) the kernel runs at ~80 % of its large-N rate; that is a latency
// effect of ~14 warps per SM, not the rounding of CTAs over SMs: 64-thread CTAs (7 x 64 = 448 targets on
// the fullest SM instead of 4 x 128 = 512) measure the same or slower (profiles/r2a_nbody_small_sweep.txt:
// 9.69 ms for 128/2/24, 9.71 ms for 64/2/24, 10.29 ms for 64/2/28), so the shape stays fixed.
static void auto_shape(int count, int sms, int *block, int *unroll, int *warps) {
    (void)count; (void)sms;
    *block = 128; *unroll = 2; *warps = 24;
}

static int launch_step(const hgb_nbody *h, const NbShard &s, const double *src, int t0, int count,
                       const int *tlist, const PeerOut &out, double *out_list,
                       const int *run_if = nullptr) {
    if (count <= 0) return HGB_OK;
    const bool fast = h->mode == HGB_NBODY_FAST;
    int block = 0, unroll = 0, warps = 0;
    auto_shape(count, s.dev->sms, &block, &unroll, &warps);
    if (h->block) { block = h->block; if (!h->warps && block != 64) warps = 24; }
    if (h->unroll) unroll = h->unroll;
    if (h->warps) warps = h->warps;
    const int grid = (count + block - 1) / block;
    cudaStream_t st = s.dev->stream;
#define HGB_V(B, U, W)                                                                             \
    if (block == B && unroll == U && warps == W) {                                                 \
        launch_variant<B, U, W>(fast, grid, st, src, h->npad, h->n, t0, count, tlist, out, out_list, run_if); \
        HGB_LAUNCHED();                                                                            \
        return HGB_OK;                                                                             \
    }
#define HGB_VW(B, U) HGB_V(B, U, 24) HGB_V(B, U, 32) HGB_V(B, U, 40)
    HGB_VW(128, 1) HGB_VW(128, 2) HGB_VW(128, 4)
    HGB_V(128, 1, 28) HGB_V(128, 2, 28) HGB_V(128, 4, 28)   // 7 CTAs per SM: 1036 slots
    HGB_VW(256, 1) HGB_VW(256, 2) HGB_VW(256, 4)
#ifdef HGB_EXPERIMENTAL_VARIANTS
    HGB_V(64, 2, 28) HGB_V(64, 2, 24) HGB_V(64, 4, 24) HGB_V(64, 1, 28)   // small target counts: measured the same or slower (profiles/r2a_nbody_small_sweep.txt)
#endif
#undef HGB_VW
#undef HGB_V
    set_error("no nbody kernel variant block=%d unroll=%d warps=%d", block, unroll, warps);
    return HGB_ERR_ARG;
}

// ---- pair-symmetric step (nbody_sym.cuh): one GPU, EXACT mode, all n targets --------------
static int g_default_sym = -1, g_default_sym_min_n = -1;
static bool g_default_sym_auto = false;                    // HGB_NBODY_SYM unset or 1: the variant follows the GPU count   // -1: read HGB_NBODY_SYM / HGB_NBODY_SYM_MIN_N
constexpr int kSymDefaultVariant = 242;                    // R = 2, U = 4, 2 CTAs (8 warps) per SM: profiles/r1e_nbody_sym_sweep_*.txt
constexpr int kSymDefaultMinN = 393216;                    // measured: slower than the standard kernel at 2^18, 26 % faster at 2^19

// third generation (nbody_crow.cuh, a CTA of W warps per row): 30PC (W = 4) / 31PC (W = 8), P = profiling, C = CTAs per SM
static bool crow_variant(int v) {
    const int c = v % 10, p = v / 10 % 10, w = v / 100 % 10;   // p: 1 profiling, 2 row pass fused into the next force block, 3 both
#ifdef HGB_EXPERIMENTAL_VARIANTS   // measured and rejected: row pass fused into the next force block (p >= 2: -20 %), W = 8 warps per row (w == 1)
    return v >= 3001 && v <= 3132 && p <= 3 && ((w == 0 && (c == 1 || c == 2)) || (w == 1 && c == 1 && p <= 1));
#else
    return v >= 3001 && v <= 3012 && p <= 1 && w == 0 && (c == 1 || c == 2);
#endif
}
// second or third generation: 200C / 201C (profiling), or a cooperative-row variant (same buffers, plan and multi-GPU dealing)
// 202C: row pass fused into the next minitile's force block (ROWPIPE)
#ifdef HGB_EXPERIMENTAL_VARIANTS
constexpr int kTriLastVariant = 2023;
#else
constexpr int kTriLastVariant = 2013;   // 202C (+0.3 %) is not built
#endif
static bool tri_variant(int v) { return (v >= 2001 && v <= kTriLastVariant && (v % 10) >= 1 && (v % 10) <= 3) || crow_variant(v); }

// 4242 (5242 with span counters): the first generation's systolic warp (R = 2, U = 4, 2 CTAs per SM) on the segmented schedule
static bool seg_variant(int v) {   // 6242 / 7242: pipelined groups
    if (v == 4242 || v == 5242 || v == 6242 || v == 7242) return true;
#ifdef HGB_EXPERIMENTAL_VARIANTS       // measured and rejected: one CTA per SM (6241 / 7241, diagnostic), U = 2 with 3 / 4 CTAs per SM
    if (v == 6241 || v == 7241 || v == 6223 || v == 6224) return true;
#endif
    return false;
}

static bool sym_variant_known(int v) {
    if (tri_variant(v) || seg_variant(v)) return true;
    switch (v) {
        case 124: case 146: case 148: case 224: case 226: case 228: case 243: case 413: case 423: case 422: case 242: case 183: case 182: case 144: case 241: case 181: case 421: case 1226: case 1242:
            return true;
    }
    return false;
}

template <int R, int U, int CTAS, bool PROF = false>
static int launch_sym_variant(cudaStream_t st, int sms, const double *src, int npad, int n, ulonglong2 *ll,
                              unsigned epoch_tag, int *ctl) {
    const size_t smem = (size_t)kSymWarps * 6 * 3 * 32 * sizeof(double);   // per warp: 6 arrays x 3 chunks of 32 (18 KB per CTA)
    // a worker that has caught up with its predecessor falls one chunk behind before it goes on: 1384-1394 ms against
    // 1413-1420 at N = 2^20 in three runs (profiles/r3r_sym_hysteresis.txt, r3s_*); more than one chunk costs time
    static const int hyst = getenv("HGB_NBODY_SYM_HYST") ? atoi(getenv("HGB_NBODY_SYM_HYST")) : 1;
    const char *h_now = getenv("HGB_NBODY_SYM_HYST_NOW");   // experiments: re-read at every launch
    nbody_sym_kernel<R, U, CTAS, PROF><<<sms * CTAS, kSymWarps * 32, smem, st>>>(src, npad, n, ll, epoch_tag, ctl, h_now ? atoi(h_now) : hyst);
    HGB_LAUNCHED();
    return HGB_OK;
}

// The hand-off tag is (epoch << 21) | (worker + 1): the worker count ceil(n / (32 R)) must stay below
// 2^21 or worker ids spill into the epoch bits and tags of different workers alias (R >= 1, so
// n < 2^26 is safe for every variant); larger sets take the standard kernel.
constexpr int kSymMaxN = 1 << 26;
static_assert((kSymMaxN / 32) < (1 << 21) + 1, "worker id must fit the 21-bit tag field");

// Multi-GPU default: the second-generation kernel, when every GPU has at least this many bodies' worth of
// rows to overlap the wavefront's ramp with (measured at N = 2^20: +32 % / +29 % / +11 % over the standard
// kernel with ncclAllGather on 2 / 4 / 8 GPUs, profiles/r2j_tri_multi_rows_weight_g8_g4.txt; the ramp grows
// with N and the work per GPU with N^2 / g, so smaller sets on many GPUs stay with the standard kernel)
constexpr int kTriDefaultVariant = 2002;
constexpr int kTriMinNPerGpu = 131072;

// Automatic choice (HGB_NBODY_SYM unset or 1), measured on B200 (profiles/r3g_small_n.txt, r3c_*, r2j_*).  What
// decides is the number of bodies per GPU: it sets how long the wavefronts' serial chains are against the work.
//   one GPU:   n < 49 152 standard kernel | < 229 376 cooperative rows (nbody_crow_kernel: +19 % at 64k, +31 % at 96k)
//              | < 655 360 one warp per row (nbody_tri_kernel: +33 % at 384k, 374 vs 405 ms at 512k)
//              | else the first generation's systolic warp on the segmented schedule, groups software-pipelined
//                (nbody_seg_kernel: 1364 ms at 2^20 against 1410 for nbody_sym_kernel and 1440 for nbody_tri_kernel; 789 vs 830 at 786 432)
//   g GPUs:    n / g < 65 536 standard kernel with ncclAllGather | < 262 144 cooperative rows | else one warp per row
constexpr int kCrowDefaultVariant = 3002;
constexpr int kSegDefaultVariant = 6242;
static int auto_variant(const hgb_nbody *h) {
    const long long per = h->n / h->world;
    if (h->world == 1) return h->n >= 655360 ? kSegDefaultVariant : h->n >= 229376 ? kTriDefaultVariant : h->n >= 49152 ? kCrowDefaultVariant : 0;
    return per >= 262144 ? kTriDefaultVariant : per >= 65536 ? kCrowDefaultVariant : 0;
}

// the pair-symmetric variant the next step of this handle runs (0 = the standard kernel)
static int sym_variant_in_effect(const hgb_nbody *h) {
    if (h->sym == 0 || h->mode != HGB_NBODY_EXACT || h->n >= kSymMaxN) return 0;
    // sym_auto 2: the library's default (the whole ladder); 1: hgb_nbody_set_symmetric(h, 1, min_n) -- the first
    // generation on one GPU from min_n up (what the caller asked for), the ladder across GPUs
    const int v = (h->sym_auto == 2 || (h->sym_auto == 1 && h->world > 1)) ? auto_variant(h) : h->sym;
    if (v == 0 || h->n < h->sym_min_n) return 0;
    if (!tri_variant(v) && !seg_variant(v)) return h->world == 1 ? v : 0;   // first generation: one GPU
    if (h->world == 1) return v;
    // rows dealt over GPUs: peer-mapped buffers -- of this process (peer-store handles) or of the other ranks (CUDA IPC)
    if (!h->rank_mode) return h->exchange == HGB_EXCHANGE_P2P ? v : 0;
    return h->tri_connected >= 0 ? v : 0;
}
static bool sym_applies(const hgb_nbody *h) { return sym_variant_in_effect(h) != 0; }
// runs through the multi-GPU wavefront framework (peer-mapped double-buffered chain words, polling finish kernel): the second and
// third generation always, the segmented first generation when its blocks are dealt over several GPUs
static bool tri_framework(const hgb_nbody *h, int v) { return tri_variant(v) || (seg_variant(v) && h->world > 1); }

// ---- second-generation pair-symmetric step (nbody_tri.cuh) ------------------------------------------
constexpr int kTriDefaultSeg = 16384;     // sources per item
// rows per tile (rows of a tile start together and then stagger themselves; a tile-row is the unit dealt over
// GPUs) and weight of the segment index in the ticket order (key = tile-row + weight * segment): one GPU /
// several (profiles/r2e_tri_sweep.txt, r2j_tri_multi_rows_weight_g8_g4.txt)
// cooperative rows (nbody_crow_kernel, 4x fewer workers): rows handed out ONE at a time -- rows of a tile take
// consecutive tickets and then follow each other at the minimum distance, where every small delay of the row
// above costs a full L2 round trip (profiles/r3i_crow_schedule_*.txt: 1537 vs 1618 ms at N = 2^20)
static int tri_default_rows(const hgb_nbody *h, int v) { return seg_variant(v) ? 4 : crow_variant(v) ? (h->world > 1 ? 8 : 1) : h->world > 1 ? 16 : 8; }
static int tri_default_weight(const hgb_nbody *h, int v) { return seg_variant(v) ? 64 : crow_variant(v) ? (h->world > 1 ? 32 : 128) : h->world > 1 ? 64 : 32; }
// rows per turn of a GPU (a multiple of the rows per tile): chain words cross NVLink once per turn.  Across GPUs the
// single-row hand-out does NOT carry over (8 GPUs, N = 2^20: 1122 ms with rows 1 / weight 128 against 220 ms with
// rows 8 / weight 32 / turns of 8, profiles/r3k_multi_g8.txt): a GPU then holds only every 8th turn of rows and the
// ticket order of one GPU no longer follows the anti-diagonal the others are on.
static int tri_default_deal(const hgb_nbody *h, int v) { return h->world > 1 ? (seg_variant(v) || crow_variant(v) ? 8 : 16) : 1; }   // seg: blocks of 64 targets
struct TriPlanParams { int seg, tr, weight, deal; };
static TriPlanParams tri_params(const hgb_nbody *h, int v) {
    TriPlanParams p;
    // cooperative rows on one GPU: 8192-source items (7.93 vs 8.13 ms at 64k, 26.6 vs 26.7 at 128k, 1527 vs 1537 at 2^20)
    p.seg = h->tri_seg > 0 ? h->tri_seg : (crow_variant(v) && h->world == 1) ? 8192 : kTriDefaultSeg;
    p.tr = h->tri_rows > 0 ? h->tri_rows : tri_default_rows(h, v);
    p.weight = h->tri_weight > 0 ? h->tri_weight : tri_default_weight(h, v);
    p.deal = h->tri_deal > 0 ? h->tri_deal : std::max(p.tr, tri_default_deal(h, v));
    p.deal = std::max(p.tr, p.deal / p.tr * p.tr);   // a tile never straddles two GPUs
    return p;
}
constexpr long long kTriSpinLimit = 10000000000ll;   // ~5 s of SM clock: a poll that long means a lost hand-off

// Tiles (b, k): the `tr` rows of tile-row b against segment k, for every segment some row of b reaches.
// Item (I, k) needs (I, k-1) finished and streams the words of (I-1, k), or -- where row I-1 ends before
// segment k -- of row I-1's last item (I-1, k-1).  Ticket order: key b + weight * k ascending (then k): all
// three predecessors have a smaller key, so the order is a linear extension of the dependencies for any
// weight >= 1; the weight spaces the segments of one row far enough apart in the ticket stream that item
// (I, k-1) is finished when (I, k) is handed out.  Tile-row b lives on shard b mod world.
static int tri_plan(hgb_nbody *h, NbShard &s, int variant) {
    const TriPlanParams pp = tri_params(h, variant);
    const int seg = pp.seg, tr = pp.tr, weight = pp.weight, deal = pp.deal;
    const int RS = seg_variant(variant) ? 64 : 32;   // targets per row (a block of nbody_seg_kernel has two per lane)
    if (s.tri_tiles && s.tri_plan_seg == seg && s.tri_plan_rows == tr && s.tri_plan_weight == weight && s.tri_plan_deal == deal && s.tri_plan_rs == RS) return HGB_OK;
    const int T = (h->n + RS - 1) / RS, TB = (T + tr - 1) / tr;
    std::vector<std::pair<long long, int2>> keyed;
    for (int b = 0; b < TB; ++b) {
        if (((long long)b * tr / deal) % h->world != s.index) continue;   // rows I with I / deal = index (mod GPUs)
        const long long last_row = std::min(T - 1, b * tr + tr - 1);
        const int kmax = (int)((long long)RS * last_row / seg);
        for (int k = 0; k <= kmax; ++k) keyed.push_back({((long long)b + (long long)weight * k) * 4096 + k, make_int2(b, k)});
    }
    std::sort(keyed.begin(), keyed.end(), [](const auto &x, const auto &y) { return x.first < y.first; });
    std::vector<int2> tiles(keyed.size());
    for (size_t i = 0; i < keyed.size(); ++i) tiles[i] = keyed[i].second;
    if (s.tri_tiles) cudaFree(s.tri_tiles);
    s.tri_tiles = nullptr;
    HGB_CUDA(cudaMalloc(&s.tri_tiles, sizeof(int2) * (tiles.size() + 1)));
    HGB_CUDA(cudaMemcpyAsync(s.tri_tiles, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, s.dev->stream));
    HGB_CUDA(cudaStreamSynchronize(s.dev->stream));   // `tiles` goes out of scope
    s.tri_ntiles = (int)tiles.size();
    s.tri_plan_seg = seg; s.tri_plan_rows = tr; s.tri_plan_weight = weight; s.tri_plan_deal = deal; s.tri_plan_rs = RS;
    return HGB_OK;
}

static int tri_alloc(hgb_nbody *h) {
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        if (!s.tri_ctl) {
            const size_t words = 48ull * h->npad;
            HGB_CUDA(cudaMalloc(&s.tri_ctl, sizeof(int) * kTriCtlInts));
            HGB_CUDA(cudaMemsetAsync(s.tri_ctl, 0, sizeof(int) * kTriCtlInts, s.dev->stream));
            for (int b = 0; b < 2; ++b) {
                HGB_CUDA(cudaMalloc(&s.tri_ll[b], words));
                HGB_CUDA(cudaMemsetAsync(s.tri_ll[b], 0, words, s.dev->stream));   // tag 0 is never expected
            }
            HGB_CUDA(cudaMalloc(&s.tri_racc, words));
            HGB_CUDA(cudaMemsetAsync(s.tri_racc, 0, words, s.dev->stream));
        }
    }
    // peers store into these buffers: they must be zeroed before anybody's first step
    for (NbShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// what one rank tells the others about its buffers (CUDA IPC handles are plain 64-byte tokens)
struct TriExport {
    cudaIpcMemHandle_t ll[2], ctl;
    int ok;
    int pad[15];
};

// Fill the peer tables.  One process per GPU: every rank exports its two chain-word buffers and its control
// words (cudaIpcGetMemHandle), the handles travel by ncclAllGather on the rank communicator, and every rank
// maps the others' (cudaIpcOpenMemHandle, peer access enabled lazily).  COLLECTIVE over the ranks.  When any
// rank cannot export or map (a container without shared IPC namespace, no peer access) all ranks agree to
// keep the standard kernel: tri_connected = -1, not an error.
static int tri_connect(hgb_nbody *h) {
    if (h->tri_connected) return HGB_OK;
    HGB_TRY(tri_alloc(h));
    if (!h->rank_mode || h->world == 1) {
        for (NbShard &s : h->sh) {
            h->tri_peer_ll[0][s.index] = s.tri_ll[0]; h->tri_peer_ll[1][s.index] = s.tri_ll[1];
            h->tri_peer_ctl[s.index] = s.tri_ctl;
        }
        h->tri_connected = 1;
        return HGB_OK;
    }
    Runtime &r = rt();
    NbShard &s = h->sh[0];
    cudaStream_t st = s.dev->stream;
    HGB_CUDA(cudaSetDevice(s.dev->device));
    std::vector<TriExport> all(h->world);
    TriExport mine;
    memset(&mine, 0, sizeof mine);
    mine.ok = cudaIpcGetMemHandle(&mine.ll[0], s.tri_ll[0]) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.ll[1], s.tri_ll[1]) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine.ctl, s.tri_ctl) == cudaSuccess;
    (void)cudaGetLastError();
    TriExport *d = nullptr;
    HGB_CUDA(cudaMalloc(&d, sizeof(TriExport) * h->world));
    HGB_CUDA(cudaMemcpyAsync(d + r.rank, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
    HGB_NCCL(nccl().AllGather(d + r.rank, d, sizeof(TriExport), ncclChar, r.rank_comm, st));
    HGB_CUDA(cudaMemcpyAsync(all.data(), d, sizeof(TriExport) * h->world, cudaMemcpyDeviceToHost, st));
    HGB_CUDA(cudaStreamSynchronize(st));
    bool ok = true;
    for (const TriExport &e : all) ok = ok && e.ok;
    std::vector<void *> opened;
    if (ok) {
        for (int p = 0; p < h->world && ok; ++p) {
            if (p == r.rank) {
                h->tri_peer_ll[0][p] = s.tri_ll[0]; h->tri_peer_ll[1][p] = s.tri_ll[1]; h->tri_peer_ctl[p] = s.tri_ctl;
                continue;
            }
            void *m[3] = {nullptr, nullptr, nullptr};
            const cudaIpcMemHandle_t *hd[3] = {&all[p].ll[0], &all[p].ll[1], &all[p].ctl};
            for (int i = 0; i < 3 && ok; ++i) {
                ok = cudaIpcOpenMemHandle(&m[i], *hd[i], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
                if (ok) opened.push_back(m[i]);
            }
            h->tri_peer_ll[0][p] = (ulonglong2 *)m[0]; h->tri_peer_ll[1][p] = (ulonglong2 *)m[1]; h->tri_peer_ctl[p] = (int *)m[2];
        }
        (void)cudaGetLastError();
    }
    // every rank must take the same path: sum of the failures over the ranks (also the barrier behind which
    // every rank's buffers are known to be zeroed and mapped)
    double fails = ok ? 0.0 : 1.0;
    double *dsum = reinterpret_cast<double *>(d);
    HGB_CUDA(cudaMemcpyAsync(dsum, &fails, sizeof fails, cudaMemcpyHostToDevice, st));
    HGB_NCCL(nccl().AllReduce(dsum, dsum, 1, ncclDouble, ncclSum, r.rank_comm, st));
    HGB_CUDA(cudaMemcpyAsync(&fails, dsum, sizeof fails, cudaMemcpyDeviceToHost, st));
    HGB_CUDA(cudaStreamSynchronize(st));
    HGB_CUDA(cudaFree(d));
    if (fails != 0.0) {
        for (void *m : opened) cudaIpcCloseMemHandle(m);
        (void)cudaGetLastError();
        memset(h->tri_peer_ll, 0, sizeof h->tri_peer_ll);
        memset(h->tri_peer_ctl, 0, sizeof h->tri_peer_ctl);
        h->tri_connected = -1;
        return HGB_OK;
    }
    h->tri_connected = 1;
    return HGB_OK;
}

// unmap the other ranks' buffers (before the owners free them: hgb_nbody_destroy puts a barrier in between)
static void tri_disconnect(hgb_nbody *h) {
    if (h->tri_connected == 1 && h->rank_mode && h->world > 1) {
        for (int p = 0; p < h->world; ++p) {
            if (p == rt().rank) continue;
            for (int b = 0; b < 2; ++b) if (h->tri_peer_ll[b][p]) cudaIpcCloseMemHandle(h->tri_peer_ll[b][p]);
            if (h->tri_peer_ctl[p]) cudaIpcCloseMemHandle(h->tri_peer_ctl[p]);
        }
        (void)cudaGetLastError();
    }
    h->tri_connected = 0;
}

// shards that share one physical device (HGB_VIRTUAL_GPUS) must fit their persistent grids side by side
static int tri_shards_on_device(const hgb_nbody *h, const NbShard &s) {
    int c = 0;
    for (const NbShard &o : h->sh) c += o.dev->device == s.dev->device;
    return c;
}

template <int W, int CTAS, bool PROF, bool ROWPIPE = false>
static int launch_crow_variant(const TriArgs &a, int grid, cudaStream_t st) {
    auto kern = nbody_crow_kernel<W, CTAS, PROF, ROWPIPE>;
    static bool attr_done[16] = {false};   // per device: the attribute belongs to the function ON a device
    int dev = 0;
    HGB_CUDA(cudaGetDevice(&dev));
    if (!attr_done[dev & 15]) {
        HGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)crow_smem_bytes<W>()));
        attr_done[dev & 15] = true;
    }
    kern<<<grid, W * 32, crow_smem_bytes<W>(), st>>>(a);
    HGB_LAUNCHED();
    return HGB_OK;
}

template <int CTAS, bool PROF, bool ROWPIPE = false>
static int launch_tri_variant(const TriArgs &a, int grid, cudaStream_t st) {
    auto kern = nbody_tri_kernel<CTAS, PROF, ROWPIPE>;
    static bool attr_done[16] = {false};   // per device: the attribute belongs to the function ON a device
    int dev = 0;
    HGB_CUDA(cudaGetDevice(&dev));
    if (!attr_done[dev & 15]) {
        HGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTriSmem));
        attr_done[dev & 15] = true;
    }
    kern<<<grid, kTriWarps * 32, kTriSmem, st>>>(a);
    HGB_LAUNCHED();
    return HGB_OK;
}

// One step on every shard of this process: tickets reset, wavefront kernels (all launched before any finish
// kernel), G*acc of all n bodies from the final words (read from the shard that ran the last row),
// conditional redo by the standard kernel.  Every shard ends with the complete next state: no exchange follows.
//
// Nothing a PEER writes is ever reset while that peer could still write it, so the shards (or ranks) need no
// barrier between steps: chain words carry the step's epoch and live in the buffer of the step's parity; a
// shard can only be a step ahead of another (it needs the final words, which need every row); the range flag
// of step s lives in slot 4 + parity(s), and slot 4 + parity(s+1) is cleared at the START of step s -- after
// step s-1 has finished everywhere (its last user) and before anybody can be in step s+1.
static int launch_tri_step(hgb_nbody *h, int variant, int cur, int nxt) {
    const unsigned step = h->tri_step++;
    const int parity = step & 1;
    const unsigned epoch_tag = (step % 2047 + 1) << 21;
    const int RS = h->sh[0].tri_plan_rs;   // targets per row: 32, or 64 for the segmented first-generation blocks
    const int seg = h->sh[0].tri_plan_seg, rows = h->sh[0].tri_plan_rows, deal = h->sh[0].tri_plan_deal, T = (h->n + RS - 1) / RS;
    const int ctas = variant % 10;
    const bool prof = (variant / 10) % 10 == 1 || (variant / 10) % 10 == 3;
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaMemsetAsync(s.tri_ctl, 0, sizeof(int), s.dev->stream));                           // ticket
        HGB_CUDA(cudaMemsetAsync(s.tri_ctl + 4 + (parity ^ 1), 0, sizeof(int), s.dev->stream));        // next step's range flag
        HGB_CUDA(cudaMemsetAsync(s.tri_ctl + kCtlSpans, 0, kCtlSpanInts * sizeof(int), s.dev->stream)); // profiling counters (all 7 that hgb_nbody_sym_counters reads)
        if (seg_variant(variant) && (variant / 1000 == 5 || variant / 1000 == 7))   // 5242 / 7242 count polls too: per step, like the spans
            HGB_CUDA(cudaMemsetAsync(s.tri_ctl + kCtlSegProf0, 0, kCtlSegProfInts * sizeof(int), s.dev->stream));
    }
    // Shards that SHARE a device (HGB_VIRTUAL_GPUS, tests on a one-GPU box) do join their streams between
    // steps: a shard's finish / redo / memset kernels (little shared memory) are not scheduled onto SMs
    // that already run another shard's next wavefront kernel (large dynamic shared-memory carve-out), which
    // in turn waits for this shard's next wavefront kernel -- a deadlock of the block scheduler, not of the
    // algorithm (measured: the lagging shard's kernel did not start within the 5 s spin limit).
    static const bool step_sync_env = getenv("HGB_NBODY_TRI_STEP_SYNC") != nullptr;
    bool shared_device = false;
    for (const NbShard &s : h->sh) shared_device = shared_device || tri_shards_on_device(h, s) > 1;
    if ((step_sync_env || shared_device) && h->sh.size() > 1) {
        for (NbShard &s : h->sh) {
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaEventRecord(s.dev->ev_step, s.dev->stream));
        }
        for (NbShard &s : h->sh)
            for (NbShard &o : h->sh)
                if (&o != &s) HGB_CUDA(cudaStreamWaitEvent(s.dev->stream, o.dev->ev_step, 0));
    }
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        TriArgs a;
        memset(&a, 0, sizeof a);
        a.src = s.state[cur]; a.npad = h->npad; a.n = h->n; a.T = T; a.seg = seg; a.tr = rows; a.deal = deal;
        a.tiles = s.tri_tiles; a.ntiles = s.tri_ntiles;
        a.ll = s.tri_ll[parity];
        a.ll_down = h->tri_peer_ll[parity][(s.index + 1) % h->world];   // tile-row b+1 lives on the next shard
        a.npeers = h->world;
        for (int p = 0; p < h->world; ++p) a.ctl_all[p] = h->tri_peer_ctl[p];
        a.range_slot = 4 + parity;
        a.racc = s.tri_racc; a.epoch_tag = epoch_tag; a.ctl = s.tri_ctl; a.spin_limit = kTriSpinLimit;
        const int grid = std::max(1, s.dev->sms * ctas / tri_shards_on_device(h, s));
        int rc = HGB_ERR_ARG;
        if (seg_variant(variant)) {
            SegArgs g;
            memset(&g, 0, sizeof g);
            g.src = a.src; g.npad = a.npad; g.n = a.n; g.T = T; g.seg = seg; g.tr = rows; g.tiles = a.tiles; g.ntiles = a.ntiles;
            g.ll = a.ll; g.racc = a.racc; g.epoch_tag = a.epoch_tag; g.ctl = a.ctl; g.spin_limit = a.spin_limit;
            g.hyst = h->seg_hyst;
            g.ll_down = a.ll_down; g.deal = deal; g.npeers = a.npeers; g.range_slot = a.range_slot;
            for (int p = 0; p < h->world; ++p) g.ctl_all[p] = a.ctl_all[p];
            const size_t smem = (size_t)kSymWarps * 6 * 3 * 32 * sizeof(double);
            const int sgrid = std::max(1, s.dev->sms * 2 / tri_shards_on_device(h, s));
            switch (variant) {
                case 4242: nbody_seg_kernel<2, 4, 2, false, false><<<sgrid, kSymWarps * 32, smem, s.dev->stream>>>(g); break;
                case 5242: nbody_seg_kernel<2, 4, 2, true, false><<<sgrid, kSymWarps * 32, smem, s.dev->stream>>>(g); break;
                case 7242: nbody_seg_kernel<2, 4, 2, true, true><<<sgrid, kSymWarps * 32, smem, s.dev->stream>>>(g); break;
                default: nbody_seg_kernel<2, 4, 2, false, true><<<sgrid, kSymWarps * 32, smem, s.dev->stream>>>(g); break;
            }
            HGB_LAUNCHED();
            rc = HGB_OK;
        } else if (crow_variant(variant)) {
            switch (variant) {
                case 3001: rc = launch_crow_variant<4, 1, false>(a, grid, s.dev->stream); break;
                case 3002: rc = launch_crow_variant<4, 2, false>(a, grid, s.dev->stream); break;
                case 3011: rc = launch_crow_variant<4, 1, true>(a, grid, s.dev->stream); break;
                case 3012: rc = launch_crow_variant<4, 2, true>(a, grid, s.dev->stream); break;
#ifdef HGB_EXPERIMENTAL_VARIANTS
                case 3021: rc = launch_crow_variant<4, 1, false, true>(a, grid, s.dev->stream); break;
                case 3022: rc = launch_crow_variant<4, 2, false, true>(a, grid, s.dev->stream); break;
                case 3031: rc = launch_crow_variant<4, 1, true, true>(a, grid, s.dev->stream); break;
                case 3032: rc = launch_crow_variant<4, 2, true, true>(a, grid, s.dev->stream); break;
                case 3101: rc = launch_crow_variant<8, 1, false>(a, grid, s.dev->stream); break;
                case 3111: rc = launch_crow_variant<8, 1, true>(a, grid, s.dev->stream); break;
#endif
                default: set_error("no nbody_crow variant %d", variant);
            }
        } else
#ifdef HGB_EXPERIMENTAL_VARIANTS
        if ((variant / 10) % 10 == 2) {
            switch (ctas) {
                case 1: rc = launch_tri_variant<1, false, true>(a, grid, s.dev->stream); break;
                case 2: rc = launch_tri_variant<2, false, true>(a, grid, s.dev->stream); break;
                case 3: rc = launch_tri_variant<3, false, true>(a, grid, s.dev->stream); break;
                default: set_error("no nbody_tri variant %d", variant);
            }
        } else
#endif
        switch (ctas * 2 + (prof ? 1 : 0)) {
            case 2: rc = launch_tri_variant<1, false>(a, grid, s.dev->stream); break;
            case 3: rc = launch_tri_variant<1, true>(a, grid, s.dev->stream); break;
            case 4: rc = launch_tri_variant<2, false>(a, grid, s.dev->stream); break;
            case 5: rc = launch_tri_variant<2, true>(a, grid, s.dev->stream); break;
            case 6: rc = launch_tri_variant<3, false>(a, grid, s.dev->stream); break;
            case 7: rc = launch_tri_variant<3, true>(a, grid, s.dev->stream); break;
            default: set_error("no nbody_tri variant %d", variant);
        }
        HGB_TRY(rc);
    }
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        // a small persistent grid: while it polls for the final words it must not crowd out a wavefront
        // kernel that shares the device (virtual shards)
        const int fgrid = std::min((h->n + 255) / 256, std::max(1, s.dev->sms / tri_shards_on_device(h, s)));
        // the finished sums are the words of the LAST row, in the buffer of the shard that ran it
        const ulonglong2 *final_words = h->tri_peer_ll[parity][((T - 1) / deal) % h->world];
        nbody_tri_finish_kernel<<<fgrid, 256, 0, s.dev->stream>>>(final_words, h->npad, h->n,
                                                                 epoch_tag | (unsigned)T, s.state[nxt], s.tri_ctl,
                                                                 kTriSpinLimit);
        HGB_LAUNCHED();
        // never taken for sane input: an operand left the fast-path range -> the ordinary kernel redoes
        // the whole step on every shard from the untouched source state
        PeerOut out;
        out.count = 1;
        out.base[0] = s.state[nxt];
        HGB_TRY(launch_step(h, s, s.state[cur], 0, h->n, nullptr, out, nullptr, s.tri_ctl + 4 + parity));
    }
    return HGB_OK;
}

// ---- first-generation warp on the segmented schedule (nbody_seg.cuh): one GPU -----------------------------
constexpr int kSegDefaultSeg = 16384, kSegDefaultRows = 2, kSegDefaultWeight = 128;
static int seg_plan(hgb_nbody *h, NbShard &s) {
    const int seg = h->tri_seg > 0 ? h->tri_seg : kSegDefaultSeg;
    const int tr = h->tri_rows > 0 ? h->tri_rows : kSegDefaultRows;
    const int weight = h->tri_weight > 0 ? h->tri_weight : kSegDefaultWeight;
    if (s.seg_tiles && s.seg_plan_seg == seg && s.seg_plan_rows == tr && s.seg_plan_weight == weight) return HGB_OK;
    constexpr int BS = 64;   // targets per block (R = 2)
    const int T = (h->n + BS - 1) / BS, TB = (T + tr - 1) / tr;
    std::vector<std::pair<long long, int2>> keyed;
    for (int b = 0; b < TB; ++b) {
        const long long last_row = std::min(T - 1, b * tr + tr - 1);
        const int kmax = (int)(BS * last_row / seg);
        for (int k = 0; k <= kmax; ++k) keyed.push_back({((long long)b + (long long)weight * k) * 4096 + k, make_int2(b, k)});
    }
    std::sort(keyed.begin(), keyed.end(), [](const auto &x, const auto &y) { return x.first < y.first; });
    std::vector<int2> tiles(keyed.size());
    for (size_t i = 0; i < keyed.size(); ++i) tiles[i] = keyed[i].second;
    if (s.seg_tiles) cudaFree(s.seg_tiles);
    s.seg_tiles = nullptr;
    HGB_CUDA(cudaMalloc(&s.seg_tiles, sizeof(int2) * (tiles.size() + 1)));
    HGB_CUDA(cudaMemcpyAsync(s.seg_tiles, tiles.data(), sizeof(int2) * tiles.size(), cudaMemcpyHostToDevice, s.dev->stream));
    HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    s.seg_ntiles = (int)tiles.size();
    s.seg_plan_seg = seg; s.seg_plan_rows = tr; s.seg_plan_weight = weight;
    return HGB_OK;
}

// ctl[0 .. count) = 0 unless the abort flag ctl[2] is up; inject != 0 (test aid) raises it after the reset
__global__ void nbody_ctl_reset_unless_aborted_kernel(int *ctl, int count, int inject) {
    const bool aborted = ctl[2] != 0;
    __syncthreads();
    if (!aborted)
        for (int i = threadIdx.x; i < count; i += blockDim.x) ctl[i] = (inject && i == 2) ? 1 : 0;
}

static int launch_seg(hgb_nbody *h, int variant, NbShard &s, const double *src, const PeerOut &out) {
    cudaStream_t st = s.dev->stream;
    if (!s.sym_ctl) {
        HGB_CUDA(cudaMalloc(&s.sym_ctl, sizeof(int) * kSymCtlInts));
        HGB_CUDA(cudaMemsetAsync(s.sym_ctl, 0, sizeof(int) * kSymCtlInts, st));
        HGB_CUDA(cudaMalloc(&s.sym_ll, 48ull * h->npad));
        HGB_CUDA(cudaMemsetAsync(s.sym_ll, 0, 48ull * h->npad, st));   // tag 0 is never expected
    }
    if (!s.seg_racc) {
        HGB_CUDA(cudaMalloc(&s.seg_racc, 48ull * h->npad));
        HGB_CUDA(cudaMemsetAsync(s.seg_racc, 0, 48ull * h->npad, st));
    }
    HGB_TRY(seg_plan(h, s));
    // Per-step reset of ticket, range flag and counters -- unless an earlier step of this advance gave a poll up
    // (ctl[2]): then everything stays as that step left it, this step's workers see the flag and leave at once,
    // and nb_advance reports the FIRST abandoned poll after the last step (it reads ctl only there).
    nbody_ctl_reset_unless_aborted_kernel<<<1, kSymCtlInts, 0, st>>>(s.sym_ctl, kSymCtlInts,
                                                                    h->seg_inject_abort == (int)h->seg_step ? 1 : 0);
    HGB_LAUNCHED();
    ++h->seg_step;
    s.sym_epoch = s.sym_epoch % 2047 + 1;
    SegArgs a;
    memset(&a, 0, sizeof a);
    a.src = src; a.npad = h->npad; a.n = h->n; a.T = (h->n + 63) / 64; a.seg = s.seg_plan_seg; a.tr = s.seg_plan_rows;
    a.tiles = s.seg_tiles; a.ntiles = s.seg_ntiles; a.ll = s.sym_ll; a.racc = s.seg_racc;
    a.epoch_tag = s.sym_epoch << 21; a.ctl = s.sym_ctl; a.spin_limit = kTriSpinLimit;
    a.hyst = h->seg_hyst;
    a.ll_down = a.ll; a.deal = a.T + 1; a.npeers = 1; a.ctl_all[0] = s.sym_ctl; a.range_slot = 1;   // one GPU: nothing leaves it
    const size_t smem = (size_t)kSymWarps * 6 * 3 * 32 * sizeof(double);
    const int grid = s.dev->sms * 2;
#ifdef HGB_EXPERIMENTAL_VARIANTS
    if (variant == 6223) nbody_seg_kernel<2, 2, 3, false, true><<<s.dev->sms * 3, kSymWarps * 32, smem, st>>>(a);
    else if (variant == 6224) nbody_seg_kernel<2, 2, 4, false, true><<<s.dev->sms * 4, kSymWarps * 32, smem, st>>>(a);
    else if (variant == 6241) nbody_seg_kernel<2, 4, 1, false, true><<<s.dev->sms, kSymWarps * 32, smem, st>>>(a);
    else if (variant == 7241) nbody_seg_kernel<2, 4, 1, true, true><<<s.dev->sms, kSymWarps * 32, smem, st>>>(a);
    else
#endif
    if (variant == 5242) nbody_seg_kernel<2, 4, 2, true><<<grid, kSymWarps * 32, smem, st>>>(a);
    else if (variant == 6242) nbody_seg_kernel<2, 4, 2, false, true><<<grid, kSymWarps * 32, smem, st>>>(a);
    else if (variant == 7242) nbody_seg_kernel<2, 4, 2, true, true><<<grid, kSymWarps * 32, smem, st>>>(a);
    else nbody_seg_kernel<2, 4, 2, false><<<grid, kSymWarps * 32, smem, st>>>(a);
    HGB_LAUNCHED();
    nbody_sym_finish_kernel<<<(h->n + 255) / 256, 256, 0, st>>>(s.sym_ll, h->npad, h->n, out);
    HGB_LAUNCHED();
    return launch_step(h, s, src, 0, h->n, nullptr, out, nullptr, s.sym_ctl + 1);
}

// whole step on shard s: tickets/flags reset, wavefront kernel, G*acc, conditional redo
static int launch_sym(const hgb_nbody *h, int variant, NbShard &s, const double *src, const PeerOut &out) {
    cudaStream_t st = s.dev->stream;
    if (!s.sym_ctl) {
        HGB_CUDA(cudaMalloc(&s.sym_ctl, sizeof(int) * kSymCtlInts));
        HGB_CUDA(cudaMalloc(&s.sym_ll, 48ull * h->npad));
        HGB_CUDA(cudaMemsetAsync(s.sym_ll, 0, 48ull * h->npad, st));   // tag 0 is never expected
    }
    HGB_CUDA(cudaMemsetAsync(s.sym_ctl, 0, sizeof(int) * kSymCtlInts, st));
    s.sym_epoch = s.sym_epoch % 2047 + 1;                            // 1..2047: every launch its own tags
    const unsigned epoch_tag = s.sym_epoch << 21;
    ulonglong2 *ll = s.sym_ll;
    int sms = 0;
    HGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, s.dev->device));
    int rc = HGB_ERR_ARG;
    switch (variant) {
#define HGB_SYM(R_, U_, C_) case R_ * 100 + U_ * 10 + C_: rc = launch_sym_variant<R_, U_, C_>(st, sms, src, h->npad, h->n, ll, epoch_tag, s.sym_ctl); break;
        HGB_SYM(1, 2, 4) HGB_SYM(1, 4, 6) HGB_SYM(1, 4, 8) HGB_SYM(2, 2, 4) HGB_SYM(2, 2, 6) HGB_SYM(2, 2, 8)
        HGB_SYM(2, 4, 3) HGB_SYM(4, 1, 3) HGB_SYM(4, 2, 3) HGB_SYM(4, 2, 2) HGB_SYM(2, 4, 2) HGB_SYM(1, 8, 3)
        HGB_SYM(1, 8, 2) HGB_SYM(1, 4, 4) HGB_SYM(2, 4, 1) HGB_SYM(1, 8, 1) HGB_SYM(4, 2, 1)
#undef HGB_SYM
        // diagnostic instantiations (clock64 spans -> hgb_nbody_sym_counters): variant + 1000
        case 1226: rc = launch_sym_variant<2, 2, 6, true>(st, sms, src, h->npad, h->n, ll, epoch_tag, s.sym_ctl); break;
        case 1242: rc = launch_sym_variant<2, 4, 2, true>(st, sms, src, h->npad, h->n, ll, epoch_tag, s.sym_ctl); break;
        default: set_error("no pair-symmetric nbody variant %d", variant);
    }
    HGB_TRY(rc);
    nbody_sym_finish_kernel<<<(h->n + 255) / 256, 256, 0, st>>>(ll, h->npad, h->n, out);
    HGB_LAUNCHED();
    // never taken for sane input: an operand left the fast-path range -> ordinary kernel redoes the step
    return launch_step(h, s, src, 0, h->n, nullptr, out, nullptr, s.sym_ctl + 1);
}

static int nb_alloc(hgb_nbody *h) {
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const size_t bytes = 3ull * h->npad * sizeof(double);
        for (int b = 0; b < 2; ++b) {
            HGB_CUDA(cudaMalloc(&s.state[b], bytes));
            HGB_CUDA(cudaMemsetAsync(s.state[b], 0, bytes, s.dev->stream));
        }
        HGB_CUDA(cudaMalloc(&s.stage, bytes));
        HGB_CUDA(cudaMalloc(&s.partials, sizeof(double) * 2 * (h->slice / 64 + 2)));
    }
    return HGB_OK;
}

static void sym_defaults(hgb_nbody *h) {
    if (g_default_sym < 0) {
        const char *e = getenv("HGB_NBODY_SYM");   // unset: on (default variant); 0: off; else a variant
        int v = e ? atoi(e) : 1;
        g_default_sym_auto = v == 1;
        if (v == 1) v = kSymDefaultVariant;
        if (v == 2) v = 2002;
        g_default_sym = sym_variant_known(v) ? v : 0;
    }
    if (g_default_sym_min_n < 0) {
        const char *e = getenv("HGB_NBODY_SYM_MIN_N");
        g_default_sym_min_n = e ? atoi(e) : (g_default_sym_auto ? 0 : kSymDefaultMinN);   // automatic: the ladder above decides
        if (g_default_sym_min_n < 0) g_default_sym_min_n = 0;
    }
    h->sym = g_default_sym;
    h->sym_auto = g_default_sym_auto ? 2 : 0;
    h->sym_min_n = g_default_sym_min_n;
    if (const char *e = getenv("HGB_NBODY_SEG_HYST")) h->seg_hyst = atoi(e);
    if (const char *e = getenv("HGB_NBODY_SEG_INJECT_ABORT")) h->seg_inject_abort = atoi(e);
}

extern "C" {

int hgb_nbody_set_symmetric(hgb_nbody *h, int variant, int min_n) {
    HGB_ARG(h, "handle");
    const bool automatic = variant == 1;
    if (variant == 1) variant = kSymDefaultVariant;
    if (variant == 2) variant = 2002;
    HGB_ARG(variant == 0 || sym_variant_known(variant), "pair-symmetric variant (R*100 + U*10 + CTAs per SM, or 2000 + CTAs per SM)");
    HGB_ARG(min_n >= 0, "min_n");
    h->sym = variant;
    h->sym_auto = automatic ? 1 : 0;
    h->sym_min_n = min_n;
    return HGB_OK;
}

int hgb_nbody_symmetric(const hgb_nbody *h) { return h ? sym_variant_in_effect(h) : 0; }

int hgb_nbody_set_segment(hgb_nbody *h, int seg, int rows_per_tile, int order_weight) {
    HGB_ARG(h, "handle");
    HGB_ARG(seg == 0 || (seg >= 32 && seg % 32 == 0), "segment: a multiple of 32 sources (0 = default)");
    HGB_ARG(rows_per_tile >= 0 && rows_per_tile <= 4096 && order_weight >= 0 && order_weight <= 4096, "rows per tile / order weight");
    h->tri_seg = seg;
    h->tri_rows = rows_per_tile;
    h->tri_weight = order_weight;
    return HGB_OK;
}

int hgb_nbody_set_deal(hgb_nbody *h, int rows_per_turn) {
    HGB_ARG(h, "handle");
    HGB_ARG(rows_per_turn >= 0 && rows_per_turn <= 65536, "rows per turn of a GPU");
    h->tri_deal = rows_per_turn;
    return HGB_OK;
}

int hgb_nbody_describe(const hgb_nbody *h, char *buf, int capacity) {
    HGB_ARG(h && buf && capacity > 0, "handle / buffer");
    char kern[160];
    const bool fast = h->mode == HGB_NBODY_FAST;
    const int v = sym_variant_in_effect(h);
    if (crow_variant(v)) {
        snprintf(kern, sizeof kern, "nbody_crow_kernel<%d,%d%s> seg=%d rows=%d weight=%d", v / 100 % 10 ? 8 : 4, v % 10, (v / 10) % 10 >= 2 ? ",rowpipe" : "", tri_params(h, v).seg,
                 tri_params(h, v).tr, tri_params(h, v).weight);
    } else if (tri_variant(v)) {
        snprintf(kern, sizeof kern, "nbody_tri_kernel<%d%s> seg=%d rows=%d weight=%d", v % 10, (v / 10) % 10 == 2 ? ",rowpipe" : "", tri_params(h, v).seg,
                 tri_params(h, v).tr, tri_params(h, v).weight);
    } else if (v) {
        if (seg_variant(v))
            snprintf(kern, sizeof kern, "nbody_seg_kernel<2,4,2%s> seg=%d rows=%d weight=%d", v >= 6000 ? ",pipelined" : "",
                     h->world > 1 ? tri_params(h, v).seg : h->tri_seg > 0 ? h->tri_seg : kSegDefaultSeg,
                     h->world > 1 ? tri_params(h, v).tr : h->tri_rows > 0 ? h->tri_rows : kSegDefaultRows,
                     h->world > 1 ? tri_params(h, v).weight : h->tri_weight > 0 ? h->tri_weight : kSegDefaultWeight);
        else
            snprintf(kern, sizeof kern, "nbody_sym_kernel<%d,%d,%d>", v / 100, v / 10 % 10, v % 10);
    } else {
        int block = 0, unroll = 0, warps = 0;
        const int count = h->n < h->slice ? h->n : h->slice;
        auto_shape(count, h->sh[0].dev->sms, &block, &unroll, &warps);
        if (h->block) { block = h->block; if (!h->warps && block != 64) warps = 24; }
        if (h->unroll) unroll = h->unroll;
        if (h->warps) warps = h->warps;
        snprintf(kern, sizeof kern, "nbody_step_kernel<%d,%d,%d,%s>", block, unroll, warps, fast ? "true" : "false");
    }
    const char *xchg = h->world == 1 ? "one GPU, no exchange"
                     : tri_framework(h, v) ? (h->rank_mode ? "pair-symmetric: tile-rows of targets dealt round-robin over the ranks, chain words stored into the next rank's buffer over NVLink (CUDA IPC mapping), finished sums read from the last row's rank, no position exchange"
                                                      : "pair-symmetric: tile-rows of targets dealt round-robin, chain words stored into the next GPU's buffer over NVLink, finished sums read from the last row's GPU, no position exchange")
                     : h->exchange == HGB_EXCHANGE_P2P ? "targets sharded, results stored into every peer's next state over NVLink (fused epilogue)"
                                                       : "targets sharded, ncclAllGather of positions per step";
    snprintf(buf, (size_t)capacity, "%s | %s", kern, xchg);
    return HGB_OK;
}

double hgb_nbody_fp64_instr(const hgb_nbody *h) {
    if (!h) return 0.0;
    if (h->mode == HGB_NBODY_FAST) return 16.0;    // FMA-contracted rsqrt-Newton sequence
    if (tri_variant(sym_variant_in_effect(h))) return 16.5;   // 264 per 32 x 8 minitile: 26 + 3 per pair + 32 column instructions
    return sym_applies(h) ? 16.0 : 29.0;           // 32 per PAIR (26 + 3 + 3) vs 17 + 8 + 4 per interaction
}

int hgb_nbody_sym_counters(hgb_nbody *h, unsigned long long *out7) {
    HGB_ARG(h && out7, "handle / out");
    NbShard &s = h->sh[0];
    int *ctl = tri_framework(h, sym_variant_in_effect(h)) ? s.tri_ctl : s.sym_ctl;
    if (!ctl) { set_error("the pair-symmetric kernel has not run on this handle"); return HGB_ERR_STATE; }
    HGB_CUDA(cudaSetDevice(s.dev->device));
    HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    HGB_CUDA(cudaMemcpy(out7, ctl + kCtlSpans, 7 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
    if (getenv("HGB_NBODY_SYM_HIST") && ctl == s.sym_ctl && seg_variant(sym_variant_in_effect(h))) {
        unsigned long long polled[5] = {0};
        HGB_CUDA(cudaMemcpy(polled, ctl + kCtlSegPolled, sizeof polled[0], cudaMemcpyDeviceToHost));
        HGB_CUDA(cudaMemcpy(polled + 1, ctl + kCtlSegPolledQuarter, 4 * sizeof polled[0], cudaMemcpyDeviceToHost));
        fprintf(stderr, "seg: chunk opens that had to poll: %llu; by quarter of the blocks: %llu %llu %llu %llu\n", polled[0], polled[1], polled[2],
                polled[3], polled[4]);
        unsigned long long ahead[8] = {0};
        HGB_CUDA(cudaMemcpy(ahead, ctl + kCtlSegAhead, sizeof ahead, cudaMemcpyDeviceToHost));
        fprintf(stderr, "seg: chunks the block above has published when an item starts (0, 1, 2, 3, 4-7, 8-15, 16-31, 32+):");
        for (int e = 0; e < 8; ++e) fprintf(stderr, " %llu", ahead[e]);
        fprintf(stderr, "\n");
    } else if (getenv("HGB_NBODY_SYM_HIST") && ctl == s.sym_ctl) {   // diagnostic: wait cycles by eighth of the row x quarter of the workers
        unsigned long long hist[40];
        HGB_CUDA(cudaMemcpy(hist, ctl + kCtlSymHist, sizeof hist, cudaMemcpyDeviceToHost));
        fprintf(stderr, "sym: chunks the predecessor is ahead at a chunk open (0, 1, 2, 3, 4-7, 8-15, 16-31, 32+):");
        for (int e = 0; e < 8; ++e) fprintf(stderr, " %llu", hist[32 + e]);
        fprintf(stderr, "\n");
        for (int q = 0; q < 4; ++q) {
            fprintf(stderr, "sym wait cycles, workers in quarter %d, by eighth of the row:", q);
            for (int e = 0; e < 8; ++e) fprintf(stderr, " %.3e", (double)hist[q * 8 + e]);
            fprintf(stderr, "\n");
        }
    }
    return HGB_OK;
}

int hgb_set_nbody_mode(int mode) {
    HGB_ARG(mode == HGB_NBODY_EXACT || mode == HGB_NBODY_FAST, "mode");
    g_default_mode = mode;
    return HGB_OK;
}

int hgb_set_exchange(int kind) {
    HGB_ARG(kind == HGB_EXCHANGE_NCCL || kind == HGB_EXCHANGE_P2P, "exchange kind");
    g_default_exchange = kind;
    return HGB_OK;
}

int hgb_nbody_destroy(hgb_nbody *h) {
    if (!h) return HGB_OK;
    if (h->tri_connected == 1 && h->rank_mode && h->world > 1) {
        // the other ranks have this rank's buffers mapped: unmap theirs, then wait until everybody has
        // (COLLECTIVE over the ranks, like the steps that used the mapping)
        for (NbShard &s : h->sh) { cudaSetDevice(s.dev->device); cudaStreamSynchronize(s.dev->stream); }
        tri_disconnect(h);
        if (rt().rank_comm) hgb_comm_barrier();
    }
    for (NbShard &s : h->sh) {
        cudaSetDevice(s.dev->device);
        cudaStreamSynchronize(s.dev->stream);
        cudaFree(s.state[0]);
        cudaFree(s.state[1]);
        cudaFree(s.stage);
        cudaFree(s.partials);
        if (s.vel) cudaFree(s.vel);
        if (s.mass) cudaFree(s.mass);
        if (s.tlist) cudaFree(s.tlist);
        if (s.sym_ctl) cudaFree(s.sym_ctl);
        if (s.sym_ll) cudaFree(s.sym_ll);
        if (s.tri_ctl) cudaFree(s.tri_ctl);
        if (s.tri_ll[0]) cudaFree(s.tri_ll[0]);
        if (s.tri_ll[1]) cudaFree(s.tri_ll[1]);
        if (s.tri_racc) cudaFree(s.tri_racc);
        if (s.tri_tiles) cudaFree(s.tri_tiles);
        if (s.seg_racc) cudaFree(s.seg_racc);
        if (s.seg_tiles) cudaFree(s.seg_tiles);
    }
    delete h;
    return HGB_OK;
}

int hgb_nbody_create(hgb_nbody **out, int n, int ngpus) {
    HGB_ARG(out && n > 0, "out == NULL or n <= 0");
    int g = 1;
    HGB_TRY(resolve_ngpus(ngpus, &g));
    HGB_ARG(g <= kMaxPeers, "more than 8 GPUs");
    HGB_TRY(ensure_devices(g));
    hgb_nbody *h = new hgb_nbody();
    h->n = n;
    h->world = g;
    h->slice = shard_size(n, g, kAlign);
    h->npad = h->slice * g;
    h->mode = g_default_mode;
    h->exchange = g_default_exchange;
    sym_defaults(h);
    h->sh.resize(g);
    for (int i = 0; i < g; ++i) {
        h->sh[i].dev = &rt().devs[i];
        h->sh[i].index = i;
    }
    int rc = nb_alloc(h);
    if (rc == HGB_OK && g > 1 && (h->exchange == HGB_EXCHANGE_P2P || oversubscribed(g))) {
        h->exchange = HGB_EXCHANGE_P2P;   // virtual shards share a device: NCCL cannot be used
        rc = ensure_peer_access(g);
    }
    if (rc != HGB_OK) { hgb_nbody_destroy(h); return rc; }
    *out = h;
    return HGB_OK;
}

int hgb_nbody_create_rank(hgb_nbody **out, int n) {
    HGB_ARG(out && n > 0, "out == NULL or n <= 0");
    Runtime &r = rt();
    HGB_TRY(ensure_rank_dev());
    if (r.world > 1 && !r.rank_comm) {
        set_error("hgb_comm_init_rank must be called before hgb_nbody_create_rank");
        return HGB_ERR_STATE;
    }
    hgb_nbody *h = new hgb_nbody();
    h->n = n;
    h->world = r.world;
    h->rank_mode = true;
    h->slice = shard_size(n, r.world, kAlign);
    h->npad = h->slice * r.world;
    h->mode = g_default_mode;
    h->exchange = HGB_EXCHANGE_NCCL;
    sym_defaults(h);                 // applies when the communicator has a single rank
    h->sh.resize(1);
    h->sh[0].dev = &r.rank_dev;
    h->sh[0].index = r.rank;
    int rc = nb_alloc(h);
    if (rc != HGB_OK) { hgb_nbody_destroy(h); return rc; }
    *out = h;
    return HGB_OK;
}

int hgb_nbody_config(hgb_nbody *h, int mode, int exchange) {
    HGB_ARG(h, "handle");
    HGB_ARG(mode == HGB_NBODY_EXACT || mode == HGB_NBODY_FAST, "mode");
    HGB_ARG(exchange == HGB_EXCHANGE_NCCL || exchange == HGB_EXCHANGE_P2P, "exchange");
    if (!h->rank_mode && h->world > 1 && oversubscribed(h->world)) exchange = HGB_EXCHANGE_P2P;
    if (exchange == HGB_EXCHANGE_P2P && h->world > 1) {
        if (h->rank_mode) {
            set_error("peer-store exchange needs the single-process multi-GPU handle");
            return HGB_ERR_STATE;
        }
        HGB_TRY(ensure_peer_access(h->world));
    }
    h->mode = mode;
    h->exchange = exchange;
    return HGB_OK;
}

int hgb_nbody_tune(hgb_nbody *h, int block, int unroll, int warps) {
    HGB_ARG(h, "handle");
#ifdef HGB_EXPERIMENTAL_VARIANTS
    HGB_ARG(block == 0 || block == 64 || block == 128 || block == 256, "block in {64,128,256}");
#else
    HGB_ARG(block == 0 || block == 128 || block == 256, "block in {128,256} (64-thread CTAs: build with -DHGB_EXPERIMENTAL_VARIANTS)");
#endif
    HGB_ARG(unroll == 0 || unroll == 1 || unroll == 2 || unroll == 4, "unroll in {1,2,4}");
    HGB_ARG(warps == 0 || warps == 24 || warps == 28 || warps == 32 || warps == 40, "warps in {24,28,32,40}");
    HGB_ARG(warps != 28 || block != 256, "warps = 28 needs 64- or 128-thread CTAs");
    if (block == 64) {
        const int u = unroll ? unroll : 2, w = warps ? warps : 28;
        HGB_ARG((u == 2 && (w == 28 || w == 24)) || (u == 4 && w == 24) || (u == 1 && w == 28),
                "64-thread CTAs: (unroll, warps) in {(2,28), (2,24), (4,24), (1,28)}");
    }
    h->block = block;
    h->unroll = unroll;
    h->warps = warps;
    if (block || unroll || warps) h->sym = 0;   // an explicit variant of the standard kernel was asked for
    return HGB_OK;
}

int hgb_nbody_upload(hgb_nbody *h, const double *pos_aos) {
    HGB_ARG(h && pos_aos, "handle / positions");
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaMemcpyAsync(s.stage, pos_aos, 3ull * h->n * sizeof(double),
                                 cudaMemcpyHostToDevice, s.dev->stream));
        aos_to_soa_kernel<<<(h->npad + 255) / 256, 256, 0, s.dev->stream>>>(s.stage, s.state[0],
                                                                            h->n, h->npad);
        HGB_LAUNCHED();
    }
    for (NbShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    h->cur = 0;
    h->uploaded = true;
    return HGB_OK;
}

// `iterations` x { step kernel on every shard; exchange the new positions }.  dyn: the integrator
// extension (mass-weighted sources, velocity/position update in the epilogue) instead of accel().
static int nb_advance(hgb_nbody *h, int iterations, float *ms, bool dyn) {
    const bool multi = h->world > 1;
    const bool p2p = multi && h->exchange == HGB_EXCHANGE_P2P;
    if (multi && !p2p && !h->rank_mode) HGB_TRY(ensure_sp_comm(h->world));
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaEventRecord(s.dev->ev_begin, s.dev->stream));
    }
    if (!dyn && tri_framework(h, sym_variant_in_effect(h))) {
        HGB_TRY(tri_connect(h));   // first use: buffers, peer tables (collective over the ranks); may turn the variant off
        const int v = sym_variant_in_effect(h);
        if (tri_framework(h, v)) for (NbShard &s : h->sh) HGB_TRY(tri_plan(h, s, v));
    }
    const int variant = dyn ? 0 : sym_variant_in_effect(h);
    const bool tri = tri_framework(h, variant);
    for (int it = 0; it < iterations; ++it) {
        const int cur = h->cur, nxt = cur ^ 1;
        if (tri) {   // rows dealt over the shards, every shard ends with the whole next state
            HGB_TRY(launch_tri_step(h, variant, cur, nxt));
            h->cur = nxt;
            continue;
        }
        for (NbShard &s : h->sh) {
            HGB_CUDA(cudaSetDevice(s.dev->device));
            PeerOut out;
            out.count = 1;
            out.base[0] = s.state[nxt];
            if (p2p) {
                out.count = h->world;
                for (int p = 0; p < h->world; ++p) out.base[p] = h->sh[p].state[nxt];
            }
            const int t0 = s.index * h->slice;
            const int count = max(0, min(h->n, t0 + h->slice) - t0);
            if (dyn) {
                if (count > 0) {
                    const Dynamics d{s.mass, s.vel, h->dt};
                    nbody_step_kernel<128, 2, 24, false, true><<<(count + 127) / 128, 128, 0, s.dev->stream>>>(
                        s.state[cur], h->npad, h->n, t0, count, nullptr, out, nullptr, d);
                    HGB_LAUNCHED();
                }
            } else if (seg_variant(variant)) {   // first-generation warp, segmented schedule
                HGB_TRY(launch_seg(h, variant, s, s.state[cur], out));
            } else if (variant) {   // first-generation wavefront (variants R*100 + U*10 + CTAs)
                HGB_TRY(launch_sym(h, variant, s, s.state[cur], out));
            } else {
                HGB_TRY(launch_step(h, s, s.state[cur], t0, count, nullptr, out, nullptr));
            }
        }
        if (p2p) {
            // every device's next step reads what all peers stored: join the streams
            for (NbShard &s : h->sh) {
                HGB_CUDA(cudaSetDevice(s.dev->device));
                HGB_CUDA(cudaEventRecord(s.dev->ev_step, s.dev->stream));
            }
            for (NbShard &s : h->sh)
                for (NbShard &o : h->sh)
                    if (&o != &s) HGB_CUDA(cudaStreamWaitEvent(s.dev->stream, o.dev->ev_step, 0));
        } else if (multi) {
            HGB_NCCL(nccl().GroupStart());
            for (NbShard &s : h->sh) {
                ncclComm_t comm = h->rank_mode ? rt().rank_comm : sp_comm(h->world, s.index);
                for (int c = 0; c < 3; ++c) {
                    double *recv = s.state[nxt] + (size_t)c * h->npad;
                    HGB_NCCL(nccl().AllGather(recv + (size_t)s.index * h->slice, recv, h->slice,
                                              ncclDouble, comm, s.dev->stream));
                }
            }
            HGB_NCCL(nccl().GroupEnd());
        }
        h->cur = nxt;
    }
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaEventRecord(s.dev->ev_end, s.dev->stream));
    }
    float worst = 0.f;
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaEventSynchronize(s.dev->ev_end));
        float t = 0.f;
        HGB_CUDA(cudaEventElapsedTime(&t, s.dev->ev_begin, s.dev->ev_end));
        worst = t > worst ? t : worst;
    }
    HGB_CUDA(cudaGetLastError());
    if (ms) *ms = worst;
    if (seg_variant(variant) && !tri && iterations > 0) {
        NbShard &s0 = h->sh[0];
        int c[32] = {0};
        HGB_CUDA(cudaMemcpy(c, s0.sym_ctl, sizeof c, cudaMemcpyDeviceToHost));
        if (c[2]) {
            HGB_CUDA(cudaMemset(s0.sym_ctl + kCtlAbort, 0, 2 * sizeof(int)));   // reported: the handle can be uploaded to and stepped again
            set_error("segmented pair-symmetric wavefront abandoned: a hand-off did not arrive within the spin limit; %d tickets of %d items, "
                      "wanted tag epoch %u id %d, found tag %#x, chain word %d, row-sum word %d", c[0], s0.seg_ntiles * s0.seg_plan_rows,
                      (unsigned)c[6] >> 21, c[6] & 0x1fffff, (unsigned)c[7], c[25], c[26]);
            return HGB_ERR_STATE;
        }
    }
    if (tri && iterations > 0) {   // a hand-off that never arrived: the kernels gave up instead of hanging
        for (NbShard &s : h->sh) {
            int flag = 0;
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaMemcpy(&flag, s.tri_ctl + 2, sizeof(int), cudaMemcpyDeviceToHost));
            if (flag) {
                // what every shard was waiting for when it gave up (tickets handed out; first abandoned poll)
                char msg[900];
                int at = snprintf(msg, sizeof msg, "pair-symmetric wavefront abandoned: a hand-off between rows did not arrive within the spin limit;");
                for (NbShard &o : h->sh) {
                    int c[kTriCtlInts];
                    HGB_CUDA(cudaSetDevice(o.dev->device));
                    HGB_CUDA(cudaMemcpy(c, o.tri_ctl, sizeof c, cudaMemcpyDeviceToHost));
                    HGB_CUDA(cudaMemset(o.tri_ctl + 2, 0, 2 * sizeof(int)));
                    if (at < (int)sizeof msg - 1)
                        at += snprintf(msg + at, sizeof msg - at, " shard %d: %d tickets of %d items%s", o.index, c[0], o.tri_ntiles * o.tri_plan_rows,
                                       c[3] ? "" : ", no poll abandoned here;");
                    if (c[3] && at < (int)sizeof msg - 1)
                        at += snprintf(msg + at, sizeof msg - at, ", wanted tag epoch %u id %d, found tag %#x, chain word %d, row-accumulator word %d;",
                                       (unsigned)c[6] >> 21, c[6] & 0x1fffff, (unsigned)c[24], c[25], c[26]);
                }
                set_error("%s", msg);
                return HGB_ERR_STATE;
            }
        }
    }
    return HGB_OK;
}

int hgb_nbody_step(hgb_nbody *h, int iterations, float *ms) {
    HGB_ARG(h && iterations >= 0, "handle / iterations");
    if (!h->uploaded) { set_error("hgb_nbody_step before hgb_nbody_upload"); return HGB_ERR_STATE; }
    return nb_advance(h, iterations, ms, false);
}

// ---- integrator extension (the reference has positions only) -------------------
int hgb_nbody_set_dynamics(hgb_nbody *h, const double *vel_aos, const double *mass, double dt) {
    HGB_ARG(h, "handle");
    HGB_ARG(dt == dt && dt - dt == 0.0, "dt must be finite");
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const size_t bytes = 3ull * h->npad * sizeof(double);
        if (!s.vel) HGB_CUDA(cudaMalloc(&s.vel, bytes));
        if (!s.mass) HGB_CUDA(cudaMalloc(&s.mass, (size_t)h->npad * sizeof(double)));
        if (vel_aos) {
            HGB_CUDA(cudaMemcpyAsync(s.stage, vel_aos, 3ull * h->n * sizeof(double), cudaMemcpyHostToDevice,
                                     s.dev->stream));
            aos_to_soa_kernel<<<(h->npad + 255) / 256, 256, 0, s.dev->stream>>>(s.stage, s.vel, h->n, h->npad);
            HGB_LAUNCHED();
        } else {
            HGB_CUDA(cudaMemsetAsync(s.vel, 0, bytes, s.dev->stream));
        }
        HGB_CUDA(cudaMemsetAsync(s.mass, 0, (size_t)h->npad * sizeof(double), s.dev->stream));
        if (mass) {
            HGB_CUDA(cudaMemcpyAsync(s.mass, mass, (size_t)h->n * sizeof(double), cudaMemcpyHostToDevice,
                                     s.dev->stream));
        } else {
            fill_value_kernel<<<(h->n + 255) / 256, 256, 0, s.dev->stream>>>(s.mass, h->n, 1.0);
            HGB_LAUNCHED();
        }
    }
    for (NbShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    h->dt = dt;
    h->dynamics = true;
    return HGB_OK;
}

int hgb_nbody_integrate(hgb_nbody *h, int steps, float *ms) {
    HGB_ARG(h && steps >= 0, "handle / steps");
    if (!h->uploaded || !h->dynamics) {
        set_error("hgb_nbody_integrate needs hgb_nbody_upload and hgb_nbody_set_dynamics first");
        return HGB_ERR_STATE;
    }
    return nb_advance(h, steps, ms, true);
}

int hgb_nbody_download_velocities(hgb_nbody *h, double *vel_aos) {
    HGB_ARG(h && vel_aos, "handle / velocities");
    if (!h->dynamics) { set_error("no velocities: hgb_nbody_set_dynamics was not called"); return HGB_ERR_STATE; }
    if (h->rank_mode && h->world > 1) {
        // every rank owns one slice: gather them (in place) so each rank can return all n
        NbShard &s = h->sh[0];
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_NCCL(nccl().GroupStart());
        for (int c = 0; c < 3; ++c) {
            double *recv = s.vel + (size_t)c * h->npad;
            HGB_NCCL(nccl().AllGather(recv + (size_t)s.index * h->slice, recv, h->slice, ncclDouble,
                                      rt().rank_comm, s.dev->stream));
        }
        HGB_NCCL(nccl().GroupEnd());
    }
    // single process: shard i holds the live velocities of slice i
    for (NbShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const int t0 = h->rank_mode ? 0 : s.index * h->slice;
        const int t1 = h->rank_mode ? h->n : min(h->n, t0 + h->slice);
        if (t1 <= t0) continue;
        soa_to_aos_kernel<<<(h->n + 255) / 256, 256, 0, s.dev->stream>>>(s.vel, s.stage, h->n, h->npad);
        HGB_LAUNCHED();
        HGB_CUDA(cudaMemcpyAsync(vel_aos + 3ull * t0, s.stage + 3ull * t0, 3ull * (t1 - t0) * sizeof(double),
                                 cudaMemcpyDeviceToHost, s.dev->stream));
    }
    for (NbShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

int hgb_nbody_download(hgb_nbody *h, double *cur_aos, double *prev_aos) {
    HGB_ARG(h, "handle");
    if (!h->uploaded) { set_error("hgb_nbody_download before upload"); return HGB_ERR_STATE; }
    NbShard &s = h->sh[0];  // after the exchange every shard holds all n bodies
    HGB_CUDA(cudaSetDevice(s.dev->device));
    double *dst[2] = {cur_aos, prev_aos};
    for (int w = 0; w < 2; ++w) {
        if (!dst[w]) continue;
        soa_to_aos_kernel<<<(h->n + 255) / 256, 256, 0, s.dev->stream>>>(
            s.state[w == 0 ? h->cur : h->cur ^ 1], s.stage, h->n, h->npad);
        HGB_LAUNCHED();
        HGB_CUDA(cudaMemcpyAsync(dst[w], s.stage, 3ull * h->n * sizeof(double),
                                 cudaMemcpyDeviceToHost, s.dev->stream));
        HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    }
    return HGB_OK;
}

int hgb_nbody_accel_subset(hgb_nbody *h, const int *targets, int count, double *out_aos) {
    HGB_ARG(h && targets && out_aos && count >= 0, "handle / targets / out");
    if (!h->uploaded) { set_error("accel_subset before upload"); return HGB_ERR_STATE; }
    if (count == 0) return HGB_OK;
    HGB_ARG(3ll * count <= 3ll * h->npad, "count larger than the body set");
    for (int i = 0; i < count; ++i) HGB_ARG(targets[i] >= 0 && targets[i] < h->n, "target index");
    NbShard &s = h->sh[0];
    HGB_CUDA(cudaSetDevice(s.dev->device));
    if (s.tlist_cap < count) {
        if (s.tlist) cudaFree(s.tlist);
        HGB_CUDA(cudaMalloc(&s.tlist, sizeof(int) * count));
        s.tlist_cap = count;
    }
    HGB_CUDA(cudaMemcpyAsync(s.tlist, targets, sizeof(int) * count, cudaMemcpyHostToDevice,
                             s.dev->stream));
    PeerOut out;
    out.count = 0;
    HGB_TRY(launch_step(h, s, s.state[h->cur], 0, count, s.tlist, out, s.stage));
    HGB_CUDA(cudaMemcpyAsync(out_aos, s.stage, 3ull * count * sizeof(double), cudaMemcpyDeviceToHost,
                             s.dev->stream));
    HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// potential (and, with dynamics, kinetic) energy of the current state; sums per shard on the
// device (warp shuffles + fixed-order block partials), shards added in index order on the host
static int nb_energy(hgb_nbody *h, bool dyn, double *potential, double *kinetic) {
    std::vector<double> part(2 * h->sh.size(), 0.0);
    for (size_t i = 0; i < h->sh.size(); ++i) {
        NbShard &s = h->sh[i];
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const int t0 = s.index * h->slice;
        const int count = max(0, min(h->n, t0 + h->slice) - t0);
        const int grid = (count + 127) / 128;
        double *blk = s.partials + 2, *kin = blk + (h->slice / 64 + 1);
        HGB_CUDA(cudaMemsetAsync(s.partials, 0, 2 * sizeof(double), s.dev->stream));
        if (count > 0) {
            if (dyn) {
                const Dynamics d{s.mass, s.vel, h->dt};
                nbody_energy_kernel<128, true><<<grid, 128, 0, s.dev->stream>>>(s.state[h->cur], h->npad, h->n,
                                                                             t0, count, blk, d, kin);
                HGB_LAUNCHED();
                sum_partials_kernel<<<1, 256, 0, s.dev->stream>>>(kin, grid, s.partials + 1);
                HGB_LAUNCHED();
            } else {
                nbody_energy_kernel<128><<<grid, 128, 0, s.dev->stream>>>(s.state[h->cur], h->npad, h->n, t0,
                                                                         count, blk);
                HGB_LAUNCHED();
            }
            sum_partials_kernel<<<1, 256, 0, s.dev->stream>>>(blk, grid, s.partials);
            HGB_LAUNCHED();
        }
        if (h->rank_mode && h->world > 1)
            HGB_NCCL(nccl().AllReduce(s.partials, s.partials, 2, ncclDouble, ncclSum, rt().rank_comm,
                                      s.dev->stream));
        HGB_CUDA(cudaMemcpyAsync(&part[2 * i], s.partials, 2 * sizeof(double), cudaMemcpyDeviceToHost,
                                 s.dev->stream));
    }
    for (NbShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    double pot = 0.0, kin = 0.0;
    for (size_t i = 0; i < h->sh.size(); ++i) { pot += part[2 * i]; kin += part[2 * i + 1]; }
    *potential = -kG * 0.5 * pot;  // every pair was counted twice
    if (kinetic) *kinetic = 0.5 * kin;
    return HGB_OK;
}

int hgb_nbody_energy(hgb_nbody *h, double *energy) {
    HGB_ARG(h && energy, "handle / energy");
    if (!h->uploaded) { set_error("energy before upload"); return HGB_ERR_STATE; }
    return nb_energy(h, false, energy, nullptr);
}

int hgb_nbody_total_energy(hgb_nbody *h, double *kinetic, double *potential) {
    HGB_ARG(h && kinetic && potential, "handle / kinetic / potential");
    if (!h->uploaded || !h->dynamics) {
        set_error("hgb_nbody_total_energy needs hgb_nbody_upload and hgb_nbody_set_dynamics first");
        return HGB_ERR_STATE;
    }
    return nb_energy(h, true, potential, kinetic);
}

// ---- the drop-in ---------------------------------------------------------
static hgb_nbody *g_cached = nullptr;
static int g_cached_n = 0, g_cached_g = 0;

int hgb_nbody_run(double *planets_aos, double *buffer_aos, int n, int iterations, int ngpus) {
    HGB_ARG(n >= 0 && iterations >= 0, "negative size or iteration count");
    int g = 1;
    HGB_TRY(resolve_ngpus(ngpus, &g));            // fails loudly without a GPU, even for n == 0
    if (n == 0 || iterations == 0) return HGB_OK; // the reference's loops do nothing either
    HGB_ARG(planets_aos && buffer_aos, "NULL body array");
    if (!g_cached || g_cached_n != n || g_cached_g != g) {
        hgb_nbody_destroy(g_cached);
        g_cached = nullptr;
        HGB_TRY(hgb_nbody_create(&g_cached, n, g));
        g_cached_n = n;
        g_cached_g = g;
    }
    hgb_nbody *h = g_cached;
    HGB_TRY(hgb_nbody_config(h, g_default_mode, g_default_exchange));
    HGB_TRY(hgb_nbody_upload(h, planets_aos));
    HGB_TRY(hgb_nbody_step(h, iterations, nullptr));
    // reference swaps LOCAL pointers: even states live in planets, odd ones in buffer
    if (iterations % 2 == 0) return hgb_nbody_download(h, planets_aos, buffer_aos);
    return hgb_nbody_download(h, buffer_aos, iterations > 1 ? planets_aos : nullptr);
}

}  // extern "C"


// ---- self test of the branch-free IEEE sequences -------------------------------
// Compares inv_sqrt_ieee(t) with __ddiv_rn(1.0, __dsqrt_rn(t)) bit for bit on
// `count` operands: random doubles spread over the whole fast-path range, operands
// clustered the way the nbody kernel produces them (cubes of small sums), and hard
// cases for both roundings (t next to (m+1/2)^2, roots with all-ones significands).
__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

__global__ void __launch_bounds__(256)
inv_sqrt_selftest_kernel(unsigned long long count, unsigned long long seed,
                         unsigned long long *mismatches, double *first_bad) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    const unsigned long long stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (; i < count; i += stride) {
        const unsigned long long r = mix64(seed + i), r2 = mix64(r);
        double t;
        switch (i & 3) {
            case 0: {  // any exponent in the fast-path range, random significand
                unsigned long long e = 0x036 + (r2 % (0x7fe - 0x036));
                t = __longlong_as_double((long long)((e << 52) | (r >> 12)));
                break;
            }
            case 1: {  // what the kernel sees: t = s^3, s = small sum + EPS
                double s = (double)(r >> 11) * (1.0 / 9007199254740992.0) * ((r2 & 7) + 1) + kEps;
                t = __dmul_rn(__dmul_rn(s, s), s);
                break;
            }
            case 2: {  // sqrt hard cases: t adjacent to (m + 1/2)^2
                double m = __longlong_as_double((long long)((0x3ffull << 52) | (r >> 12)));
                double h = m + 1.1102230246251565e-16;   // m + ulp/2 (may round; still near-tie)
                t = __dmul_rn(h, h);
                long long b = __double_as_longlong(t) + (long long)(r2 % 5) - 2;
                t = __longlong_as_double(b);
                if (r2 & 64) t *= 4503599627370496.0;
                break;
            }
            default: {  // roots with all-ones (and nearly all-ones) significands
                unsigned long long e = 0x200 + (r2 % 0x3ff);
                unsigned long long sig = 0xFFFFFFFFFFFFFull - (r % 3);
                double g = __longlong_as_double((long long)((e << 52) | sig));
                t = __dmul_rn(g, g);
                long long b = __double_as_longlong(t) + (long long)((r >> 20) % 5) - 2;
                t = __longlong_as_double(b);
                break;
            }
        }
        if (range_key(t) >= kRangeLimit) continue;
        const double want = __ddiv_rn(1.0, __dsqrt_rn(t));
        const double got = inv_sqrt_ieee(t);
        if (__double_as_longlong(want) != __double_as_longlong(got)) {
            if (!bad) first_bad[0] = t;
            ++bad;
        }
    }
    if (bad) atomicAdd(mismatches, bad);
}

// pair_wise_accel (reference src/nbody/nbody-run.c:6-15) for `count` sources in order:
// acc += d * factor, no G scaling.  One thread; this is the symbol-level drop-in, not a fast path.
__global__ void nbody_pairwise_kernel(const double *__restrict__ p1, const double *__restrict__ p2,
                                      int count, double *__restrict__ acc) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double ax = acc[0], ay = acc[1], az = acc[2];
    unsigned key = 0;
    for (int j = 0; j < count; ++j)
        interact<false, true>(p1[0], p1[1], p1[2], p2[3 * j], p2[3 * j + 1], p2[3 * j + 2], ax, ay, az, key);
    acc[0] = ax; acc[1] = ay; acc[2] = az;
}

extern "C" int hgb_nbody_pairwise(const double *p1, const double *p2_aos, int count, double *acc_inout) {
    HGB_ARG(p1 && p2_aos && acc_inout && count >= 0, "pairwise arguments");
    HGB_TRY(ensure_rank_dev());
    DevCtx &d = rt().rank_dev;
    HGB_CUDA(cudaSetDevice(d.device));
    double *buf = nullptr;
    const size_t doubles = 6 + 3ull * count;
    HGB_CUDA(cudaMalloc(&buf, doubles * sizeof(double)));
    HGB_CUDA(cudaMemcpyAsync(buf, p1, 24, cudaMemcpyHostToDevice, d.stream));
    HGB_CUDA(cudaMemcpyAsync(buf + 3, acc_inout, 24, cudaMemcpyHostToDevice, d.stream));
    if (count) HGB_CUDA(cudaMemcpyAsync(buf + 6, p2_aos, 24ull * count, cudaMemcpyHostToDevice, d.stream));
    nbody_pairwise_kernel<<<1, 32, 0, d.stream>>>(buf, buf + 6, count, buf + 3);
    HGB_LAUNCHED();
    HGB_CUDA(cudaMemcpyAsync(acc_inout, buf + 3, 24, cudaMemcpyDeviceToHost, d.stream));
    HGB_CUDA(cudaStreamSynchronize(d.stream));
    cudaFree(buf);
    return HGB_OK;
}

extern "C" int hgb_selftest_inv_sqrt(unsigned long long count, unsigned long long seed,
                                     unsigned long long *mismatches, double *first_bad) {
    HGB_ARG(mismatches, "mismatches == NULL");
    HGB_TRY(ensure_rank_dev());
    DevCtx &d = rt().rank_dev;
    HGB_CUDA(cudaSetDevice(d.device));
    unsigned long long *dm = nullptr;
    double *db = nullptr;
    HGB_CUDA(cudaMalloc(&dm, 8));
    HGB_CUDA(cudaMalloc(&db, 8));
    HGB_CUDA(cudaMemsetAsync(dm, 0, 8, d.stream));
    HGB_CUDA(cudaMemsetAsync(db, 0, 8, d.stream));
    inv_sqrt_selftest_kernel<<<148 * 8, 256, 0, d.stream>>>(count, seed, dm, db);
    HGB_LAUNCHED();
    double fb = 0.0;
    HGB_CUDA(cudaMemcpyAsync(mismatches, dm, 8, cudaMemcpyDeviceToHost, d.stream));
    HGB_CUDA(cudaMemcpyAsync(&fb, db, 8, cudaMemcpyDeviceToHost, d.stream));
    HGB_CUDA(cudaStreamSynchronize(d.stream));
    if (first_bad) *first_bad = fb;
    cudaFree(dm);
    cudaFree(db);
    return HGB_OK;
}

namespace hgb {
void nbody_drop_cache() {
    hgb_nbody_destroy(g_cached);
    g_cached = nullptr;
}
}  // namespace hgb
