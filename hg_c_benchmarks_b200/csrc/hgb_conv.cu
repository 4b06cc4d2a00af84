// hgb_conv.cu -- host side of the convolution path: device-resident bands,
// row-band sharding with halo exchange (ncclSend/ncclRecv or peer copies) and the
// drop-in for run_on_padded_image() (reference src/convolution/convolution-run.c:41-49).
#include <algorithm>
#include <vector>

#include "hgb_common.cuh"
#include "conv_kernels.cuh"

using namespace hgb;

namespace {
constexpr int kTW = 256;   // strip width of the streaming kernels

struct CvShard {
    DevCtx *dev = nullptr;
    int index = 0;
    int row0 = 0, rows = 0;
    double *img[2] = {nullptr, nullptr};
    double *taps_dev = nullptr;
    double *scratch = nullptr;
    double *bak = nullptr;       // hgb_conv_backup: copy of one state (copy_padded_image on the device)
};
}  // namespace

struct hgb_conv {
    int W = 0, H = 0, p = 0, world = 1, active = 1;  // active = ranks that own rows
    bool rank_mode = false;
    int dp = 0, halo = 0;     // halo = rows stored above/below a band (addressing); >= depth * p
    int depth_req = 0;        // iterations per halo exchange asked for (0 = automatic)
    int valid_xr = 0;         // halo rows of the current state that hold the neighbours' data
    int kw = 0, kh = 0;
    Taps taps;
    std::vector<double> taps_host;
    bool have_kernel = false;
    int variant = 0, chunk = 0, exchange = HGB_EXCHANGE_NCCL;
    bool overlap = true;   // boundary/interior split with the exchange under the interior kernel
    bool fuse2 = true;     // 3x3: two iterations per pass where the schedule allows
    bool fuse_warp = false; // warp-independent variant of that kernel (shuffles instead of a shared-memory row)
    int fuse_cols = 2;     // columns per thread of that kernel (2 or 4; measured equal, 2 slightly ahead)
    int cur = 0;
    bool uploaded = false;
    bool have_backup = false;
    int bak_valid_xr = 0;
    int mode = HGB_CONV_EXACT;    // HGB_CONV_FAST: fused multiply-add taps, floating-point images only
    bool integer_input = false;   // the uploaded image was integer-valued: exact kernels even in FAST mode
    bool bak_integer_input = false;
    std::vector<CvShard> sh;
};

static BandGeom geom(const hgb_conv *h, const CvShard &s) {
    BandGeom g;
    g.W = h->W; g.H = h->H; g.p = h->p;
    g.row0 = s.row0; g.rows = s.rows;
    g.crow0 = s.row0; g.crows = s.rows;
    g.halo = h->halo; g.dp = h->dp;
    return g;
}

// FAST arithmetic applies to floating-point images only (north_star: integer images stay bit-exact)
static bool use_fast(const hgb_conv *h) { return h->mode == HGB_CONV_FAST && !h->integer_input; }

// band i of `active` equal bands (first H % active bands get one extra row)
static void band_of(int H, int active, int i, int *row0, int *rows) {
    if (i >= active) { *row0 = H; *rows = 0; return; }
    int base = H / active, rem = H % active;
    *row0 = i * base + (i < rem ? i : rem);
    *rows = base + (i < rem ? 1 : 0);
}

constexpr int kBnd = 32;   // boundary rows per side computed ahead of the exchange (and halo capacity)

// Communication-avoiding halo: exchange depth*p rows once every `depth` iterations and let a band
// recompute the shrinking fringe in between (iteration t of a cycle covers (depth-1-t)*p rows beyond
// the band).  Same arithmetic per pixel => bit-identical; exchange latency is amortised depth-fold.
static int eff_depth(const hgb_conv *h) {
    if (h->active <= 1 || h->p <= 0) return 1;
    int minrows = h->H / h->active;
    // automatic: 8; 16 for 3x3, whose non-final iterations of a cycle run two per pass
    int d = h->depth_req > 0 ? h->depth_req : (h->p == 1 && h->fuse2 ? 16 : 8);
    if (d > h->halo / h->p) d = h->halo / h->p;
    if (d > minrows / h->p) d = minrows / h->p;
    return d < 1 ? 1 : d;
}

static int cv_setup(hgb_conv *h) {
    const int p = h->p;
    h->halo = (h->world > 1 && p > 0 && p < kBnd) ? kBnd : p;
    // every band must be at least p rows tall so halos come from direct neighbours only
    int act = h->world;
    const int need = p > 0 ? p : 1;
    if (act > h->H / need) act = h->H / need;
    if (act < 1) act = 1;
    h->active = act;
    int wpad = (h->W + kTW - 1) / kTW * kTW;
    h->dp = kLead + wpad + 16;            // staged strips may over-read up to LP columns
    for (CvShard &s : h->sh) {
        band_of(h->H, act, s.index, &s.row0, &s.rows);
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const size_t bytes = (size_t)(s.rows + 2 * h->halo + 1) * h->dp * sizeof(double);
        for (int b = 0; b < 2; ++b) {
            HGB_CUDA(cudaMalloc(&s.img[b], bytes));
            HGB_CUDA(cudaMemsetAsync(s.img[b], 0, bytes, s.dev->stream));
        }
        HGB_CUDA(cudaMalloc(&s.taps_dev, sizeof(double) * 1024));
        HGB_CUDA(cudaMalloc(&s.scratch, 64));
    }
    return HGB_OK;
}

// rows [crow0, crow0+crows) of the band on stream `st`
template <int K, int NST>
static int launch_stream(const hgb_conv *h, const CvShard &s, const double *in, double *out, int crow0,
                         int crows, cudaStream_t st) {
    using Cfg = ConvCfg<K, kTW, 2, NST>;
    const bool fast = use_fast(h);
    auto kern = fast ? conv_stream_kernel<K, kTW, 2, NST, true> : conv_stream_kernel<K, kTW, 2, NST, false>;
    static bool attr_done[2][16] = {{false}};
    if (!attr_done[fast][s.dev->device & 15]) {
        HGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
        attr_done[fast][s.dev->device & 15] = true;
    }
    // output rows per CTA: small chunks keep the concurrently streamed rows close together in
    // DRAM (3x3 is HBM-bound); larger ones amortise the K-1 extra input rows (5x5, 7x7 are
    // FP64-bound).  Measured on B200: profiles/r1_conv_chunk_sweep.txt
    int chunk = h->chunk > 0 ? h->chunk : (K == 3 ? 8 : K == 5 ? 32 : 128);
    if (chunk > crows) chunk = crows;
    BandGeom g = geom(h, s);
    g.crow0 = crow0;
    g.crows = crows;
    dim3 grid((h->W + kTW - 1) / kTW, (crows + chunk - 1) / chunk);
    kern<<<grid, Cfg::THREADS, Cfg::SMEM, st>>>(in, out, g, chunk, h->taps);
    HGB_LAUNCHED();
    return HGB_OK;
}

// rows [crow0, crow0+crows) advanced by TWO iterations in one pass (3x3 only)
static int launch_fused2(const hgb_conv *h, const CvShard &s, const double *in, double *out, int crow0,
                         int crows, cudaStream_t st) {
    if (crows <= 0) return HGB_OK;
    // rows per CTA, measured on B200: 64 on a 16384-row image (profiles/r1c_conv3_fused_sweep.txt); on the
    // 2048-row band of an 8-way split 64 leaves 1.56 waves of long CTAs (1.20 x an eighth of the full image),
    // 48 gives 1.09 x (profiles/r2q_conv_band_sweep.txt): shorter CTAs, shorter tail
    int chunk = h->chunk > 0 ? h->chunk : (crows <= 4096 ? 48 : 64);
    if (chunk > crows) chunk = crows;
    BandGeom g = geom(h, s);
    g.crow0 = crow0;
    g.crows = crows;
    if (h->fuse_warp && !use_fast(h)) {
        dim3 gridw((h->W + kF2wOut - 1) / kF2wOut, (crows + chunk - 1) / chunk);
        conv3_fused2w_kernel<<<gridw, kF2wWarps * 32, kF2wSmem, st>>>(in, out, g, chunk, h->taps);
        HGB_LAUNCHED();
        return HGB_OK;
    }
    dim3 grid((h->W + kF2Out - 1) / kF2Out, (crows + chunk - 1) / chunk);
    if (use_fast(h)) {
        if (h->fuse_cols == 2) conv3_fused2_kernel<2, true><<<grid, kF2TW / 2, kF2Smem, st>>>(in, out, g, chunk, h->taps);
        else conv3_fused2_kernel<4, true><<<grid, kF2TW / 4, kF2Smem, st>>>(in, out, g, chunk, h->taps);
    } else if (h->fuse_cols == 2) conv3_fused2_kernel<2><<<grid, kF2TW / 2, kF2Smem, st>>>(in, out, g, chunk, h->taps);
    else conv3_fused2_kernel<4><<<grid, kF2TW / 4, kF2Smem, st>>>(in, out, g, chunk, h->taps);
    HGB_LAUNCHED();
    return HGB_OK;
}

static int launch_conv(const hgb_conv *h, const CvShard &s, const double *in, double *out, int crow0,
                       int crows, cudaStream_t st) {
    if (crows <= 0) return HGB_OK;
    const bool square = h->kw == h->kh && h->kw == 2 * h->p + 1;
    if (h->variant != 1 && square) {
        if (h->kw == 3) return launch_stream<3, 4>(h, s, in, out, crow0, crows, st);
        if (h->kw == 5) return launch_stream<5, 3>(h, s, in, out, crow0, crows, st);
        if (h->kw == 7) return launch_stream<7, 3>(h, s, in, out, crow0, crows, st);
    }
    BandGeom g = geom(h, s);
    g.crow0 = crow0;
    g.crows = crows;
    dim3 grid((h->W + 255) / 256, crows);
    if (use_fast(h)) conv_generic_kernel<true><<<grid, 256, 0, st>>>(in, out, g, s.taps_dev, h->kw, h->kh);
    else conv_generic_kernel<false><<<grid, 256, 0, st>>>(in, out, g, s.taps_dev, h->kw, h->kh);
    HGB_LAUNCHED();
    return HGB_OK;
}

// exchange the p edge rows of state `b` between neighbouring bands
static int halo_exchange(hgb_conv *h, int b, bool side = false) {
    // side: run on the boundary stream (stream2) so the interior kernel overlaps the exchange
    auto xs = [side](const CvShard &s) { return side ? s.dev->stream2 : s.dev->stream; };
    if (h->active <= 1 || h->p == 0) return HGB_OK;
    const int xr = eff_depth(h) * h->p;          // rows exchanged per side
    const size_t cnt = (size_t)xr * h->dp;       // full device rows (pad columns included)
    if (!h->rank_mode && h->exchange == HGB_EXCHANGE_P2P) {
        // peer copies on the RECEIVER's stream after the sender's kernel finished
        for (CvShard &s : h->sh) {
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaEventRecord(s.dev->ev_step, xs(s)));
        }
        for (CvShard &s : h->sh) {
            if (s.rows <= 0) continue;
            HGB_CUDA(cudaSetDevice(s.dev->device));
            if (s.index > 0) {
                CvShard &up = h->sh[s.index - 1];
                HGB_CUDA(cudaStreamWaitEvent(xs(s), up.dev->ev_step, 0));
                HGB_CUDA(cudaMemcpyPeerAsync(s.img[b] + (size_t)(h->halo - xr) * h->dp, s.dev->device,
                                             up.img[b] + (size_t)(h->halo + up.rows - xr) * h->dp,
                                             up.dev->device, cnt * sizeof(double), xs(s)));
            }
            if (s.index < h->active - 1) {
                CvShard &dn = h->sh[s.index + 1];
                HGB_CUDA(cudaStreamWaitEvent(xs(s), dn.dev->ev_step, 0));
                HGB_CUDA(cudaMemcpyPeerAsync(s.img[b] + (size_t)(h->halo + s.rows) * h->dp,
                                             s.dev->device, dn.img[b] + (size_t)h->halo * h->dp,
                                             dn.dev->device, cnt * sizeof(double), xs(s)));
            }
        }
        // a band may only be overwritten (next iteration writes the OTHER buffer, but the
        // one after that writes this one) once its neighbours have pulled their halos
        for (CvShard &s : h->sh) {
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaEventRecord(s.dev->ev_step, xs(s)));
        }
        for (CvShard &s : h->sh)
            for (int d = -1; d <= 1; d += 2) {
                int o = s.index + d;
                if (o < 0 || o >= h->active) continue;
                HGB_CUDA(cudaStreamWaitEvent(xs(s), h->sh[o].dev->ev_step, 0));
            }
        h->valid_xr = xr;
        return HGB_OK;
    }
    HGB_NCCL(nccl().GroupStart());
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        ncclComm_t comm = h->rank_mode ? rt().rank_comm : sp_comm(h->world, s.index);
        double *base = s.img[b];
        if (s.index > 0) {
            HGB_NCCL(nccl().Send(base + (size_t)h->halo * h->dp, cnt, ncclDouble, s.index - 1, comm,
                                 xs(s)));
            HGB_NCCL(nccl().Recv(base + (size_t)(h->halo - xr) * h->dp, cnt, ncclDouble, s.index - 1, comm,
                                 xs(s)));
        }
        if (s.index < h->active - 1) {
            HGB_NCCL(nccl().Send(base + (size_t)(h->halo + s.rows - xr) * h->dp, cnt, ncclDouble,
                                 s.index + 1, comm, xs(s)));
            HGB_NCCL(nccl().Recv(base + (size_t)(h->halo + s.rows) * h->dp, cnt, ncclDouble,
                                 s.index + 1, comm, xs(s)));
        }
    }
    HGB_NCCL(nccl().GroupEnd());
    h->valid_xr = xr;
    return HGB_OK;
}

extern "C" {

int hgb_conv_destroy(hgb_conv *h) {
    if (!h) return HGB_OK;
    for (CvShard &s : h->sh) {
        cudaSetDevice(s.dev->device);
        cudaStreamSynchronize(s.dev->stream);
        cudaFree(s.img[0]);
        cudaFree(s.img[1]);
        cudaFree(s.taps_dev);
        cudaFree(s.scratch);
        if (s.bak) cudaFree(s.bak);
    }
    delete h;
    return HGB_OK;
}

static int cv_classify(hgb_conv *h);   // FAST mode: integer-valued images keep the exact kernels

static int cv_check_dims(int w, int hh, int p) {
    HGB_ARG(w > 0 && hh > 0, "image size must be positive");
    HGB_ARG(p >= 0 && p <= kLead - 4, "padding must be in [0, 12]");
    return HGB_OK;
}

int hgb_conv_create(hgb_conv **out, int inner_w, int inner_h, int padding, int ngpus) {
    HGB_ARG(out, "out == NULL");
    HGB_TRY(cv_check_dims(inner_w, inner_h, padding));
    int g = 1;
    HGB_TRY(resolve_ngpus(ngpus, &g));
    HGB_TRY(ensure_devices(g));
    hgb_conv *h = new hgb_conv();
    h->W = inner_w; h->H = inner_h; h->p = padding; h->world = g;
    h->sh.resize(g);
    for (int i = 0; i < g; ++i) { h->sh[i].dev = &rt().devs[i]; h->sh[i].index = i; }
    int rc = cv_setup(h);
    if (rc == HGB_OK && g > 1 && oversubscribed(g)) rc = hgb_conv_set_exchange(h, HGB_EXCHANGE_P2P);
    if (rc != HGB_OK) { hgb_conv_destroy(h); return rc; }
    *out = h;
    return HGB_OK;
}

int hgb_conv_create_rank(hgb_conv **out, int inner_w, int inner_h, int padding) {
    HGB_ARG(out, "out == NULL");
    HGB_TRY(cv_check_dims(inner_w, inner_h, padding));
    Runtime &r = rt();
    HGB_TRY(ensure_rank_dev());
    if (r.world > 1 && !r.rank_comm) {
        set_error("hgb_comm_init_rank must be called before hgb_conv_create_rank");
        return HGB_ERR_STATE;
    }
    hgb_conv *h = new hgb_conv();
    h->W = inner_w; h->H = inner_h; h->p = padding; h->world = r.world; h->rank_mode = true;
    h->sh.resize(1);
    h->sh[0].dev = &r.rank_dev;
    h->sh[0].index = r.rank;
    int rc = cv_setup(h);
    if (rc != HGB_OK) { hgb_conv_destroy(h); return rc; }
    *out = h;
    return HGB_OK;
}

int hgb_conv_set_kernel(hgb_conv *h, const double *kernel, int kw, int kh) {
    HGB_ARG(h && kernel, "handle / kernel");
    HGB_ARG(kw >= 1 && kh >= 1, "kernel size must be positive");
    // the reference reads image[y+ky-p][x+kx-p]; larger kernels index outside its rows
    HGB_ARG(kw <= 2 * h->p + 1 && kh <= 2 * h->p + 1, "kernel larger than 2*padding+1");
    h->kw = kw; h->kh = kh;
    h->taps_host.assign(kernel, kernel + (size_t)kw * kh);
    memset(&h->taps, 0, sizeof(h->taps));
    if ((size_t)kw * kh <= (size_t)kMaxTaps) memcpy(h->taps.k, kernel, sizeof(double) * kw * kh);
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaMemcpyAsync(s.taps_dev, h->taps_host.data(), sizeof(double) * kw * kh,
                                 cudaMemcpyHostToDevice, s.dev->stream));
        HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    }
    h->have_kernel = true;
    return HGB_OK;
}

int hgb_conv_tune(hgb_conv *h, int variant, int band_rows) {
    HGB_ARG(h && variant >= 0 && variant <= 31 && band_rows >= 0, "handle / variant / band_rows");
    h->fuse_warp = (variant & 16) != 0;
    h->fuse_cols = (variant & 8) ? 4 : 2;
    h->variant = variant & 1;
    h->overlap = !(variant & 2);
    h->fuse2 = !(variant & 4);
    h->chunk = band_rows;
    return HGB_OK;
}

int hgb_conv_set_depth(hgb_conv *h, int depth) {
    HGB_ARG(h && depth >= 0, "handle / depth");
    h->depth_req = depth;
    return eff_depth(h);
}

int hgb_conv_set_exchange(hgb_conv *h, int kind) {
    HGB_ARG(h && (kind == HGB_EXCHANGE_NCCL || kind == HGB_EXCHANGE_P2P), "handle / kind");
    if (!h->rank_mode && h->world > 1 && oversubscribed(h->world)) kind = HGB_EXCHANGE_P2P;
    if (kind == HGB_EXCHANGE_P2P && h->world > 1) {
        if (h->rank_mode) { set_error("peer exchange needs the single-process handle"); return HGB_ERR_STATE; }
        HGB_TRY(ensure_peer_access(h->world));
    }
    h->exchange = kind;
    return HGB_OK;
}

int hgb_conv_band(hgb_conv *h, int *row0, int *rows) {
    HGB_ARG(h && row0 && rows, "handle / out pointers");
    if (h->rank_mode) { *row0 = h->sh[0].row0; *rows = h->sh[0].rows; }
    else { *row0 = 0; *rows = h->H; }
    return HGB_OK;
}

// Host rows: pitch W+2p.  Single-process handle: `cur` is the whole image.  Rank
// handle: `cur` is this rank's band only (rows [row0,row0+rows)); halos are filled
// by an exchange.
int hgb_conv_upload(hgb_conv *h, const double *cur) {
    HGB_ARG(h && cur, "handle / image");
    const size_t hp = (size_t)h->W + 2 * h->p;
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        HGB_CUDA(cudaSetDevice(s.dev->device));
        int g0 = s.row0, g1 = s.row0 + s.rows;
        const double *src = cur;
        if (!h->rank_mode) {   // halos straight from the host image
            g0 = max(0, s.row0 - h->halo);
            g1 = min(h->H, s.row0 + s.rows + h->halo);
            src = cur + (size_t)g0 * hp;
        }
        double *dst = s.img[0] + (size_t)(g0 - s.row0 + h->halo) * h->dp + kLead - h->p;
        HGB_CUDA(cudaMemcpy2DAsync(dst, h->dp * sizeof(double), src, hp * sizeof(double),
                                   hp * sizeof(double), g1 - g0, cudaMemcpyHostToDevice,
                                   s.dev->stream));
    }
    h->cur = 0;
    h->uploaded = true;
    h->valid_xr = h->rank_mode ? 0 : h->halo;
    if (h->rank_mode && h->world > 1) HGB_TRY(halo_exchange(h, 0));
    HGB_TRY(cv_classify(h));
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// FAST mode: classify the image just uploaded into state 0 (integer-valued images keep the exact kernels).
// One pass over the band; skipped entirely in EXACT mode.
static int cv_classify(hgb_conv *h) {
    h->integer_input = false;
    if (h->mode != HGB_CONV_FAST) return HGB_OK;
    bool all_int = true;
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        HGB_CUDA(cudaSetDevice(s.dev->device));
        int *flag = reinterpret_cast<int *>(s.scratch) + 8;
        HGB_CUDA(cudaMemsetAsync(flag, 0, sizeof(int), s.dev->stream));
        dim3 grid((h->W + 255) / 256, std::min(s.rows, 64));
        conv_not_integer_kernel<<<grid, 256, 0, s.dev->stream>>>(s.img[0], geom(h, s), s.row0, s.row0 + s.rows, flag);
        HGB_LAUNCHED();
        int host_flag = 0;
        HGB_CUDA(cudaMemcpyAsync(&host_flag, flag, sizeof(int), cudaMemcpyDeviceToHost, s.dev->stream));
        HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
        all_int = all_int && host_flag == 0;
    }
    if (h->rank_mode && h->world > 1) {
        // every rank must take the same kernels: the image is integer-valued only if every band is
        double v = all_int ? 0.0 : 1.0, *d = reinterpret_cast<double *>(h->sh[0].scratch);
        cudaStream_t st = h->sh[0].dev->stream;
        HGB_CUDA(cudaMemcpyAsync(d, &v, sizeof v, cudaMemcpyHostToDevice, st));
        HGB_NCCL(nccl().AllReduce(d, d, 1, ncclDouble, ncclSum, rt().rank_comm, st));
        HGB_CUDA(cudaMemcpyAsync(&v, d, sizeof v, cudaMemcpyDeviceToHost, st));
        HGB_CUDA(cudaStreamSynchronize(st));
        all_int = v == 0.0;
    }
    h->integer_input = all_int;
    return HGB_OK;
}

int hgb_conv_set_mode(hgb_conv *h, int mode) {
    HGB_ARG(h && (mode == HGB_CONV_EXACT || mode == HGB_CONV_FAST), "handle / mode");
    if (mode == HGB_CONV_FAST && h->mode != HGB_CONV_FAST && h->uploaded) {
        set_error("hgb_conv_set_mode(FAST) must precede the upload (the image is classified there)");
        return HGB_ERR_STATE;
    }
    h->mode = mode;
    return HGB_OK;
}

int hgb_conv_mode_in_effect(const hgb_conv *h) { return h && use_fast(h) ? HGB_CONV_FAST : HGB_CONV_EXACT; }

int hgb_conv_step(hgb_conv *h, int iterations, float *ms) {
    HGB_ARG(h && iterations >= 0, "handle / iterations");
    if (!h->uploaded || !h->have_kernel) {
        set_error("hgb_conv_step needs hgb_conv_upload and hgb_conv_set_kernel first");
        return HGB_ERR_STATE;
    }
    if (h->active > 1 && !h->rank_mode && h->exchange == HGB_EXCHANGE_NCCL)
        HGB_TRY(ensure_sp_comm(h->world));
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaEventRecord(s.dev->ev_begin, s.dev->stream));
    }
    // Multi-GPU schedule, per cycle of d <= depth iterations (see eff_depth):
    //   iterations 0..d-2: one launch over the band plus its shrinking fringe (main stream);
    //   iteration d-1:     the band's first/last kBnd rows on stream2 followed by the halo exchange,
    //                      the interior rows on the main stream -- the exchange runs under the interior.
    // ev_pre (main -> stream2): everything before the last iteration is done;
    // ev_x   (stream2 -> main): boundary rows and halos of the finished cycle are in place.
    bool overlap = h->overlap && h->active > 1 && h->p > 0;
    for (CvShard &s : h->sh)
        if (s.rows > 0 && s.rows < 4 * kBnd) overlap = false;
    const int depth = eff_depth(h);
    if (h->active > 1 && h->valid_xr < depth * h->p && iterations > 0)
        HGB_TRY(halo_exchange(h, h->cur));   // the depth was raised after the last exchange
    if (overlap)
        for (CvShard &s : h->sh) {   // stream2 starts after whatever the main stream did before
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaStreamWaitEvent(s.dev->stream2, s.dev->ev_begin, 0));
        }
    // One band, 3x3: two iterations per pass (conv3_fused2_kernel) while at least one plain iteration
    // remains at the end, so that the state before the last iteration -- what the caller gets as
    // `prev` -- exists in memory.
    int fused = 0;
    const bool fuse_ok = h->kw == 3 && h->kh == 3 && h->p == 1 && h->variant != 1 && h->fuse2;
    if (h->active == 1 && fuse_ok) {
        const int tail = (iterations & 1) ? 1 : 2;
        for (; iterations - fused > tail; fused += 2) {
            for (CvShard &s : h->sh) {
                if (s.rows <= 0) continue;
                HGB_CUDA(cudaSetDevice(s.dev->device));
                HGB_TRY(launch_fused2(h, s, s.img[h->cur], s.img[h->cur ^ 1], s.row0, s.rows, s.dev->stream));
            }
            h->cur ^= 1;
        }
    }
    for (int it = fused, cycle = 0; it < iterations; ++cycle) {
        const int d = min(depth, iterations - it);
        if (overlap && cycle > 0)
            for (CvShard &s : h->sh) {
                if (s.rows <= 0) continue;
                HGB_CUDA(cudaSetDevice(s.dev->device));
                HGB_CUDA(cudaStreamWaitEvent(s.dev->stream, s.dev->ev_bnd[0], 0));   // ev_x
            }
        for (int t = 0; t < d; ++t, ++it) {
            const int cur = h->cur, nxt = cur ^ 1;
            // 3x3: iterations t and t+1 as ONE pass (conv3_fused2_kernel) unless t+1 is the final
            // iteration of the whole step (the state before it must exist in memory: it is `prev`)
            const bool pair = fuse_ok && t + 1 < d && it + 2 < iterations;
            const int steps = pair ? 2 : 1;
            const bool last = t + steps == d;            // this launch ends the cycle: exchange follows
            const int ext = (d - t - steps) * h->p;      // fringe rows beyond the band (0 when last)
            auto run = [&](const CvShard &s, int row0, int nrows, cudaStream_t st) {
                return pair ? launch_fused2(h, s, s.img[cur], s.img[nxt], row0, nrows, st)
                            : launch_conv(h, s, s.img[cur], s.img[nxt], row0, nrows, st);
            };
            if (!last || !overlap) {
                for (CvShard &s : h->sh) {
                    if (s.rows <= 0) continue;
                    HGB_CUDA(cudaSetDevice(s.dev->device));
                    const int lo = max(0, s.row0 - ext), hi = min(h->H, s.row0 + s.rows + ext);
                    HGB_TRY(run(s, lo, hi - lo, s.dev->stream));
                }
                if (last) HGB_TRY(halo_exchange(h, nxt));
            } else {
                for (CvShard &s : h->sh) {
                    if (s.rows <= 0) continue;
                    HGB_CUDA(cudaSetDevice(s.dev->device));
                    cudaStream_t sb = s.dev->stream2;
                    HGB_CUDA(cudaEventRecord(s.dev->ev_int[0], s.dev->stream));      // ev_pre
                    HGB_CUDA(cudaStreamWaitEvent(sb, s.dev->ev_int[0], 0));
                    HGB_TRY(run(s, s.row0, kBnd, sb));
                    HGB_TRY(run(s, s.row0 + s.rows - kBnd, kBnd, sb));
                }
                HGB_TRY(halo_exchange(h, nxt, true));
                for (CvShard &s : h->sh) {
                    if (s.rows <= 0) continue;
                    HGB_CUDA(cudaSetDevice(s.dev->device));
                    HGB_CUDA(cudaEventRecord(s.dev->ev_bnd[0], s.dev->stream2));     // ev_x
                    HGB_TRY(run(s, s.row0 + kBnd, s.rows - 2 * kBnd, s.dev->stream));
                }
            }
            h->cur = nxt;
            t += steps - 1;
            it += steps - 1;                             // the loop header adds the other one
        }
    }
    if (overlap)
        for (CvShard &s : h->sh) {   // join: the main stream (and its timing event) covers stream2
            HGB_CUDA(cudaSetDevice(s.dev->device));
            HGB_CUDA(cudaEventRecord(s.dev->ev_join, s.dev->stream2));
            HGB_CUDA(cudaStreamWaitEvent(s.dev->stream, s.dev->ev_join, 0));
        }
    float worst = 0.f;
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        HGB_CUDA(cudaEventRecord(s.dev->ev_end, s.dev->stream));
    }
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaEventSynchronize(s.dev->ev_end));
        float t = 0.f;
        HGB_CUDA(cudaEventElapsedTime(&t, s.dev->ev_begin, s.dev->ev_end));
        worst = t > worst ? t : worst;
    }
    HGB_CUDA(cudaGetLastError());
    if (ms) *ms = worst;
    return HGB_OK;
}

int hgb_conv_download(hgb_conv *h, double *cur, double *prev) {
    HGB_ARG(h, "handle");
    if (!h->uploaded) { set_error("hgb_conv_download before upload"); return HGB_ERR_STATE; }
    const size_t hp = (size_t)h->W + 2 * h->p;
    double *dsts[2] = {cur, prev};
    for (int w = 0; w < 2; ++w) {
        if (!dsts[w]) continue;
        const int b = w == 0 ? h->cur : h->cur ^ 1;
        for (CvShard &s : h->sh) {
            if (s.rows <= 0) continue;
            HGB_CUDA(cudaSetDevice(s.dev->device));
            double *dst = dsts[w] + (h->rank_mode ? 0 : (size_t)s.row0 * hp);
            const double *src = s.img[b] + (size_t)h->halo * h->dp + kLead - h->p;
            HGB_CUDA(cudaMemcpy2DAsync(dst, hp * sizeof(double), src, h->dp * sizeof(double),
                                       hp * sizeof(double), s.rows, cudaMemcpyDeviceToHost,
                                       s.dev->stream));
        }
    }
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// ---- device-resident image lifecycle ---------------------------------------
// add_padding (reference src/convolution/convolution-util.c:258-269) on the device: `dense` is the
// unpadded image, inner_h rows of inner_w doubles (rank mode: this rank's band of it); the pad
// columns are produced here instead of being marshalled from the host.
int hgb_conv_upload_dense(hgb_conv *h, const double *dense) {
    HGB_ARG(h && dense, "handle / image");
    const size_t w = (size_t)h->W;
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        HGB_CUDA(cudaSetDevice(s.dev->device));
        int g0 = s.row0, g1 = s.row0 + s.rows;
        const double *src = dense;
        if (!h->rank_mode) {
            g0 = max(0, s.row0 - h->halo);
            g1 = min(h->H, s.row0 + s.rows + h->halo);
            src = dense + (size_t)g0 * w;
        }
        double *dst = s.img[0] + (size_t)(g0 - s.row0 + h->halo) * h->dp + kLead;
        HGB_CUDA(cudaMemcpy2DAsync(dst, h->dp * sizeof(double), src, w * sizeof(double), w * sizeof(double),
                                   g1 - g0, cudaMemcpyHostToDevice, s.dev->stream));
        if (h->p > 0) {
            conv_pad_columns_kernel<<<(g1 - g0 + 15) / 16, 128, 0, s.dev->stream>>>(s.img[0], geom(h, s), g0, g1);
            HGB_LAUNCHED();
        }
    }
    h->cur = 0;
    h->uploaded = true;
    h->valid_xr = h->rank_mode ? 0 : h->halo;
    if (h->rank_mode && h->world > 1) HGB_TRY(halo_exchange(h, 0));
    HGB_TRY(cv_classify(h));
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// remove_padding (reference src/convolution/convolution-util.c:271-279) of the newest state
int hgb_conv_download_dense(hgb_conv *h, double *dense) {
    HGB_ARG(h && dense, "handle / image");
    if (!h->uploaded) { set_error("hgb_conv_download_dense before upload"); return HGB_ERR_STATE; }
    const size_t w = (size_t)h->W;
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        HGB_CUDA(cudaSetDevice(s.dev->device));
        double *dst = dense + (h->rank_mode ? 0 : (size_t)s.row0 * w);
        const double *src = s.img[h->cur] + (size_t)h->halo * h->dp + kLead;
        HGB_CUDA(cudaMemcpy2DAsync(dst, w * sizeof(double), src, h->dp * sizeof(double), w * sizeof(double),
                                   s.rows, cudaMemcpyDeviceToHost, s.dev->stream));
    }
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    return HGB_OK;
}

// copy_padded_image(img, backup) / copy_padded_image(backup, img) (reference
// src/convolution/convolution-util.c:357-373, the REPETITION loop of src/2d-convolution.c:63-68)
// as device-to-device copies: the newest state is saved, restore makes it the newest state again.
int hgb_conv_backup(hgb_conv *h) {
    HGB_ARG(h, "handle");
    if (!h->uploaded) { set_error("hgb_conv_backup before upload"); return HGB_ERR_STATE; }
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const size_t bytes = (size_t)(s.rows + 2 * h->halo + 1) * h->dp * sizeof(double);
        if (!s.bak) HGB_CUDA(cudaMalloc(&s.bak, bytes));
        HGB_CUDA(cudaMemcpyAsync(s.bak, s.img[h->cur], bytes, cudaMemcpyDeviceToDevice, s.dev->stream));
    }
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    h->have_backup = true;
    h->bak_valid_xr = h->valid_xr;
    h->bak_integer_input = h->integer_input;
    return HGB_OK;
}

int hgb_conv_restore(hgb_conv *h) {
    HGB_ARG(h, "handle");
    if (!h->have_backup) { set_error("hgb_conv_restore without hgb_conv_backup"); return HGB_ERR_STATE; }
    for (CvShard &s : h->sh) {
        HGB_CUDA(cudaSetDevice(s.dev->device));
        const size_t bytes = (size_t)(s.rows + 2 * h->halo + 1) * h->dp * sizeof(double);
        HGB_CUDA(cudaMemcpyAsync(s.img[h->cur], s.bak, bytes, cudaMemcpyDeviceToDevice, s.dev->stream));
    }
    for (CvShard &s : h->sh) HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    h->valid_xr = h->bak_valid_xr;
    h->integer_input = h->bak_integer_input;
    return HGB_OK;
}

int hgb_conv_checksum(hgb_conv *h, double *checksum) {
    HGB_ARG(h && checksum, "handle / checksum");
    if (!h->uploaded) { set_error("checksum before upload"); return HGB_ERR_STATE; }
    if (h->rank_mode && h->world > 1) {
        set_error("sequential checksum is only available on the single-process handle");
        return HGB_ERR_STATE;
    }
    double carry = 0.0;   // the sum is sequential: bands are chained in row order
    for (CvShard &s : h->sh) {
        if (s.rows <= 0) continue;
        HGB_CUDA(cudaSetDevice(s.dev->device));
        conv_checksum_seq_kernel<<<1, 32, 0, s.dev->stream>>>(s.img[h->cur], geom(h, s), carry, s.scratch);
        HGB_LAUNCHED();
        HGB_CUDA(cudaMemcpyAsync(&carry, s.scratch, sizeof(double), cudaMemcpyDeviceToHost, s.dev->stream));
        HGB_CUDA(cudaStreamSynchronize(s.dev->stream));
    }
    *checksum = carry;
    return HGB_OK;
}

// ---- the drop-in ---------------------------------------------------------
static hgb_conv *g_cv = nullptr;
static int g_cv_key[4] = {0, 0, 0, 0};
static int g_cv_exchange = HGB_EXCHANGE_NCCL;
static int g_cv_mode = -1;   // -1: read HGB_CONV_FAST from the environment on first use

int hgb_set_conv_mode(int mode) {
    HGB_ARG(mode == HGB_CONV_EXACT || mode == HGB_CONV_FAST, "mode");
    g_cv_mode = mode;
    return HGB_OK;
}

// ---- streamed drop-in: the PCIe copies of row chunks run under the iterations of other chunks ----------
// One GPU, EXACT mode, images much taller than a chunk.  The image is cut into chunks of `Hs` rows that
// overlap by 2F, F = iterations * p: after `iterations` passes over a chunk taken as an image of its own
// (its artificial top/bottom edges clamp like real ones, which is wrong), exactly the rows closer than F to
// an artificial edge are wrong, so the chunks' valid rows tile the image and every valid pixel went through
// the same arithmetic as in the one-piece run: bit-identical.  Three streams: H2D of chunk c+1 (only rows
// not yet on the device; the overlap is copied device-to-device from the previous staging buffer), the
// iterations of chunk c on the sub-handle's stream, D2H of chunk c-1.  In-place (cur is input and output)
// is safe: a chunk's result rows end F rows above the first row the next H2D reads.
namespace {
struct CvPipe {
    hgb_conv *sub = nullptr;
    int W = 0, Hs = 0, p = 0;
    size_t cap = 0;                                  // doubles per staging buffer
    double *din[2] = {nullptr, nullptr}, *dout[2] = {nullptr, nullptr}, *dprev[2] = {nullptr, nullptr};
    cudaStream_t s_in = nullptr, s_out = nullptr;
    cudaEvent_t ev_in[2] = {}, ev_loaded[2] = {}, ev_res[2] = {}, ev_out[2] = {};
    bool events = false;
    void *reg[2] = {nullptr, nullptr};               // caller buffers page-locked by cudaHostRegister
    size_t reg_bytes[2] = {0, 0};
};
CvPipe g_pipe;

void pipe_unpin() {
    for (int i = 0; i < 2; ++i) {
        if (g_pipe.reg[i]) cudaHostUnregister(g_pipe.reg[i]);
        g_pipe.reg[i] = nullptr; g_pipe.reg_bytes[i] = 0;
    }
    (void)cudaGetLastError();
}

void pipe_free() {
    pipe_unpin();
    hgb_conv_destroy(g_pipe.sub);
    g_pipe.sub = nullptr;
    for (int b = 0; b < 2; ++b) {
        cudaFree(g_pipe.din[b]); cudaFree(g_pipe.dout[b]); cudaFree(g_pipe.dprev[b]);
        g_pipe.din[b] = g_pipe.dout[b] = g_pipe.dprev[b] = nullptr;
    }
    if (g_pipe.events) {
        for (int b = 0; b < 2; ++b) {
            cudaEventDestroy(g_pipe.ev_in[b]); cudaEventDestroy(g_pipe.ev_loaded[b]);
            cudaEventDestroy(g_pipe.ev_res[b]); cudaEventDestroy(g_pipe.ev_out[b]);
        }
        cudaStreamDestroy(g_pipe.s_in); cudaStreamDestroy(g_pipe.s_out);
        g_pipe.events = false;
    }
    g_pipe.W = g_pipe.Hs = g_pipe.p = 0; g_pipe.cap = 0;
}

// async copies overlap only from page-locked memory: is [ptr, ptr+bytes) page-locked already (cudaHostAlloc,
// torch pinned memory, the caller's own cudaHostRegister)?  Then nothing to do.  Otherwise register it for THIS
// call only (slot 0: cur, 1: prev): conv_run_streamed drops the registration before it returns, on every path,
// because the caller is free to free() the buffer the moment hgb_conv_run is back (the reference's
// run_on_padded_image callers do) and a registration that outlives its pages poisons the next buffer malloc
// places at the same address.
int pipe_pin(int slot, void *ptr, size_t bytes) {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, ptr) == cudaSuccess && at.type == cudaMemoryTypeHost) {
        cudaPointerAttributes end;
        if (cudaPointerGetAttributes(&end, (char *)ptr + bytes - 1) == cudaSuccess && end.type == cudaMemoryTypeHost) return HGB_OK;
    }
    (void)cudaGetLastError();
    if (g_pipe.reg[slot]) { cudaHostUnregister(g_pipe.reg[slot]); g_pipe.reg[slot] = nullptr; g_pipe.reg_bytes[slot] = 0; }
    (void)cudaGetLastError();
    if (cudaHostRegister(ptr, bytes, cudaHostRegisterDefault) != cudaSuccess) {
        (void)cudaGetLastError();
        return HGB_ERR_STATE;   // the caller falls back to the one-piece path
    }
    g_pipe.reg[slot] = ptr; g_pipe.reg_bytes[slot] = bytes;
    return HGB_OK;
}

int pipe_setup(int W, int Hs, int p) {
    if (g_pipe.sub && g_pipe.W == W && g_pipe.Hs == Hs && g_pipe.p == p) return HGB_OK;
    void *keep[2] = {g_pipe.reg[0], g_pipe.reg[1]};
    size_t keep_b[2] = {g_pipe.reg_bytes[0], g_pipe.reg_bytes[1]};
    g_pipe.reg[0] = g_pipe.reg[1] = nullptr;          // registrations survive a change of geometry
    pipe_free();
    g_pipe.reg[0] = keep[0]; g_pipe.reg[1] = keep[1]; g_pipe.reg_bytes[0] = keep_b[0]; g_pipe.reg_bytes[1] = keep_b[1];
    HGB_TRY(hgb_conv_create(&g_pipe.sub, W, Hs, p, 1));
    g_pipe.W = W; g_pipe.Hs = Hs; g_pipe.p = p;
    g_pipe.cap = (size_t)Hs * (W + 2 * p);
    CvShard &s = g_pipe.sub->sh[0];
    HGB_CUDA(cudaSetDevice(s.dev->device));
    for (int b = 0; b < 2; ++b) {
        HGB_CUDA(cudaMalloc(&g_pipe.din[b], g_pipe.cap * sizeof(double)));
        HGB_CUDA(cudaMalloc(&g_pipe.dout[b], g_pipe.cap * sizeof(double)));
    }
    HGB_CUDA(cudaStreamCreateWithFlags(&g_pipe.s_in, cudaStreamNonBlocking));
    HGB_CUDA(cudaStreamCreateWithFlags(&g_pipe.s_out, cudaStreamNonBlocking));
    for (int b = 0; b < 2; ++b) {
        HGB_CUDA(cudaEventCreateWithFlags(&g_pipe.ev_in[b], cudaEventDisableTiming));
        HGB_CUDA(cudaEventCreateWithFlags(&g_pipe.ev_loaded[b], cudaEventDisableTiming));
        HGB_CUDA(cudaEventCreateWithFlags(&g_pipe.ev_res[b], cudaEventDisableTiming));
        HGB_CUDA(cudaEventCreateWithFlags(&g_pipe.ev_out[b], cudaEventDisableTiming));
    }
    g_pipe.events = true;
    return HGB_OK;
}

void copy_rows(double *dst, size_t dpitch, const double *src, size_t spitch, int cols, int rows, cudaStream_t st) {
    if (rows <= 0) return;
    dim3 grid((unsigned)std::min((cols + 255) / 256, 64), (unsigned)std::min(rows, 1024));
    conv_copy_rows_kernel<<<grid, 256, 0, st>>>(dst, dpitch, src, spitch, cols, rows);
}

static int g_stream_rows = -1;   // -1: HGB_CONV_STREAM_ROWS from the environment (0 = off); else rows per chunk
}  // namespace

int hgb_set_conv_streaming(int chunk_rows) {
    HGB_ARG(chunk_rows >= 0, "chunk rows (0 = never stream, 1 = automatic)");
    g_stream_rows = chunk_rows;
    return HGB_OK;
}

int hgb_conv_unpin(void) {
    pipe_unpin();
    return HGB_OK;
}

// rows per chunk for this call, or 0: run in one piece
static int stream_plan(int inner_w, int inner_h, int padding, int iterations, int g, const double *prev, int *Hs_out) {
    (void)prev;
    if (g_stream_rows < 0) {
        const char *e = getenv("HGB_CONV_STREAM_ROWS");
        g_stream_rows = e ? atoi(e) : 1;
        if (g_stream_rows < 0) g_stream_rows = 0;
    }
    if (g_stream_rows == 0 || g != 1 || g_cv_mode == HGB_CONV_FAST) return 0;   // FAST classifies the WHOLE image first
    const long F = (long)iterations * padding;
    long valid = g_stream_rows;
    if (valid == 1) {   // automatic: ~32 MiB of valid rows per chunk, never less than 8 F (fringe work <= 25 %)
        valid = (32l << 20) / (8l * (inner_w + 2 * padding));
        valid = std::max(valid, std::max(64l, 8 * F));
    }
    const long Hs = valid + 2 * F;
    if (Hs > (1 << 20) || inner_h < 3 * Hs) return 0;       // fewer than ~3 chunks: nothing to overlap
    *Hs_out = (int)Hs;
    return 1;
}

static int conv_run_streamed_pinned(double *cur, double *prev, int W, int H, int p, const double *kernel, int kw,
                                    int kh, int iterations, int Hs);

// page-lock what is not page-locked yet, run, and give the registrations back before returning
static int conv_run_streamed(double *cur, double *prev, int W, int H, int p, const double *kernel, int kw, int kh,
                             int iterations, int Hs) {
    const size_t bytes = (size_t)H * ((size_t)W + 2 * p) * sizeof(double);
    int rc = pipe_pin(0, cur, bytes);
    if (rc == HGB_OK && prev) rc = pipe_pin(1, prev, bytes);
    if (rc == HGB_OK) {
        rc = conv_run_streamed_pinned(cur, prev, W, H, p, kernel, kw, kh, iterations, Hs);
        if (rc != HGB_OK) {   // copies from / to the caller's pages may still be queued: drain before unpinning
            if (g_pipe.events) { cudaStreamSynchronize(g_pipe.s_in); cudaStreamSynchronize(g_pipe.s_out); }
            if (g_pipe.sub) cudaStreamSynchronize(g_pipe.sub->sh[0].dev->stream);
            (void)cudaGetLastError();
        }
    }
    pipe_unpin();
    return rc;
}

static int conv_run_streamed_pinned(double *cur, double *prev, int W, int H, int p, const double *kernel, int kw,
                                    int kh, int iterations, int Hs) {
    const size_t hp = (size_t)W + 2 * p;
    const int F = iterations * p;
    HGB_TRY(pipe_setup(W, Hs, p));
    CvPipe &P = g_pipe;
    hgb_conv *sub = P.sub;
    CvShard &s = sub->sh[0];
    cudaStream_t sc = s.dev->stream;
    HGB_CUDA(cudaSetDevice(s.dev->device));
    if (prev)
        for (int b = 0; b < 2; ++b)
            if (!P.dprev[b]) HGB_CUDA(cudaMalloc(&P.dprev[b], P.cap * sizeof(double)));
    sub->uploaded = false;
    HGB_TRY(hgb_conv_set_mode(sub, HGB_CONV_EXACT));
    HGB_TRY(hgb_conv_set_kernel(sub, kernel, kw, kh));
    // chunk c: sub-image rows [s0, s0 + Hs), valid result rows [v0, v1)
    struct Chunk { int s0, v0, v1; };
    std::vector<Chunk> chunks;
    for (int done = 0; done < H;) {
        Chunk c;
        c.s0 = std::min(done == 0 ? 0 : done - F, H - Hs);
        c.v0 = done;
        c.v1 = (c.s0 + Hs >= H) ? H : c.s0 + Hs - F;
        chunks.push_back(c);
        done = c.v1;
    }
    const int nc = (int)chunks.size();
    auto enqueue_h2d = [&](int c) -> int {   // rows of chunk c into din[c & 1]
        const int b = c & 1;
        const Chunk &k = chunks[c];
        if (c >= 2) HGB_CUDA(cudaStreamWaitEvent(P.s_in, P.ev_loaded[b], 0));   // chunk c-2 has left this buffer
        int fresh0 = k.s0;
        if (c > 0) {   // rows already on the device: the tail of the previous chunk's staging buffer
            const Chunk &q = chunks[c - 1];
            const int have1 = q.s0 + Hs;
            if (have1 > k.s0) {
                copy_rows(P.din[b], hp, P.din[b ^ 1] + (size_t)(k.s0 - q.s0) * hp, hp, (int)hp, have1 - k.s0, P.s_in);
                HGB_LAUNCHED();
                fresh0 = have1;
            }
        }
        if (fresh0 < k.s0 + Hs)
            HGB_CUDA(cudaMemcpyAsync(P.din[b] + (size_t)(fresh0 - k.s0) * hp, cur + (size_t)fresh0 * hp,
                                     (size_t)(k.s0 + Hs - fresh0) * hp * sizeof(double), cudaMemcpyHostToDevice, P.s_in));
        HGB_CUDA(cudaEventRecord(P.ev_in[b], P.s_in));
        return HGB_OK;
    };
    HGB_TRY(enqueue_h2d(0));
    for (int c = 0; c < nc; ++c) {
        const int b = c & 1;
        const Chunk &k = chunks[c];
        if (c + 1 < nc) HGB_TRY(enqueue_h2d(c + 1));
        // the chunk becomes state 0 of the sub-handle (device-to-device, host pitch -> device pitch)
        HGB_CUDA(cudaStreamWaitEvent(sc, P.ev_in[b], 0));
        copy_rows(s.img[0] + (size_t)sub->halo * sub->dp + kLead - p, sub->dp, P.din[b], hp, (int)hp, Hs, sc);
        HGB_LAUNCHED();
        HGB_CUDA(cudaEventRecord(P.ev_loaded[b], sc));
        sub->cur = 0; sub->uploaded = true; sub->valid_xr = sub->halo; sub->integer_input = false;
        HGB_TRY(hgb_conv_step(sub, iterations, nullptr));
        // valid rows -> staging (device pitch -> host pitch), then home on the D2H stream
        if (c >= 2) HGB_CUDA(cudaStreamWaitEvent(sc, P.ev_out[b], 0));
        const int nrow = k.v1 - k.v0;
        const size_t first = (size_t)(k.v0 - k.s0 + sub->halo) * sub->dp + kLead - p;
        copy_rows(P.dout[b], hp, s.img[sub->cur] + first, sub->dp, (int)hp, nrow, sc);
        HGB_LAUNCHED();
        if (prev) {
            copy_rows(P.dprev[b], hp, s.img[sub->cur ^ 1] + first, sub->dp, (int)hp, nrow, sc);
            HGB_LAUNCHED();
        }
        HGB_CUDA(cudaEventRecord(P.ev_res[b], sc));
        HGB_CUDA(cudaStreamWaitEvent(P.s_out, P.ev_res[b], 0));
        HGB_CUDA(cudaMemcpyAsync(cur + (size_t)k.v0 * hp, P.dout[b], (size_t)nrow * hp * sizeof(double), cudaMemcpyDeviceToHost, P.s_out));
        if (prev)
            HGB_CUDA(cudaMemcpyAsync(prev + (size_t)k.v0 * hp, P.dprev[b], (size_t)nrow * hp * sizeof(double), cudaMemcpyDeviceToHost, P.s_out));
        HGB_CUDA(cudaEventRecord(P.ev_out[b], P.s_out));
    }
    HGB_CUDA(cudaStreamSynchronize(P.s_out));
    HGB_CUDA(cudaStreamSynchronize(P.s_in));
    HGB_CUDA(cudaGetLastError());
    return HGB_OK;
}

int hgb_conv_run(double *cur, double *prev, int inner_w, int inner_h, int padding,
                 const double *kernel, int kw, int kh, int iterations, int ngpus) {
    HGB_ARG(iterations >= 0, "negative iteration count");
    int g = 1;
    HGB_TRY(resolve_ngpus(ngpus, &g));          // fails loudly without a GPU
    HGB_TRY(cv_check_dims(inner_w, inner_h, padding));
    HGB_ARG(kernel && kw >= 1 && kh >= 1, "kernel");
    HGB_ARG(kw <= 2 * padding + 1 && kh <= 2 * padding + 1, "kernel larger than 2*padding+1");
    if (iterations == 0) return HGB_OK;         // the reference's loop body never runs
    HGB_ARG(cur, "NULL image");                 // prev == NULL: the caller does not want state n-1 back
    if (g_cv_mode < 0) {
        const char *e = getenv("HGB_CONV_FAST");
        g_cv_mode = (e && atoi(e) != 0) ? HGB_CONV_FAST : HGB_CONV_EXACT;
    }
    int Hs = 0;
    if (stream_plan(inner_w, inner_h, padding, iterations, g, prev, &Hs)) {
        hgb_conv_destroy(g_cv);   // the two paths do not keep two images' worth of device memory around
        g_cv = nullptr;
        const int rc = conv_run_streamed(cur, prev, inner_w, inner_h, padding, kernel, kw, kh, iterations, Hs);
        if (rc != HGB_ERR_STATE) return rc;     // HGB_ERR_STATE: the buffers could not be page-locked -> one piece
    }
    if (!g_cv || g_cv_key[0] != inner_w || g_cv_key[1] != inner_h || g_cv_key[2] != padding ||
        g_cv_key[3] != g) {
        hgb_conv_destroy(g_cv);
        g_cv = nullptr;
        HGB_TRY(hgb_conv_create(&g_cv, inner_w, inner_h, padding, g));
        g_cv_key[0] = inner_w; g_cv_key[1] = inner_h; g_cv_key[2] = padding; g_cv_key[3] = g;
    }
    g_cv->uploaded = false;   // a new image follows: the mode may change
    HGB_TRY(hgb_conv_set_mode(g_cv, g_cv_mode));
    HGB_TRY(hgb_conv_set_exchange(g_cv, g_cv_exchange));
    HGB_TRY(hgb_conv_set_kernel(g_cv, kernel, kw, kh));
    HGB_TRY(hgb_conv_upload(g_cv, cur));
    HGB_TRY(hgb_conv_step(g_cv, iterations, nullptr));
    return hgb_conv_download(g_cv, cur, prev);
}

}  // extern "C"

namespace hgb {
void conv_drop_cache() {
    hgb_conv_destroy(g_cv);
    g_cv = nullptr;
    pipe_free();
}
}  // namespace hgb
