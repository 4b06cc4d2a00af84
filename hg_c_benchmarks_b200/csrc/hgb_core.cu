// hgb_core.cu -- runtime of libhgb200.so: devices, streams, NCCL loading,
// communicators, L2 flush and the two roofline probes (FP64 pipe, HBM copy).
#include <dlfcn.h>

#include "hgb_common.cuh"

namespace hgb {

static thread_local char g_err[1024] = "";
unsigned long long g_launches = 0;

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static NcclApi g_nccl;
NcclApi &nccl() { return g_nccl; }

int nccl_load() {
    if (g_nccl.loaded) return HGB_OK;
    // If the host process already mapped an NCCL (torch bundles one) this returns it.
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) {
        set_error("NCCL not found: %s", dlerror());
        return HGB_ERR_NCCL;
    }
#define HGB_SYM(field, name)                                                     \
    g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(h, #name));    \
    if (!g_nccl.field) {                                                         \
        set_error("NCCL symbol %s missing", #name);                              \
        return HGB_ERR_NCCL;                                                     \
    }
    HGB_SYM(GetUniqueId, ncclGetUniqueId)
    HGB_SYM(CommInitRank, ncclCommInitRank)
    HGB_SYM(CommInitAll, ncclCommInitAll)
    HGB_SYM(CommDestroy, ncclCommDestroy)
    HGB_SYM(AllGather, ncclAllGather)
    HGB_SYM(AllReduce, ncclAllReduce)
    HGB_SYM(Send, ncclSend)
    HGB_SYM(Recv, ncclRecv)
    HGB_SYM(GroupStart, ncclGroupStart)
    HGB_SYM(GroupEnd, ncclGroupEnd)
    HGB_SYM(GetErrorString, ncclGetErrorString)
    HGB_SYM(GetVersion, ncclGetVersion)
#undef HGB_SYM
    g_nccl.loaded = true;
    return HGB_OK;
}

static Runtime g_rt;
Runtime &rt() { return g_rt; }

int device_count_checked(int *count) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess || c <= 0) {
        cudaGetLastError();
        set_error("no CUDA device visible (%s); libhgb200 has no CPU fallback",
                  e == cudaSuccess ? "count == 0" : cudaGetErrorString(e));
        return HGB_ERR_NOGPU;
    }
    *count = c;
    return HGB_OK;
}

// HGB_VIRTUAL_GPUS=v (test aid): shard over up to v "GPUs" even when fewer devices exist; shard i
// then lives on device i % count with its own streams, and exchanges use peer copies / peer
// stores (NCCL refuses two ranks on one device).  Lets one GPU exercise all the sharding logic.
static int virtual_gpus() {
    const char *env = getenv("HGB_VIRTUAL_GPUS");
    int v = env ? atoi(env) : 0;
    return v > 8 ? 8 : v;
}

bool oversubscribed(int g) {
    int c = 0;
    return device_count_checked(&c) == HGB_OK && g > c;
}

int resolve_ngpus(int requested, int *out) {
    int c = 0;
    HGB_TRY(device_count_checked(&c));
    int g = requested < 1 ? 1 : requested;
    if (const char *env = getenv("HGB_NGPUS")) {
        int v = atoi(env);
        if (v > 0) g = v;
    }
    const int cap = virtual_gpus() > c ? virtual_gpus() : c;
    if (g > cap) g = cap;
    *out = g;
    return HGB_OK;
}

int init_dev(DevCtx &d, int device) {
    if (d.stream) return HGB_OK;
    d.device = device;
    HGB_CUDA(cudaSetDevice(device));
    HGB_CUDA(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, device));
    HGB_CUDA(cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking));
    HGB_CUDA(cudaEventCreate(&d.ev_begin));
    HGB_CUDA(cudaEventCreate(&d.ev_end));
    HGB_CUDA(cudaEventCreateWithFlags(&d.ev_step, cudaEventDisableTiming));
    {   // boundary rows + halo exchange must win SM slots against the bulk interior kernel
        int lo = 0, hi = 0;
        HGB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        HGB_CUDA(cudaStreamCreateWithPriority(&d.stream2, cudaStreamNonBlocking, hi));
    }
    for (int i = 0; i < 2; ++i) {
        HGB_CUDA(cudaEventCreateWithFlags(&d.ev_bnd[i], cudaEventDisableTiming));
        HGB_CUDA(cudaEventCreateWithFlags(&d.ev_int[i], cudaEventDisableTiming));
    }
    HGB_CUDA(cudaEventCreateWithFlags(&d.ev_join, cudaEventDisableTiming));
    return HGB_OK;
}

int ensure_devices(int g) {
    int c = 0;
    HGB_TRY(device_count_checked(&c));
    HGB_ARG(g >= 1 && (g <= c || g <= virtual_gpus()), "gpu count out of range");
    if ((int)g_rt.devs.size() < g) g_rt.devs.resize(g);
    for (int i = 0; i < g; ++i) HGB_TRY(init_dev(g_rt.devs[i], i % c));
    HGB_CUDA(cudaSetDevice(0));
    return HGB_OK;
}

int ensure_sp_comm(int g) {
    if (g <= 1) return HGB_OK;
    HGB_TRY(ensure_devices(g));
    if (g_rt.sp_comms.count(g)) return HGB_OK;
    if (oversubscribed(g)) {
        set_error("NCCL cannot put %d ranks on fewer devices (HGB_VIRTUAL_GPUS uses peer exchange)", g);
        return HGB_ERR_STATE;
    }
    HGB_TRY(nccl_load());
    std::vector<ncclComm_t> comms(g);
    std::vector<int> ids(g);
    for (int i = 0; i < g; ++i) ids[i] = i;
    HGB_NCCL(nccl().CommInitAll(comms.data(), g, ids.data()));
    g_rt.sp_comms[g] = comms;
    return HGB_OK;
}

ncclComm_t sp_comm(int g, int index) {
    auto it = g_rt.sp_comms.find(g);
    return it == g_rt.sp_comms.end() ? nullptr : it->second[index];
}

int ensure_peer_access(int g) {
    HGB_TRY(ensure_devices(g));
    int c = 0;
    HGB_TRY(device_count_checked(&c));
    if (g > c) g = c;   // virtual shards share devices
    for (int a = 0; a < g; ++a) {
        HGB_CUDA(cudaSetDevice(a));
        for (int b = 0; b < g; ++b) {
            if (a == b) continue;
            int ok = 0;
            HGB_CUDA(cudaDeviceCanAccessPeer(&ok, a, b));
            if (!ok) {
                set_error("device %d cannot access peer %d", a, b);
                return HGB_ERR_CUDA;
            }
            cudaError_t e = cudaDeviceEnablePeerAccess(b, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else HGB_CUDA(e);
        }
    }
    HGB_CUDA(cudaSetDevice(0));
    g_rt.peer_enabled = true;
    return HGB_OK;
}

int ensure_rank_dev() {
    if (g_rt.rank_dev.stream) return HGB_OK;
    int c = 0, dev = 0;
    HGB_TRY(device_count_checked(&c));
    HGB_CUDA(cudaGetDevice(&dev));
    return init_dev(g_rt.rank_dev, dev);
}

}  // namespace hgb

using namespace hgb;

// ---- roofline probes -----------------------------------------------------
// Dependent-free DFMA streams: 8 independent accumulators per thread, enough
// warps to fill every SM sub-partition.  2 flop per DFMA.
__global__ void __launch_bounds__(256) fp64_peak_kernel(double *sink, int iters, double a, double b) {
    double r0 = threadIdx.x, r1 = r0 + 1, r2 = r0 + 2, r3 = r0 + 3;
    double r4 = r0 + 4, r5 = r0 + 5, r6 = r0 + 6, r7 = r0 + 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            r0 = fma(r0, a, b); r1 = fma(r1, a, b); r2 = fma(r2, a, b); r3 = fma(r3, a, b);
            r4 = fma(r4, a, b); r5 = fma(r5, a, b); r6 = fma(r6, a, b); r7 = fma(r7, a, b);
        }
    }
    double s = ((r0 + r1) + (r2 + r3)) + ((r4 + r5) + (r6 + r7));
    if (s == 123.456) sink[0] = s;  // never true; keeps the chain live
}

__global__ void __launch_bounds__(256) copy_kernel(const double2 *__restrict__ in,
                                                   double2 *__restrict__ out, size_t n2) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n2; i += stride) out[i] = in[i];
}

__global__ void __launch_bounds__(256) fill_kernel(double2 *out, size_t n2, double v) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n2; i += stride) out[i] = make_double2(v, v);
}

extern "C" {

const char *hgb_version(void) { return "hgb200 0.1 (sm_100a)"; }
const char *hgb_last_error(void) { return g_err; }

int hgb_device_count(void) {
    int c = 0;
    int rc = device_count_checked(&c);
    return rc == HGB_OK ? c : rc;
}

int hgb_init(void) { return ensure_devices(1); }

unsigned long long hgb_launch_count(void) { return g_launches; }

void hgb_comm_destroy(void) {
    Runtime &r = rt();
    if (r.rank_comm && nccl().loaded) nccl().CommDestroy(r.rank_comm);
    r.rank_comm = nullptr;
    if (r.tiny) { cudaFree(r.tiny); r.tiny = nullptr; }
    r.rank_mode = false;
    r.rank = 0;
    r.world = 1;
}

void hgb_finalize(void) {
    Runtime &r = rt();
    nbody_drop_cache();
    conv_drop_cache();
    hgb_comm_destroy();
    for (auto &kv : r.sp_comms)
        for (ncclComm_t c : kv.second)
            if (c && nccl().loaded) nccl().CommDestroy(c);
    r.sp_comms.clear();
    auto drop = [](DevCtx &d) {
        if (!d.stream) return;
        cudaSetDevice(d.device);
        cudaStreamSynchronize(d.stream);
        if (d.flush_buf) cudaFree(d.flush_buf);
        cudaEventDestroy(d.ev_begin);
        cudaEventDestroy(d.ev_end);
        cudaEventDestroy(d.ev_step);
        cudaStreamSynchronize(d.stream2);
        for (int i = 0; i < 2; ++i) { cudaEventDestroy(d.ev_bnd[i]); cudaEventDestroy(d.ev_int[i]); }
        cudaEventDestroy(d.ev_join);
        cudaStreamDestroy(d.stream2);
        cudaStreamDestroy(d.stream);
        d = DevCtx();
    };
    for (auto &d : r.devs) drop(d);
    r.devs.clear();
    drop(r.rank_dev);
}

static int local_devs(std::vector<DevCtx *> &out) {
    Runtime &r = rt();
    if (r.rank_mode || r.devs.empty()) {
        HGB_TRY(ensure_rank_dev());
        out.push_back(&r.rank_dev);
    }
    for (auto &d : r.devs)
        if (d.stream) out.push_back(&d);
    return HGB_OK;
}

int hgb_flush_l2(void) {
    std::vector<DevCtx *> ds;
    HGB_TRY(local_devs(ds));
    const size_t bytes = 512ull << 20;  // 4x the 126 MB L2
    for (DevCtx *d : ds) {
        HGB_CUDA(cudaSetDevice(d->device));
        if (!d->flush_buf) HGB_CUDA(cudaMalloc(&d->flush_buf, bytes));
        fill_kernel<<<148 * 8, 256, 0, d->stream>>>((double2 *)d->flush_buf, bytes / 16, 1.0);
        HGB_LAUNCHED();
    }
    for (DevCtx *d : ds) HGB_CUDA(cudaStreamSynchronize(d->stream));
    return HGB_OK;
}

// Sustained figure: after ~0.2 s of untimed warm-up launches (the clock ramps from idle), one
// timed region of ~1 s of back-to-back launches.  A millisecond-scale probe reads low (round 1:
// 8 % under the clock-true rate) because it ends before the SM clock has settled.
int hgb_measure_fp64_peak(double *tflops, double *sm_mhz_est) {
    HGB_ARG(tflops, "tflops == NULL");
    HGB_TRY(ensure_rank_dev());
    DevCtx &d = rt().rank_dev;
    HGB_CUDA(cudaSetDevice(d.device));
    cudaDeviceProp prop;
    HGB_CUDA(cudaGetDeviceProperties(&prop, d.device));
    double *sink = nullptr;
    HGB_CUDA(cudaMalloc(&sink, 64));
    const int blocks = prop.multiProcessorCount * 8, iters = 40960;   // ~43 ms per launch at 1.965 GHz
    const int warm = 5, timed = 24;
    for (int rep = 0; rep < warm + timed; ++rep) {
        if (rep == warm) HGB_CUDA(cudaEventRecord(d.ev_begin, d.stream));
        fp64_peak_kernel<<<blocks, 256, 0, d.stream>>>(sink, iters, 0.999999, 1e-9);
        HGB_LAUNCHED();
    }
    HGB_CUDA(cudaEventRecord(d.ev_end, d.stream));
    HGB_CUDA(cudaEventSynchronize(d.ev_end));
    float ms = 0;
    HGB_CUDA(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
    cudaFree(sink);
    const double dfma = (double)timed * blocks * 256.0 * iters * 64.0;
    *tflops = 2.0 * dfma / (ms * 1e-3) / 1e12;
    // 64 DFMA lanes per SM per clock (4 sub-partitions x 16)
    if (sm_mhz_est) *sm_mhz_est = dfma / (ms * 1e-3) / (64.0 * prop.multiProcessorCount) / 1e6;
    return HGB_OK;
}

int hgb_measure_hbm_copy(double *gbs) {
    HGB_ARG(gbs, "gbs == NULL");
    HGB_TRY(ensure_rank_dev());
    DevCtx &d = rt().rank_dev;
    HGB_CUDA(cudaSetDevice(d.device));
    const size_t bytes = 2ull << 30;
    double2 *a = nullptr, *b = nullptr;
    HGB_CUDA(cudaMalloc(&a, bytes));
    HGB_CUDA(cudaMalloc(&b, bytes));
    fill_kernel<<<148 * 8, 256, 0, d.stream>>>(a, bytes / 16, 2.0);
    HGB_LAUNCHED();
    float best = 1e30f;
    for (int rep = 0; rep < 6; ++rep) {
        HGB_CUDA(cudaEventRecord(d.ev_begin, d.stream));
        copy_kernel<<<148 * 16, 256, 0, d.stream>>>(a, b, bytes / 16);
        HGB_LAUNCHED();
        HGB_CUDA(cudaEventRecord(d.ev_end, d.stream));
        HGB_CUDA(cudaEventSynchronize(d.ev_end));
        float ms = 0;
        HGB_CUDA(cudaEventElapsedTime(&ms, d.ev_begin, d.ev_end));
        if (rep > 0 && ms < best) best = ms;
    }
    cudaFree(a);
    cudaFree(b);
    *gbs = 2.0 * (double)bytes / (best * 1e-3) / 1e9;
    return HGB_OK;
}

// ---- one process per GPU -------------------------------------------------
int hgb_comm_unique_id(void *id_out, int capacity) {
    HGB_ARG(id_out && capacity >= (int)sizeof(ncclUniqueId), "id buffer too small (need 128 bytes)");
    HGB_TRY(nccl_load());
    ncclUniqueId id;
    HGB_NCCL(nccl().GetUniqueId(&id));
    memset(id_out, 0, capacity);
    memcpy(id_out, &id, sizeof(id));
    return HGB_OK;
}

int hgb_comm_init_rank(const void *id, int id_bytes, int rank, int world) {
    HGB_ARG(world >= 1 && rank >= 0 && rank < world, "rank/world");
    Runtime &r = rt();
    HGB_TRY(ensure_rank_dev());
    hgb_comm_destroy();
    r.rank_mode = true;
    r.rank = rank;
    r.world = world;
    if (world == 1) return HGB_OK;
    HGB_ARG(id && id_bytes >= (int)sizeof(ncclUniqueId), "unique id missing");
    HGB_TRY(nccl_load());
    ncclUniqueId uid;
    memcpy(&uid, id, sizeof(uid));
    HGB_CUDA(cudaSetDevice(r.rank_dev.device));
    HGB_NCCL(nccl().CommInitRank(&r.rank_comm, world, uid, rank));
    HGB_CUDA(cudaMalloc(&r.tiny, 64));
    HGB_CUDA(cudaMemset(r.tiny, 0, 64));
    return HGB_OK;
}

int hgb_comm_rank(void) { return rt().rank; }
int hgb_comm_world(void) { return rt().world; }

int hgb_comm_barrier(void) {
    Runtime &r = rt();
    if (!r.rank_mode || r.world == 1) {
        if (r.rank_dev.stream) HGB_CUDA(cudaStreamSynchronize(r.rank_dev.stream));
        return HGB_OK;
    }
    HGB_NCCL(nccl().AllReduce(r.tiny, r.tiny, 1, ncclDouble, ncclSum, r.rank_comm, r.rank_dev.stream));
    HGB_CUDA(cudaStreamSynchronize(r.rank_dev.stream));
    return HGB_OK;
}

}  // extern "C"
