// hgb_common.cuh -- shared host-side plumbing of libhgb200.so: error channel,
// per-device contexts (stream + events), lazily loaded NCCL, launch counter.
#pragma once
#include <cuda_runtime.h>
#include <nccl.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <string>
#include <vector>

#include "../../include/hgb200.h"

namespace hgb {

// ---- error channel (hgb_last_error) -------------------------------------
void set_error(const char *fmt, ...);
extern unsigned long long g_launches;

#define HGB_CUDA(call)                                                                  \
    do {                                                                                \
        cudaError_t e_ = (call);                                                        \
        if (e_ != cudaSuccess) {                                                        \
            hgb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,                \
                           cudaGetErrorString(e_));                                     \
            return HGB_ERR_CUDA;                                                        \
        }                                                                               \
    } while (0)

#define HGB_NCCL(call)                                                                  \
    do {                                                                                \
        ncclResult_t r_ = (call);                                                       \
        if (r_ != ncclSuccess) {                                                        \
            hgb::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call,                \
                           hgb::nccl().GetErrorString(r_));                             \
            return HGB_ERR_NCCL;                                                        \
        }                                                                               \
    } while (0)

#define HGB_TRY(call)                    \
    do {                                 \
        int rc_ = (call);                \
        if (rc_ != HGB_OK) return rc_;   \
    } while (0)

#define HGB_ARG(cond, msg)                           \
    do {                                             \
        if (!(cond)) {                               \
            hgb::set_error("bad argument: %s", msg); \
            return HGB_ERR_ARG;                      \
        }                                            \
    } while (0)

// every kernel launch goes through this so hgb_launch_count() is honest
#define HGB_LAUNCHED()                                  \
    do {                                                \
        ++hgb::g_launches;                              \
        HGB_CUDA(cudaGetLastError());                   \
    } while (0)

// ---- NCCL, resolved at run time (single-GPU use never needs it) ---------
struct NcclApi {
    bool loaded = false;
    decltype(&ncclGetUniqueId) GetUniqueId = nullptr;
    decltype(&ncclCommInitRank) CommInitRank = nullptr;
    decltype(&ncclCommInitAll) CommInitAll = nullptr;
    decltype(&ncclCommDestroy) CommDestroy = nullptr;
    decltype(&ncclAllGather) AllGather = nullptr;
    decltype(&ncclAllReduce) AllReduce = nullptr;
    decltype(&ncclSend) Send = nullptr;
    decltype(&ncclRecv) Recv = nullptr;
    decltype(&ncclGroupStart) GroupStart = nullptr;
    decltype(&ncclGroupEnd) GroupEnd = nullptr;
    decltype(&ncclGetErrorString) GetErrorString = nullptr;
    decltype(&ncclGetVersion) GetVersion = nullptr;
};
NcclApi &nccl();
int nccl_load();  // HGB_OK or HGB_ERR_NCCL

// ---- devices -------------------------------------------------------------
struct DevCtx {
    int device = -1;
    int sms = 0;                    // multiprocessor count of `device`
    cudaStream_t stream = nullptr;  // all kernels + collectives of this device
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_step = nullptr;
    cudaStream_t stream2 = nullptr; // boundary rows + halo exchange, overlapped with the interior
    cudaEvent_t ev_bnd[2] = {nullptr, nullptr}, ev_int[2] = {nullptr, nullptr}, ev_join = nullptr;
    void *flush_buf = nullptr;      // L2 flush scratch
};

struct Runtime {
    // single-process mode: devices 0..g-1.  A deque: handles keep DevCtx* into it, and growing it
    // for a handle with more GPUs must not move the contexts older handles point at.
    std::deque<DevCtx> devs;
    // ncclCommInitAll communicator sets, one per GPU count in use (a g=2 handle and a g=4 handle can
    // alternate without tearing communicators down)
    std::map<int, std::vector<ncclComm_t>> sp_comms;
    bool peer_enabled = false;
    // one-process-per-GPU mode
    ncclComm_t rank_comm = nullptr;
    int rank = 0, world = 1;
    DevCtx rank_dev;                // the calling process's device
    bool rank_mode = false;
    double *tiny = nullptr;         // barrier scratch (rank mode)
};
Runtime &rt();

int device_count_checked(int *count);       // HGB_ERR_NOGPU when there is none
int resolve_ngpus(int requested, int *out); // clamp + HGB_NGPUS override
bool oversubscribed(int g);                 // g shards on fewer devices (HGB_VIRTUAL_GPUS)
int ensure_devices(int g);                  // streams/events for devices 0..g-1
int ensure_sp_comm(int g);                  // ncclCommInitAll over devices 0..g-1 (cached per g)
ncclComm_t sp_comm(int g, int index);       // communicator of device `index` in the set of g (after ensure_sp_comm)
int ensure_peer_access(int g);
int ensure_rank_dev();                      // stream/events on the current device
int init_dev(DevCtx &d, int device);
void nbody_drop_cache();                   // cached handles of the drop-in entry points
void conv_drop_cache();

// even split of `total` units into `parts` shards of equal padded size
static inline int shard_size(int total, int parts, int align) {
    long per = ((long)total + parts - 1) / parts;
    per = (per + align - 1) / align * align;
    return (int)per;
}

}  // namespace hgb
