// hgb_noalloc.cu -- hgb_nbody_energy_static(): the potential energy of bodies that already sit in
// device memory, computed without allocating anything.
//
// Who this is for.  Every other entry point of libhgb200.so works on a handle, and creating a handle
// calls cudaMalloc.  Some callers must not do that:
//   * an in-situ diagnostic hook inside a simulation that has already handed all of the GPU's memory
//     to its own pool at start-up -- a further cudaMalloc would simply fail there;
//   * a monitoring call made while the host code's step loop is in flight -- cudaMalloc and cudaFree
//     synchronise the whole device, which stalls that loop;
//   * host code that is only allowed to pass device pointers it was given (a plugin under a
//     framework that owns allocation and accounting).
// Such a caller already has the positions on the device (Float3D AoS, the reference's layout,
// src/nbody/nbody-util.h) and one double of device memory for the answer.  The only other memory the
// energy reduction needs is one partial sum per block of 128 targets; here that lives in
// `g_energy_scratch` below, a fixed array that is part of the library image.  Cost: 64 KiB of device
// memory on every GPU that loads the library, held until the process exits.  Limit:
// HGB_STATIC_ENERGY_MAX_BODIES bodies per call.
//
// The arithmetic is that of hgb_nbody_energy (nbody_energy_kernel<128> + sum_partials_kernel in
// nbody_kernels.cuh): same tiles, same Kahan accumulation per target, same fixed-order block and
// final sums -- so on one GPU the two entry points return the same bits.  Only the addressing differs
// (AoS straight from the caller instead of the handle's SoA copy) and where the partials go.
#include <mutex>

#include "hgb_common.cuh"

namespace hgb {

// nbody_kernels.cuh defines non-template kernels and belongs to hgb_nbody.cu alone, so the two
// constants are restated here (reference src/nbody/nbody-run.h:12-13)
static constexpr double kG = 9.8;
static constexpr double kEps = 0.005;
constexpr int kEnergyBlock = 128;
constexpr int kEnergyMaxBlocks = HGB_STATIC_ENERGY_MAX_BODIES / kEnergyBlock;  // 8192 partials = 64 KiB
static_assert(HGB_STATIC_ENERGY_MAX_BODIES % kEnergyBlock == 0, "limit is a whole number of blocks");

// One partial per block of targets.  Shared by every call on this device and never cleared: a call
// with `grid` blocks writes slots [0, grid) in its first kernel and reads exactly those in its
// second, so what an earlier (possibly larger) call left behind is never looked at.
// In HOST code this name is only a handle for the runtime, not a device address -- host code gets
// the address from cudaGetSymbolAddress (static_scratch_on_current_device below).
static __device__ __align__(256) double g_energy_scratch[kEnergyMaxBlocks];

// nbody_energy_kernel<128, false> with Float3D AoS addressing: body j is aos[3j .. 3j+2]
__global__ void __launch_bounds__(kEnergyBlock)
nbody_energy_aos_kernel(const double *__restrict__ aos, int n, double *__restrict__ block_sums) {
    constexpr int TILE = kEnergyBlock;
    __shared__ double shx[TILE], shy[TILE], shz[TILE];
    __shared__ double warp_sums[kEnergyBlock / 32];
    const int k = blockIdx.x * kEnergyBlock + threadIdx.x;
    const bool active = k < n;
    const int body = active ? k : 0;
    const double px = aos[3 * (size_t)body], py = aos[3 * (size_t)body + 1], pz = aos[3 * (size_t)body + 2];
    double total = 0.0, comp = 0.0;  // Kahan over tile partials
    for (int base = 0; base < n; base += TILE) {
        const int j = base + threadIdx.x;
        shx[threadIdx.x] = j < n ? aos[3 * (size_t)j] : 0.0;
        shy[threadIdx.x] = j < n ? aos[3 * (size_t)j + 1] : 0.0;
        shz[threadIdx.x] = j < n ? aos[3 * (size_t)j + 2] : 0.0;
        __syncthreads();
        const int cnt = min(TILE, n - base);
        double part = 0.0;
#pragma unroll 4
        for (int jj = 0; jj < cnt; ++jj) {
            double dx = shx[jj] - px, dy = shy[jj] - py, dz = shz[jj] - pz;
            double s = fma(dz, dz, fma(dy, dy, fma(dx, dx, kEps)));
            double r = rsqrt(s);
            part += (base + jj == body) ? 0.0 : r;
        }
        double yk = part - comp;
        double tk = total + yk;
        comp = (tk - total) - yk;
        total = tk;
        __syncthreads();
    }
    double v = active ? total : 0.0;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int i = 0; i < kEnergyBlock / 32; ++i) s += warp_sums[i];
        block_sums[blockIdx.x] = s;
    }
}

// sum_partials_kernel's fixed order, then U = -G/2 * sum (every pair was counted twice)
__global__ void __launch_bounds__(256)
energy_finish_kernel(const double *__restrict__ parts, int m, double *__restrict__ energy) {
    __shared__ double sh[256];
    double s = 0.0;
    for (int i = threadIdx.x; i < m; i += 256) s += parts[i];
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
        if (threadIdx.x < off) sh[threadIdx.x] += sh[threadIdx.x + off];
        __syncthreads();
    }
    if (threadIdx.x == 0) energy[0] = -kG * 0.5 * (0.0 + sh[0]);
}

// Device address of g_energy_scratch on the CURRENT device, looked up once per device.  Each device
// that loads this module has its own copy of the array at its own address, hence the table.
static int static_scratch_on_current_device(double **out) {
    constexpr int kMaxDevices = 64;
    static double *addr[kMaxDevices] = {};
    static std::mutex lock;
    int dev = -1;
    HGB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("hgb_nbody_energy_static: device ordinal %d outside the table of %d", dev, kMaxDevices);
        return HGB_ERR_STATE;
    }
    std::lock_guard<std::mutex> guard(lock);
    if (!addr[dev]) {
        double *p = nullptr;
        HGB_CUDA(cudaGetSymbolAddress((void **)&p, g_energy_scratch));
        if (!p) { set_error("cudaGetSymbolAddress returned a null scratch address"); return HGB_ERR_CUDA; }
        addr[dev] = p;
    }
    *out = addr[dev];
    return HGB_OK;
}

// the caller's pointers must be memory a kernel on this device can dereference
static int require_device_pointer(const void *p, const char *what) {
    cudaPointerAttributes a{};
    cudaError_t e = cudaPointerGetAttributes(&a, p);
    if (e != cudaSuccess || (a.type != cudaMemoryTypeDevice && a.type != cudaMemoryTypeManaged)) {
        cudaGetLastError();
        set_error("hgb_nbody_energy_static: %s must point to device (or managed) memory; this entry point "
                  "copies nothing and allocates nothing", what);
        return HGB_ERR_ARG;
    }
    return HGB_OK;
}

}  // namespace hgb

using namespace hgb;

extern "C" int hgb_nbody_energy_static(const double *pos_aos_dev, int n, double *energy_dev, void *stream) {
    HGB_ARG(pos_aos_dev && energy_dev, "positions / energy (device pointers)");
    HGB_ARG(n >= 1, "n >= 1");
    if (n > HGB_STATIC_ENERGY_MAX_BODIES) {
        set_error("hgb_nbody_energy_static: n = %d exceeds the built-in scratch (HGB_STATIC_ENERGY_MAX_BODIES = %d "
                  "bodies); use hgb_nbody_create + hgb_nbody_energy, which allocate, for larger sets",
                  n, HGB_STATIC_ENERGY_MAX_BODIES);
        return HGB_ERR_ARG;
    }
    int count = 0;
    HGB_TRY(device_count_checked(&count));
    HGB_TRY(require_device_pointer(pos_aos_dev, "pos_aos_dev"));
    HGB_TRY(require_device_pointer(energy_dev, "energy_dev"));
    double *scratch = nullptr;
    HGB_TRY(static_scratch_on_current_device(&scratch));  // first call on a device: symbol lookup
    const cudaStream_t st = (cudaStream_t)stream;
    const int grid = (n + kEnergyBlock - 1) / kEnergyBlock;
    nbody_energy_aos_kernel<<<grid, kEnergyBlock, 0, st>>>(pos_aos_dev, n, scratch);
    HGB_LAUNCHED();
    energy_finish_kernel<<<1, 256, 0, st>>>(scratch, grid, energy_dev);
    HGB_LAUNCHED();
    return HGB_OK;
}
