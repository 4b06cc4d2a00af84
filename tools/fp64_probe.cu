// dev probe: FP64 pipe on sm_100a -- dependent-issue latency and the rate one warp / W warps per SM
// sub-partition reach with P independent chains.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
// -o tools/fp64_probe.bin tools/fp64_probe.cu ; run: tools/fp64_probe.bin
#include <cstdio>
#include <cuda_runtime.h>

template <int P, int KIND>   // KIND 0: DFMA, 1: DADD, 2: DMUL
__global__ void chains(double *out, long long *cyc, int iters, double seed) {
    double a[P];
#pragma unroll
    for (int p = 0; p < P; ++p) a[p] = seed + p + threadIdx.x;
    const double m = 1.0000001, c = 1e-9;
    long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
#pragma unroll
            for (int p = 0; p < P; ++p) {
                if (KIND == 0) a[p] = __fma_rn(a[p], m, c);
                else if (KIND == 1) a[p] = __dadd_rn(a[p], c);
                else a[p] = __dmul_rn(a[p], m);
            }
        }
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int p = 0; p < P; ++p) s += a[p];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int P, int KIND>
void run(int warps_per_smsp, double *out, long long *cyc) {
    const int iters = 4000;
    chains<P, KIND><<<148, 128 * warps_per_smsp>>>(out, cyc, iters, 1.0);   // 4 SMSPs x warps
    cudaDeviceSynchronize();
    long long h = 0;
    cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
    const double ops = (double)iters * 8 * P;   // per warp
    printf("kind=%d P=%2d warps/SMSP=%d: %7.2f cycles per FP64 instr per warp, %6.2f per SMSP\n", KIND, P,
           warps_per_smsp, h / ops, h / ops / warps_per_smsp);
}

int main() {
    double *out; long long *cyc;
    cudaMalloc(&out, 148 * 1024 * 8); cudaMalloc(&cyc, 8);
    for (int w : {1, 2, 3, 4, 6, 8}) {
        run<1, 0>(w, out, cyc); run<2, 0>(w, out, cyc); run<4, 0>(w, out, cyc); run<8, 0>(w, out, cyc);
        run<16, 0>(w, out, cyc);
    }
    run<1, 1>(1, out, cyc); run<4, 1>(1, out, cyc); run<8, 1>(1, out, cyc);
    run<1, 2>(1, out, cyc); run<4, 2>(1, out, cyc); run<8, 2>(1, out, cyc);
    printf("%s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
