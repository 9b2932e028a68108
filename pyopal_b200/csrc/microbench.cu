// pyopal_b200/csrc/microbench.cu -- integer-ALU-pipe microbenchmark behind the
// cell-update roofline.  Every maximum of the DP cell
// (VIADDMNMX / VIMNMX3 / VIMNMX, 16x2 or 32 bit) and the PRMT that widens the
// profile bytes issue down the SM's ALU pipe at one warp instruction per two
// clocks per SM sub-partition (64 lanes/clk/SM), whatever the mix; this kernel
// measures that rate with independent VIADDMNMX.U16x2 chains.  A 16-bit cell
// pair needs 5.5 such instructions (kernels.cuh), its two additions being
// IMADs on the FMA pipe, so roof = rate * 2 cells / 5.5.
#include "../../include/pyopal_cuda.h"
#include <cuda_runtime.h>
#include <stdint.h>

namespace {
constexpr int CHAINS = 8;
constexpr int INNER = 64;
__global__ void __launch_bounds__(256) issue_kernel(uint32_t* out, int iters, uint32_t seed, uint32_t one) {
    uint32_t e[CHAINS], f[CHAINS], g[CHAINS], b = one;
    const uint32_t negext = 0xffffffffu;
    for (int c = 0; c < CHAINS; c++) { e[c] = seed + c; f[c] = seed * 3 + c; g[c] = seed * 7 + threadIdx.x + c; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < INNER; k++) {
#pragma unroll
            for (int c = 0; c < CHAINS; c++) {
                e[c] = __viaddmax_u16x2(e[c], negext, g[c]);
                f[c] = __viaddmax_u16x2(f[c], negext, g[c]);
            }
        }
    }
    uint32_t acc = b;
    for (int c = 0; c < CHAINS; c++) acc ^= e[c] ^ f[c] ^ g[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
}  // namespace

extern "C" int pyopal_cuda_int_issue_rate(int device, double* thread_instr_per_s, double* sm_clock_mhz) {
    if (cudaSetDevice(device) != cudaSuccess) return OPAL_ERR_NO_SIMD_SUPPORT;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return PYOPAL_CUDA_ERR_CUDA;
    const int blocks = prop.multiProcessorCount * 8, threads = 256, iters = 200;
    uint32_t* d = nullptr;
    if (cudaMalloc(&d, (size_t)blocks * threads * sizeof(uint32_t)) != cudaSuccess) return PYOPAL_CUDA_ERR_NOMEM;
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    issue_kernel<<<blocks, threads>>>(d, 10, 1u, 1u);  // warm-up
    cudaDeviceSynchronize();
    float best_ms = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(a);
        issue_kernel<<<blocks, threads>>>(d, iters, 1u + rep, 1u);
        cudaEventRecord(b);
        cudaEventSynchronize(b);
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        if (ms < best_ms) best_ms = ms;
    }
    cudaError_t e = cudaGetLastError();
    cudaEventDestroy(a); cudaEventDestroy(b); cudaFree(d);
    if (e != cudaSuccess) return PYOPAL_CUDA_ERR_CUDA;
    const double instr = (double)blocks * threads * (double)iters * INNER * CHAINS * 2.0;
    if (thread_instr_per_s) *thread_instr_per_s = instr / (best_ms * 1e-3);
    if (sm_clock_mhz) *sm_clock_mhz = prop.clockRate / 1000.0;
    return 0;
}
