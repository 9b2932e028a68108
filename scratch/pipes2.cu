// scratch/pipes2.cu -- which instructions share the ALU pipe with VIADDMNMX on sm_100a?
// For each candidate X: rate of X alone, and of a 1:1 mix VIADDMNMX.U16x2 : X on independent chains.
// mix ~ 2x single rate => X issues to another pipe; mix ~ 1x => X competes for the ALU pipe.
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

constexpr int CH = 8;
constexpr int INNER = 32;

template <int X>
__device__ __forceinline__ uint32_t opx(uint32_t a, uint32_t b, uint32_t k1, uint32_t one) {
    uint32_t d = a;
    if (X == 0) d = __viaddmax_u16x2(a, k1, b);
    if (X == 1) d = __vmaxu2(a, b);
    if (X == 2) d = __vmaxs2(a, b);
    if (X == 3) d = __vadd2(a, b);
    if (X == 4) asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    if (X == 5) asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(d) : "r"(a), "r"(b));
    if (X == 6) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(d) : "r"(a), "r"(b), "r"(k1));
    if (X == 7) d = __vimax3_u16x2(a, b, k1);
    if (X == 8) asm volatile("max.s32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    if (X == 9) asm volatile("max.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    if (X == 10) { __half2 x = *reinterpret_cast<__half2*>(&a), y = *reinterpret_cast<__half2*>(&b); __half2 r = __hmax2(x, y); d = *reinterpret_cast<uint32_t*>(&r); }
    if (X == 11) { __half2 x = *reinterpret_cast<__half2*>(&a), y = *reinterpret_cast<__half2*>(&b); __half2 r = __hadd2(x, y); d = *reinterpret_cast<uint32_t*>(&r); }
    if (X == 12) { float r = fmaxf(__uint_as_float(a), __uint_as_float(b)); d = __float_as_uint(r); }
    if (X == 13) { float r = __uint_as_float(a) + __uint_as_float(b); d = __float_as_uint(r); }
    if (X == 14) asm volatile("add.s32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    if (X == 15) asm volatile("vadd2.s32.s32.s32.sat %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(0));
    if (X == 16) d = __vimax_s16x2_relu(a, b);
    if (X == 17) asm volatile("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(k1));
    if (X == 18) d = (k1 & 4) ? a : (a ^ b);
    return d;
}

// MIX: 0 = X alone, 1 = 1:1 with VIADDMNMX, 2 = 2 VIADDMNMX : 1 X
template <int X, int MIX>
__global__ void __launch_bounds__(256) probe(uint32_t* out, int iters, uint32_t k1, uint32_t one) {
    uint32_t a[CH], b[CH], c[CH], e[CH];
    for (int i = 0; i < CH; i++) { a[i] = threadIdx.x * 3 + i; b[i] = threadIdx.x * 7 + i + k1; c[i] = threadIdx.x + 11 * i; e[i] = i * 5 + threadIdx.x; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < INNER; k++) {
#pragma unroll
            for (int i = 0; i < CH; i++) {
                a[i] = opx<X>(a[i], b[i], k1, one);
                if (MIX >= 1) c[i] = __viaddmax_u16x2(c[i], k1, b[i]);
                if (MIX >= 2) e[i] = __viaddmax_u16x2(e[i], k1, b[i]);
            }
        }
    }
    uint32_t acc = 0;
    for (int i = 0; i < CH; i++) acc ^= a[i] ^ b[i] ^ c[i] ^ e[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}

template <int X, int MIX>
double run() {
    int sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 4, threads = 256, iters = 300;
    uint32_t* out;
    cudaMalloc(&out, (size_t)blocks * threads * 4);
    probe<X, MIX><<<blocks, threads>>>(out, 10, 0xffffffffu, 1u);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaEventRecord(e0);
        probe<X, MIX><<<blocks, threads>>>(out, iters, 0xffffffffu, 1u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaFree(out);
    const double n = (double)blocks * threads * iters * INNER * CH * (1 + MIX);
    return n / (best * 1e-3);
}

template <int X>
void row(const char* name) {
    const double alone = run<X, 0>(), mix1 = run<X, 1>(), mix2 = run<X, 2>();
    printf("%-28s alone %.3e   1:1 with VIADDMNMX %.3e (x%.2f of 1.82e13)   2 VIADDMNMX : 1 X %.3e (x%.2f)  %s\n", name, alone, mix1,
           mix1 / 1.82e13, mix2, mix2 / 1.82e13, cudaGetErrorString(cudaGetLastError()));
}

int main() {
    row<0>("VIADDMNMX.U16x2");
    row<1>("VIMNMX.U16x2 (2-input)");
    row<2>("VIMNMX.S16x2 (2-input)");
    row<16>("VIMNMX.S16x2.RELU");
    row<7>("VIMNMX3.U16x2");
    row<3>("VIADD.16x2 (vadd2)");
    row<4>("IMAD");
    row<5>("PRMT");
    row<6>("LOP3");
    row<17>("SHF");

    row<8>("max.s32");
    row<9>("max.u32");
    row<14>("add.s32");
    row<10>("HMNMX2 (hmax2)");
    row<11>("HADD2");
    row<12>("FMNMX");
    row<13>("FADD");
    return 0;
}
