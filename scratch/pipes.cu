// scratch/pipes.cu -- per-instruction issue-rate probe for the DP cell mix on sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/pipes scratch/pipes.cu
// Prints warp-instructions per clock per SM for each op / mix (8 independent chains per thread).
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int CH = 8;
constexpr int INNER = 32;

template <int OP>
__device__ __forceinline__ void body(uint32_t (&a)[CH], uint32_t (&b)[CH], uint32_t (&c)[CH], uint32_t k1, uint32_t k2,
                                     const uint32_t* sm, uint32_t& acc) {
#pragma unroll
    for (int i = 0; i < CH; i++) {
        if (OP == 0) a[i] = __viaddmax_s16x2(a[i], k1, b[i]);
        if (OP == 2) a[i] = __vimax3_s16x2(a[i], b[i], c[i]);
        if (OP == 3) a[i] = __vadd2(a[i], k1);
        if (OP == 4) {  // fibonacci-style so ptxas cannot strength-reduce the chain
            asm volatile("add.s32 %0, %1, %2;" : "=r"(a[i]) : "r"(a[i]), "r"(b[i]));
            asm volatile("add.s32 %0, %1, %2;" : "=r"(b[i]) : "r"(b[i]), "r"(a[i]));
        }
        if (OP == 5) asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(a[i]) : "r"(a[i]), "r"(k2), "r"(k1));
        if (OP == 6) asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(a[i]) : "r"(a[i]), "r"(b[i]));
        if (OP == 7) a[i] = __viaddmax_u16x2(a[i], k1, b[i]);
        if (OP == 8) a[i] = __vimax3_u16x2(a[i], b[i], c[i]);
        if (OP == 9) {  // current cell mix (signed): E, F, Td+relu, max relu, sub, best/2, PRMT
            uint32_t s;
            asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(s) : "r"(c[i]), "r"(a[i]));
            a[i] = __viaddmax_s16x2(a[i], k1, c[i]);
            b[i] = __viaddmax_s16x2(b[i], k1, c[(i + 1) % CH]);
            uint32_t h = __viaddmax_s16x2_relu(c[i], s, a[i]);
            h = __vimax_s16x2_relu(h, b[i]);
            c[i] = __vsub2(h, k2);
            if (i & 1) acc = __vimax3_s16x2(acc, h, c[i]);
        }
        if (OP == 10) {  // biased mix: the H-open subtract as a 32-bit IMAD (fma pipe)
            uint32_t s;
            asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(s) : "r"(c[i]), "r"(a[i]));
            a[i] = __viaddmax_u16x2(a[i], k1, c[i]);
            b[i] = __viaddmax_u16x2(b[i], k1, c[(i + 1) % CH]);
            uint32_t h = __viaddmax_u16x2(c[i], s, a[i]);
            h = __vimax3_u16x2(h, b[i], k2);
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(h), "r"(k2 >> 20), "r"(k1));
            if (i & 1) acc = __vimax3_u16x2(acc, h, c[i]);
        }
        if (OP == 11) {  // biased mix, both adds on the fma pipe: Td = Gd + S via IMAD, then max3
            uint32_t s;
            asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(s) : "r"(c[i]), "r"(a[i]));
            a[i] = __viaddmax_u16x2(a[i], k1, c[i]);
            b[i] = __viaddmax_u16x2(b[i], k1, c[(i + 1) % CH]);
            uint32_t td;
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(td) : "r"(c[i]), "r"(k2 >> 20), "r"(s));
            uint32_t h = __vimax3_u16x2(td, a[i], b[i]);
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(h), "r"(k2 >> 20), "r"(k1));
            if (i & 1) acc = __vimax3_u16x2(acc, h, c[i]);
        }
        if (OP == 12) {  // LDS.32 alone (conflict-free: lane-linear)
            uint32_t v = sm[(a[i] & 1023u)];
            a[i] = v + i;
        }
        if (OP == 13) {  // current mix + its LDS share (0.5 LDS.32 per cell pair)
            uint32_t s;
            if ((i & 1) == 0) b[i] ^= sm[(c[i] & 1023u)];
            asm volatile("prmt.b32 %0, %1, %2, 0xC480;" : "=r"(s) : "r"(c[i]), "r"(a[i]));
            a[i] = __viaddmax_s16x2(a[i], k1, c[i]);
            b[i] = __viaddmax_s16x2(b[i], k1, c[(i + 1) % CH]);
            uint32_t h = __viaddmax_s16x2_relu(c[i], s, a[i]);
            h = __vimax_s16x2_relu(h, b[i]);
            c[i] = __vsub2(h, k2);
            if (i & 1) acc = __vimax3_s16x2(acc, h, c[i]);
        }
        if (OP == 14) {  // 4 DPX + 2 IMAD, no PRMT (upper bound of the biased scheme with pre-widened S)
            a[i] = __viaddmax_u16x2(a[i], k1, c[i]);
            b[i] = __viaddmax_u16x2(b[i], k1, c[(i + 1) % CH]);
            uint32_t td;
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(td) : "r"(c[i]), "r"(k2 >> 20), "r"(b[(i + 3) % CH]));
            uint32_t h = __vimax3_u16x2(td, a[i], b[i]);
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(h), "r"(k2 >> 20), "r"(k1));
            if (i & 1) acc = __vimax3_u16x2(acc, h, c[i]);
        }
        if (OP == 15) a[i] = __vimax_s16x2_relu(a[i], b[i]);
        if (OP == 16) a[i] = __viaddmax_s16x2_relu(a[i], k1, b[i]);
        if (OP == 17) asm volatile("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(a[i]) : "r"(a[i]), "r"(b[i]), "r"(k1));
        if (OP == 18) asm volatile("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(a[i]) : "r"(a[i]), "r"(b[i]), "r"(k1));
        if (OP == 19) a[i] = __viaddmax_s32(a[i], k1, b[i]);
        if (OP == 20) a[i] = __vimax3_s32(a[i], b[i], c[i]);
        if (OP == 21) {  // half ALU (viaddmax) + half IMAD, independent
            a[i] = __viaddmax_s16x2(a[i], k1, b[i]);
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(c[i]), "r"(k2 >> 20), "r"(k1));
        }
        if (OP == 22) {  // 2 ALU : 1 IMAD
            a[i] = __viaddmax_s16x2(a[i], k1, b[i]);
            b[i] = __viaddmax_s16x2(b[i], k1, a[i]);
            asm volatile("mad.lo.s32 %0, %1, %2, %3;" : "=r"(c[i]) : "r"(c[i]), "r"(k2 >> 20), "r"(k1));
        }
    }
}

__device__ int instr_per_body(int op) {
    switch (op) {
        case 9: return CH * 6 + CH / 2;
        case 10: return CH * 6 + CH / 2;
        case 11: return CH * 6 + CH / 2;
        case 13: return CH * 6 + CH / 2 + CH / 2 * 2;
        case 14: return CH * 5 + CH / 2;
        case 12: return CH * 3;
        case 21: return CH * 2;
        case 22: return CH * 3;
        case 1: return CH * 2;
        default: return CH;
    }
}

template <int OP>
__global__ void __launch_bounds__(256) probe(uint32_t* out, long long* cyc, int iters, uint32_t k1, uint32_t k2) {
    __shared__ uint32_t sm[1024];
    for (int x = threadIdx.x; x < 1024; x += blockDim.x) sm[x] = x * 2654435761u;
    __syncthreads();
    uint32_t a[CH], b[CH], c[CH], acc = 0;
    for (int i = 0; i < CH; i++) { a[i] = threadIdx.x * 3 + i + k1; b[i] = threadIdx.x * 7 + i; c[i] = threadIdx.x + 11 * i; }
    long long t0 = clock64();
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int k = 0; k < INNER; k++) body<OP>(a, b, c, k1, k2, sm, acc);
    }
    long long t1 = clock64();
    for (int i = 0; i < CH; i++) acc ^= a[i] ^ b[i] ^ c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int OP>
void run(const char* name, int per_body, int blocks_per_sm) {
    int dev = 0, sms = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int blocks = sms * blocks_per_sm, threads = 256, iters = 400;
    uint32_t* out; long long* cyc;
    cudaMalloc(&out, (size_t)blocks * threads * 4);
    cudaMalloc(&cyc, (size_t)blocks * 8);
    probe<OP><<<blocks, threads>>>(out, cyc, 10, 0xffffffffu, 0x00100010u);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<OP><<<blocks, threads>>>(out, cyc, iters, 0xffffffffu, 0x00100010u);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    long long* h = new long long[blocks];
    cudaMemcpy(h, cyc, (size_t)blocks * 8, cudaMemcpyDeviceToHost);
    double mean = 0; for (int i = 0; i < blocks; i++) mean += (double)h[i]; mean /= blocks;
    const double warp_instr_per_block = (double)(threads / 32) * iters * INNER * per_body;
    // all blocks_per_sm blocks of an SM run concurrently for ~mean cycles
    const double ipc_sm = warp_instr_per_block * blocks_per_sm / mean;
    const double total_thread_instr = warp_instr_per_block * 32.0 * blocks;
    printf("%-34s op=%2d  warp-instr/clk/SM = %6.3f  (lanes/clk/SM %6.1f)  %7.3f ms  %.3e thread-instr/s  err=%s\n", name, OP, ipc_sm,
           ipc_sm * 32, ms, total_thread_instr / (ms * 1e-3), cudaGetErrorString(cudaGetLastError()));
    delete[] h; cudaFree(out); cudaFree(cyc);
}

int main() {
    const int B = 4;  // 4 x 256 threads = 32 warps per SM
    run<0>("VIADDMNMX.S16x2", CH, B);
    run<16>("VIADDMNMX.S16x2.RELU", CH, B);
    run<15>("VIMNMX.S16x2.RELU", CH, B);
    run<2>("VIMNMX3.S16x2", CH, B);
    run<3>("VIADD.16x2", CH, B);
    run<4>("add.s32", CH * 2, B);
    run<5>("IMAD (mad.lo.s32)", CH, B);
    run<6>("PRMT", CH, B);
    run<7>("VIADDMNMX.U16x2", CH, B);
    run<8>("VIMNMX3.U16x2", CH, B);
    run<17>("LOP3", CH, B);
    run<18>("SHF", CH, B);
    run<19>("VIADDMNMX.S32", CH, B);
    run<20>("VIMNMX3.S32", CH, B);
    run<21>("1 VIADDMNMX : 1 IMAD", CH * 2, B);
    run<22>("2 VIADDMNMX : 1 IMAD", CH * 3, B);
    run<12>("LDS.32 (+LOP+IADD)", CH * 3, B);
    run<9>("cell mix now (6.5/pair)", CH * 6 + CH / 2, B);
    run<13>("cell mix now + LDS share", CH * 6 + CH / 2 + CH, B);
    run<10>("biased: sub as IMAD", CH * 6 + CH / 2, B);
    run<11>("biased: both adds as IMAD", CH * 6 + CH / 2, B);
    run<14>("biased: 2 IMAD, no PRMT", CH * 5 + CH / 2, B);
    return 0;
}
