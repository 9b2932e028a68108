// pyopal_b200/csrc/store.h -- internal definitions shared by api.cu and pairdp.cu:
// error plumbing, device buffers and the device-resident store handle
// (the replacement of the per-sequence heap store of
// /root/reference/src/pyopal/lib.pxd:95-98 for the search).
#pragma once
#include "../../include/pyopal_cuda.h"
#include "kernels.cuh"

#include <cuda_runtime.h>
#include <algorithm>
#include <cstdint>
#include <mutex>
#include <string>
#include <vector>

namespace pyopal {

// thread-local message behind pyopal_cuda_last_error() (api.cu)
int fail(int code, const std::string& msg);

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            int code_ = (e_ == cudaErrorMemoryAllocation) ? PYOPAL_CUDA_ERR_NOMEM                  \
                        : (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver)           \
                            ? OPAL_ERR_NO_SIMD_SUPPORT                                             \
                            : PYOPAL_CUDA_ERR_CUDA;                                                \
            return ::pyopal::fail(code_, std::string(#call) + ": " + cudaGetErrorString(e_));      \
        }                                                                                          \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;
    int reserve(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 64;
        CU(cudaMalloc(&p, want * sizeof(T)));
        cap = want;
        return 0;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

constexpr int PAD_ALIGN_COLS = 8;  // pair streams start on 16-byte boundaries

// topk.cu: exact device-side selection of the k best scores (ties to the lower index)
size_t topk_workspace_words();
int topk_device(const int32_t* d_scores, int64_t n, int k, int32_t* d_vals, int32_t* d_idx, uint32_t* ws, int sm_count,
                cudaStream_t st);

}  // namespace pyopal

struct pyopal_cuda_db {
    int device = 0;
    int sm_count = 148;
    int A = 0;
    int64_t n = 0;          // sequences
    int64_t total_len = 0;  // residues
    int64_t npairs = 0;     // pairs of non-empty targets, sorted by len_a descending
    int64_t ncols = 0;      // u16 entries in the stream
    double mean_cols = 0;   // mean len_a
    // host mirrors of the pair tables (worklists, rung eligibility, fix-ups)
    std::vector<int64_t> h_pair_off;
    std::vector<int32_t> h_pair_len, h_pair_idx;
    std::vector<int32_t> h_lengths;   // database order
    std::vector<int32_t> h_zero_idx;  // zero-length targets (closed form on the host)
    // device store
    uint16_t* d_stream = nullptr;
    int64_t* d_pair_off = nullptr;
    int32_t* d_pair_len = nullptr;
    int32_t* d_pair_idx = nullptr;
    int32_t* d_identity = nullptr;  // 0..npairs-1
    // per-search workspace, serialised by `mu`
    std::mutex mu;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    pyopal::DevBuf<int32_t> wl16, wl32, ovf, out, fixes;
    pyopal::DevBuf<uint8_t> query;
    pyopal::DevBuf<int8_t> gprofile;
    pyopal::DevBuf<int32_t> wprofile;
    pyopal::DevBuf<uint32_t> bndH, bndF, bnd32H, bnd32F;
    cudaStream_t side = nullptr, side2 = nullptr;
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr;
    int32_t* d_matrix = nullptr;  // 32*32
    int32_t* d_counts = nullptr;  // [0..2] list sizes, [4..6] task counters
    int64_t stats[12] = {0};
    cudaEvent_t kev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // around k32 / k16 / rescore
    bool kev_valid[3] = {false, false, false};
    // end / full modes (pairdp.cu)
    pyopal::DevBuf<int32_t> pd_i32;     // per-target int results and job tables
    pyopal::DevBuf<uint8_t> pd_tb;      // traceback direction bytes
    pyopal::DevBuf<uint8_t> pd_ops;     // alignment operations
    pyopal::DevBuf<int32_t> pd_bnd;     // pass boundary rows (H, F)
    pyopal::DevBuf<uint8_t> pd_jobs;    // PairJob records
    pyopal::DevBuf<int32_t> pd_team;    // band bookkeeping of pairdp.cu (progress, finished bands, cut, candidates)
    // top-k selection (topk.cu)
    pyopal::DevBuf<uint32_t> tk_ws;
    pyopal::DevBuf<int32_t> tk_vals, tk_idx;
};


// end / full modes (pairdp.cu); called with db->mu held and the device selected
int pyopal_end_full_search(pyopal_cuda_db* db, const unsigned char* query, int Lq, int go, int ge, const int* M,
                           int A, int searchType, int mode, int64_t start, int64_t end, int32_t* scores,
                           int32_t* endQ, int32_t* endT, int32_t* startQ, int32_t* startT, unsigned char* ops,
                           const int64_t* ops_offsets, int32_t* ops_len);
