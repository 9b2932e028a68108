// pyopal_b200/csrc/store.h -- internal definitions shared by api.cu, pairdp.cu and multi.cu:
// error plumbing, device buffers, the device-resident store handle (the
// replacement of the per-sequence heap store of
// /root/reference/src/pyopal/lib.pxd:95-98 for the search) and the per-search
// workspaces that make concurrent searches on one handle safe (the reference's
// Aligner.align is re-entrant on a shared Database, README.md:116-144).
#pragma once
#include "../../include/pyopal_cuda.h"
#include "kernels.cuh"

#include <cuda_runtime.h>
#include <algorithm>
#include <condition_variable>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <utility>
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

// Select `device` for the calling thread and restore the caller's device on scope exit: a process that also
// drives other GPUs (torch, another handle) must not find its current device changed behind its back.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) { prev = -1; cudaGetLastError(); }
        if (prev != device) cudaSetDevice(device);
        else prev = -1;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    DeviceGuard(const DeviceGuard&) = delete;
    DeviceGuard& operator=(const DeviceGuard&) = delete;
};

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

// pinned host memory, grown on demand
template <typename T>
struct PinBuf {
    T* p = nullptr;
    size_t cap = 0;
    int reserve(size_t n) {
        if (n <= cap) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        size_t want = n + n / 4 + 64;
        CU(cudaMallocHost(&p, want * sizeof(T)));
        cap = want;
        return 0;
    }
    void release() {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
};

constexpr int PAD_ALIGN_COLS = 8;  // pair streams start on 16-byte boundaries

// topk.cu: exact device-side selection of the k best scores (ties to the lower index)
size_t topk_workspace_words();
int topk_device(const int32_t* d_scores, int64_t n, int k, int32_t* d_vals, int32_t* d_idx, uint32_t* ws, int sm_count,
                cudaStream_t st);

// Test and tuning switches.  Read ONCE when a handle is created, never on the search path.
struct Tuning {
    bool has_plan = false;
    int plan_logG = 0, plan_K = 0, plan_threads = 0;
    bool no_giant = false, no_tgiant = false, no_tail = false, debug = false, no_end16 = false, no_mid = false;
    bool no_stream = false;   // keep whole-database score searches on search_kernel (one pair per task)
    bool force_dual = false;    // ... or always, whatever the size of the database (tests)
    bool no_dual = false;       // batched searches: one query per pass of the score kernel, never two (dual.cuh)
    bool no_giant16 = false;    // giant.cu at 32 bit, one target per team, also for Smith-Waterman
    bool force_stream = false;  // ... or run them on stream_kernel whatever the tile
    bool has_tgiant = false;
    int tg_warps = 4, tg_K = 32, tg_split = -1;
    double solo = 0.0, gmul = 1.0, mid = 0.0;
    int pair_k = 0, pair_c = 0;
    int lanes = 1;   // searches in flight per batched call
    void load();
};

// Everything one search in flight owns: streams, events, device buffers, the pinned parameter block.  A handle
// keeps a small pool of these so that searches from several host threads overlap instead of serialising.
struct Workspace {
    int device = 0;
    cudaStream_t stream = nullptr;              // non-blocking; blocking entry points run here
    cudaStream_t side = nullptr, side2 = nullptr;  // high priority: the longest targets start first
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;   // device time of a blocking search
    cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_join2 = nullptr;
    cudaEvent_t ev_free = nullptr;              // end of the last enqueue that used this workspace
    cudaEvent_t ev_aux0 = nullptr, ev_aux1 = nullptr;  // fork / join of the end-mode spill launch (pairdp.cu)
    bool used = false;
    cudaEvent_t kev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};  // around k32 / k16 / rescore
    bool kev_valid[3] = {false, false, false};
    // one pinned block per search carries the query, the matrix, the counters and the work lists: one H2D copy.
    // A small ring lets several searches be enqueued back to back without waiting for the copies.
    static constexpr int NPARAM = 4;
    PinBuf<unsigned char> h_par[NPARAM];
    cudaEvent_t par_done[NPARAM] = {nullptr, nullptr, nullptr, nullptr};
    bool par_used[NPARAM] = {false, false, false, false};
    int par_next = 0;
    DevBuf<unsigned char> d_par;
    // score search
    DevBuf<int32_t> ovf, ovf2, out, out2;   // ovf2: second query of a dual search; out2: second row buffer of the stream-ordered batch
    DevBuf<int8_t> gprofile;
    DevBuf<int32_t> wprofile;
    DevBuf<uint32_t> bndH, bndF, bnd32H, bnd32F;
    DevBuf<int32_t> end_out;  // packed end mode: [score | endQ | endT] rows
    int64_t stats[12] = {0};
    int32_t* d_ovf_count = nullptr;  // device: number of 16-bit overflows of the last SW search on this workspace
    // end / full modes (pairdp.cu)
    DevBuf<int32_t> pd_i32;     // per-target int results and job tables
    DevBuf<uint8_t> pd_tb;      // traceback direction bytes
    DevBuf<uint8_t> pd_ops;     // alignment operations
    DevBuf<uint8_t> pd_jobs;    // PairJob records
    DevBuf<int32_t> pd_team;    // band bookkeeping of pairdp.cu (progress, finished bands, cut, candidates)
    DevBuf<uint8_t> pd_query;   // query + matrix of the end / full passes
    // top-k selection (topk.cu)
    DevBuf<uint32_t> tk_ws;
    DevBuf<int32_t> tk_vals, tk_idx, tk_vals2, tk_idx2;

    int init(int device);
    void destroy();
};

}  // namespace pyopal

struct pyopal_cuda_db {
    int device = 0;
    int sm_count = 148;
    int smem_optin = 0;
    int A = 0;
    int64_t n = 0;          // sequences
    int64_t total_len = 0;  // residues
    int64_t npairs = 0;     // pairs of non-empty targets, sorted by len_a descending
    int64_t ncols = 0;      // u16 entries in the stream
    double mean_cols = 0;   // mean len_a
    pyopal::Tuning tune;
    // host mirrors of the pair tables (worklists, rung eligibility, fix-ups)
    std::vector<int64_t> h_pair_off;
    std::vector<int32_t> h_pair_len, h_pair_idx;
    std::vector<int32_t> h_lengths;   // database order
    std::vector<int32_t> h_zero_idx;  // zero-length targets (closed form on the host)
    std::vector<int32_t> h_where;     // database index -> sorted position (2 * pair + half), -1 if empty
    // device store (immutable once built)
    uint16_t* d_stream = nullptr;
    int64_t* d_pair_off = nullptr;
    int32_t* d_pair_len = nullptr;
    int32_t* d_pair_idx = nullptr;
    int32_t* d_identity = nullptr;  // 0..npairs-1
    // workspace pool
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::unique_ptr<pyopal::Workspace>> pool;
    std::vector<pyopal::Workspace*> free_ws;
    int max_ws = 4;
    // statistics of the most recent search (pyopal_cuda_last_stats)
    int64_t last_stats[12] = {0};
    pyopal::Workspace* last_ws = nullptr;

    pyopal::Workspace* acquire();       // nullptr + fail() on error
    pyopal::Workspace* try_acquire();   // nullptr when none is free and the pool is full (never waits)
    void release(pyopal::Workspace* ws);
    void publish(pyopal::Workspace* ws); // make ws's statistics the handle's "last search" without releasing it
};

namespace pyopal {

// RAII lease of a workspace
struct Lease {
    pyopal_cuda_db* db;
    Workspace* ws;
    explicit Lease(pyopal_cuda_db* d) : db(d), ws(d->acquire()) {}
    ~Lease() { if (ws) db->release(ws); }
    Lease(const Lease&) = delete;
    Lease& operator=(const Lease&) = delete;
};

// api.cu: build the store of `device` from n sequences given by pointer (seq(i), lengths[i])
int db_build_from(int device, const unsigned char* residues, const int64_t* offsets, const unsigned char* const* pointers,
                  const int32_t* index, const int32_t* lengths, int64_t n, int A, pyopal_cuda_db** out);

// End-mode plan of a 16-bit search (api.cu): which targets the packed kernel located, which it could not.
struct EndPlan {
    int32_t* d_endq = nullptr;          // in: device rows of (end - start) entries, database order
    int32_t* d_endt = nullptr;
    std::vector<int32_t> spill;         // out: sorted positions (2 * pair + half) left to pairdp.cu (known on the host)
    int32_t* d_ovf = nullptr;           // out: device list of (pair << 1 | half) whose 16-bit score overflowed ...
    int32_t* d_ovf_count = nullptr;     // ... and its device-side length
    bool packed = false;                // out: a 16-bit launch was enqueued
    std::function<int()> launch_spill;  // in: called once `spill` is known and BEFORE the 16-bit launches
};

// Options of enqueue_score_search_ex for its use as one half of a dual search (api.cu, enqueue_dual_search).
struct SearchOpts {
    int plan_Lq = 0;          // in: plan tiles and giants as if the query had this many rows (the longer query of the pair)
    bool plan_only = false;   // in: report the plan below, enqueue nothing, touch nothing
    bool no_bulk = false;     // in: enqueue the giants (and the rescue of their overflows) only
    int bulk_G = 0, bulk_logG = 0, bulk_K = 0;   // out: the bulk tile
    int64_t first_bulk = 0;   // out: first pair of the bulk launch (the pairs before it are the giants)
    bool clean = false;       // out: giants + ONE 16-bit bulk launch over the pairs [first_bulk, npairs), nothing else
};

// api.cu: enqueue a SCORE search of [start, end) on `st` with workspace `ws`; scores land in d_out (device).
int enqueue_score_search(pyopal_cuda_db* db, Workspace* ws, const unsigned char* query, int Lq, int go, int ge,
                         const int* M, int A, int mode, int64_t start, int64_t end, int32_t* d_out, cudaStream_t st);
// the same with `endp` != nullptr: END mode in the 16-bit kernels (score + end cell)
int enqueue_score_search_ex(pyopal_cuda_db* db, Workspace* ws, const unsigned char* query, int Lq, int go, int ge,
                            const int* M, int A, int mode, int64_t start, int64_t end, int32_t* d_out, cudaStream_t st,
                            EndPlan* endp, SearchOpts* opts = nullptr);

// api.cu: two SW score searches of the whole database in one pass (dual.cuh); DUAL_NOT_ELIGIBLE = search them one by one
constexpr int DUAL_NOT_ELIGIBLE = -1;
int enqueue_dual_search(pyopal_cuda_db* db, Workspace* ws, const unsigned char* qa, int La, const unsigned char* qb, int Lb,
                        int go, int ge, const int* M, int A, int32_t* d_out_a, int32_t* d_out_b, cudaStream_t st);

// api.cu: the order and pairing in which the queries of a batch are searched: (first, second or -1) per unit
void plan_batch_units(const pyopal_cuda_db* db, const int32_t* q_lengths, int nq, int mode, std::vector<std::pair<int, int>>& units);
void batch_units(const int32_t* q_lengths, int nq, bool pair_queries, std::vector<std::pair<int, int>>& units);

int check_search_args(pyopal_cuda_db* db, const unsigned char* query, int Lq, int go, int ge, const int* M, int A,
                      int mode, int64_t* start, int64_t* end);

}  // namespace pyopal

// end / full modes (pairdp.cu); called with a leased workspace and the device selected
int pyopal_end_full_search(pyopal_cuda_db* db, pyopal::Workspace* ws, const unsigned char* query, int Lq, int go,
                           int ge, const int* M, int A, int searchType, int mode, int64_t start, int64_t end,
                           int32_t* scores, int32_t* endQ, int32_t* endT, int32_t* startQ, int32_t* startT,
                           unsigned char* ops, const int64_t* ops_offsets, int32_t* ops_len);
