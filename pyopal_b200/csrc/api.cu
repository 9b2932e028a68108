// pyopal_b200/csrc/api.cu -- C-ABI of the sm_100a backend (include/pyopal_cuda.h).
//
// Host side of the path that /root/reference/src/pyopal/platform/pyx.in:57-91
// drives: take the query + the database, run the DP on the device, hand back
// per-target results.  No PyTorch, no CPU fallback: without a CUDA device every
// entry point fails with OPAL_ERR_NO_SIMD_SUPPORT (pyx.in:102-103 maps it to
// RuntimeError, the same class the reference raises when no SIMD backend fits).
#include "store.h"

#include <algorithm>
#include <atomic>
#include <thread>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <numeric>
#include <string>
#include <vector>

using namespace pyopal;

namespace {
thread_local std::string g_err;
}  // namespace

int pyopal::fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}

// ---------------------------------------------------------------------------
extern "C" {

void opalInitSearchResult(OpalSearchResult* r) {
    if (!r) return;
    r->scoreSet = 0;
    r->score = 0;
    r->endLocationTarget = r->endLocationQuery = -1;
    r->startLocationTarget = r->startLocationQuery = -1;
    r->alignment = nullptr;
    r->alignmentLength = 0;
}

int opalSearchResultIsEmpty(const OpalSearchResult r) { return !r.scoreSet; }

void opalSearchResultSetScore(OpalSearchResult* r, int score) {
    r->scoreSet = 1;
    r->score = score;
}

const char* pyopal_cuda_last_error(void) { return g_err.c_str(); }

int pyopal_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int pyopal_cuda_device_info(int device, char* name, int cap, int* sm_count, int64_t* total_mem) {
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    if (name && cap > 0) {
        strncpy(name, prop.name, (size_t)cap - 1);
        name[cap - 1] = 0;
    }
    if (sm_count) *sm_count = prop.multiProcessorCount;
    if (total_mem) *total_mem = (int64_t)prop.totalGlobalMem;
    return 0;
}

void pyopal_cuda_db_destroy(pyopal_cuda_db* db) {
    if (!db) return;
    cudaSetDevice(db->device);
    cudaFree(db->d_stream);
    cudaFree(db->d_pair_off);
    cudaFree(db->d_pair_len);
    cudaFree(db->d_pair_idx);
    cudaFree(db->d_identity);
    cudaFree(db->d_matrix);
    cudaFree(db->d_counts);
    db->wl16.release(); db->wl32.release(); db->ovf.release(); db->out.release(); db->fixes.release();
    db->query.release(); db->gprofile.release(); db->wprofile.release();
    db->bndH.release(); db->bndF.release(); db->bnd32H.release(); db->bnd32F.release();
    db->pd_i32.release(); db->pd_tb.release(); db->pd_ops.release(); db->pd_bnd.release(); db->pd_jobs.release(); db->pd_team.release();
    db->tk_ws.release(); db->tk_vals.release(); db->tk_idx.release();
    if (db->ev_fork) cudaEventDestroy(db->ev_fork);
    if (db->ev_join) cudaEventDestroy(db->ev_join);
    if (db->side) cudaStreamDestroy(db->side);
    if (db->side2) cudaStreamDestroy(db->side2);
    if (db->ev_join2) cudaEventDestroy(db->ev_join2);
    for (int i = 0; i < 6; i++) if (db->kev[i]) cudaEventDestroy(db->kev[i]);
    if (db->ev0) cudaEventDestroy(db->ev0);
    if (db->ev1) cudaEventDestroy(db->ev1);
    if (db->stream) cudaStreamDestroy(db->stream);
    delete db;
}

int64_t pyopal_cuda_db_size(const pyopal_cuda_db* db) { return db ? db->n : 0; }
int64_t pyopal_cuda_db_total_length(const pyopal_cuda_db* db) { return db ? db->total_len : 0; }
int64_t pyopal_cuda_db_device_bytes(const pyopal_cuda_db* db) {
    if (!db) return 0;
    return db->ncols * 2 + db->npairs * (8 + 8 + 8 + 4);
}

int pyopal_cuda_last_stats(const pyopal_cuda_db* db, int64_t* stats, int n) {
    if (!db || !stats) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    // resolve the per-kernel events of the last search if they have completed
    pyopal_cuda_db* m = const_cast<pyopal_cuda_db*>(db);
    for (int k = 0; k < 3; k++) {
        m->stats[8 + k] = 0;
        if (m->kev_valid[k] && cudaEventQuery(m->kev[2 * k + 1]) == cudaSuccess) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, m->kev[2 * k], m->kev[2 * k + 1]) == cudaSuccess)
                m->stats[8 + k] = (int64_t)(ms * 1000.0f + 0.5f);
        }
    }
    cudaGetLastError();
    for (int i = 0; i < n && i < 12; i++) stats[i] = db->stats[i];
    return 0;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Store construction: sort by length (descending, ties by index), pair
// neighbours, interleave each pair's residues into one u16 stream.
template <typename GetSeq>
static int db_build(int device, GetSeq get_seq, const int32_t* lengths, int64_t n, int A,
                    pyopal_cuda_db** out) {
    if (!out) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output handle");
    *out = nullptr;
    if (n < 0 || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "database size out of range");
    if (A < 1 || A > 32) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "alphabet length must be in 1..32");
    if (pyopal_cuda_device_count() <= 0) return fail(OPAL_ERR_NO_SIMD_SUPPORT, "no CUDA device available");
    CU(cudaSetDevice(device));
    pyopal_cuda_db* db = new (std::nothrow) pyopal_cuda_db();
    if (!db) return fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed");
    struct Guard {
        pyopal_cuda_db* d;
        ~Guard() { if (d) pyopal_cuda_db_destroy(d); }
    } guard{db};
    db->device = device;
    db->A = A;
    db->n = n;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    db->sm_count = prop.multiProcessorCount;

    db->h_lengths.assign(lengths, lengths + n);
    std::vector<int32_t> order;
    order.reserve((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        if (lengths[i] < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative sequence length");
        db->total_len += lengths[i];
        if (lengths[i] == 0) db->h_zero_idx.push_back((int32_t)i);
        else order.push_back((int32_t)i);
    }
    std::stable_sort(order.begin(), order.end(),
                     [&](int32_t a, int32_t b) { return lengths[a] > lengths[b]; });
    const int64_t m = (int64_t)order.size();
    const int64_t npairs = (m + 1) / 2;
    db->npairs = npairs;
    db->h_pair_off.resize((size_t)npairs);
    db->h_pair_len.resize((size_t)npairs * 2);
    db->h_pair_idx.resize((size_t)npairs * 2);
    int64_t ncols = 0;
    for (int64_t k = 0; k < npairs; k++) {
        int32_t a = order[(size_t)(2 * k)];
        int32_t b = (2 * k + 1 < m) ? order[(size_t)(2 * k + 1)] : -1;
        db->h_pair_off[(size_t)k] = ncols;
        db->h_pair_len[(size_t)(2 * k)] = lengths[a];
        db->h_pair_len[(size_t)(2 * k + 1)] = b >= 0 ? lengths[b] : 0;
        db->h_pair_idx[(size_t)(2 * k)] = a;
        db->h_pair_idx[(size_t)(2 * k + 1)] = b;
        ncols += ((int64_t)lengths[a] + PAD_ALIGN_COLS - 1) / PAD_ALIGN_COLS * PAD_ALIGN_COLS;
    }
    db->ncols = ncols;
    db->mean_cols = npairs ? (double)ncols / (double)npairs : 0.0;

    // interleave on the host into a pinned staging buffer, chunk by chunk
    if (ncols > 0) {
        CU(cudaMalloc(&db->d_stream, (size_t)ncols * sizeof(uint16_t)));
        const size_t CHUNK = (size_t)16 << 20;  // columns per staging chunk
        uint16_t* stage = nullptr;
        size_t stage_cols = std::min<size_t>(CHUNK, (size_t)ncols);
        // a pair never straddles chunks: grow the staging buffer to the longest pair
        int64_t longest = ((int64_t)db->h_pair_len[0] + PAD_ALIGN_COLS - 1) / PAD_ALIGN_COLS * PAD_ALIGN_COLS;
        stage_cols = std::max<size_t>(stage_cols, (size_t)longest);
        CU(cudaMallocHost(&stage, stage_cols * sizeof(uint16_t)));
        const uint16_t PADPAIR = (uint16_t)(A | (A << 8));
        int64_t k = 0;
        int rc = 0;
        while (k < npairs && rc == 0) {
            const int64_t base = db->h_pair_off[(size_t)k];
            int64_t k2 = k;
            while (k2 < npairs) {
                int64_t endc = (k2 + 1 < npairs) ? db->h_pair_off[(size_t)(k2 + 1)] : ncols;
                if ((size_t)(endc - base) > stage_cols) break;
                k2++;
            }
            const int64_t endc = (k2 < npairs) ? db->h_pair_off[(size_t)k2] : ncols;
            // interleave the pairs of this chunk on several host threads (205 MB of residues for config 2)
            std::atomic<unsigned> bad_any{0};
            auto fill = [&](int64_t q0, int64_t q1) {
                unsigned bad = 0;  // residues must index the profile: < A
                for (int64_t q = q0; q < q1; q++) {
                    uint16_t* dst = stage + (db->h_pair_off[(size_t)q] - base);
                    const int la = db->h_pair_len[(size_t)(2 * q)], lb = db->h_pair_len[(size_t)(2 * q + 1)];
                    const unsigned char* sa = get_seq(db->h_pair_idx[(size_t)(2 * q)]);
                    const unsigned char* sb = lb > 0 ? get_seq(db->h_pair_idx[(size_t)(2 * q + 1)]) : nullptr;
                    int c = 0;
                    for (; c < lb; c++) {
                        bad |= (unsigned)(sa[c] >= A) | (unsigned)(sb[c] >= A);
                        dst[c] = (uint16_t)(sa[c] | (sb[c] << 8));
                    }
                    for (; c < la; c++) {
                        bad |= (unsigned)(sa[c] >= A);
                        dst[c] = (uint16_t)(sa[c] | (A << 8));
                    }
                    const int padded = (la + PAD_ALIGN_COLS - 1) / PAD_ALIGN_COLS * PAD_ALIGN_COLS;
                    for (; c < padded; c++) dst[c] = PADPAIR;
                }
                if (bad) bad_any.store(1);
            };
            {
                const int64_t npq = k2 - k;
                unsigned nth = std::min<unsigned>(16u, std::max(1u, std::thread::hardware_concurrency()));
                if ((endc - base) < ((int64_t)1 << 20)) nth = 1;
                if (nth <= 1) {
                    fill(k, k2);
                } else {
                    // pairs are sorted by length: deal contiguous ranges of about equal column count
                    std::vector<std::thread> pool;
                    int64_t q0 = k;
                    for (unsigned t = 0; t < nth && q0 < k2; t++) {
                        const int64_t want = base + (endc - base) * (int64_t)(t + 1) / nth;
                        int64_t q1 = (t + 1 == nth) ? k2
                                                    : std::upper_bound(db->h_pair_off.begin() + q0, db->h_pair_off.begin() + k2, want) -
                                                          db->h_pair_off.begin();
                        q1 = std::max(q1, q0 + (t + 1 == nth ? 0 : 1));
                        q1 = std::min(q1, k2);
                        pool.emplace_back(fill, q0, q1);
                        q0 = q1;
                    }
                    if (q0 < k2) fill(q0, k2);
                    for (auto& th : pool) th.join();
                }
                (void)npq;
            }
            if (bad_any.load()) rc = fail(PYOPAL_CUDA_ERR_ARGUMENT, "residue code outside the alphabet");
            cudaError_t e = cudaMemcpy(db->d_stream + base, stage, (size_t)(endc - base) * sizeof(uint16_t),
                                       cudaMemcpyHostToDevice);
            if (e != cudaSuccess) rc = fail(PYOPAL_CUDA_ERR_CUDA, std::string("stream upload: ") + cudaGetErrorString(e));
            k = k2;
        }
        cudaFreeHost(stage);
        if (rc) return rc;
    }
    if (npairs > 0) {
        CU(cudaMalloc(&db->d_pair_off, (size_t)npairs * sizeof(int64_t)));
        CU(cudaMalloc(&db->d_pair_len, (size_t)npairs * 2 * sizeof(int32_t)));
        CU(cudaMalloc(&db->d_pair_idx, (size_t)npairs * 2 * sizeof(int32_t)));
        CU(cudaMalloc(&db->d_identity, (size_t)npairs * sizeof(int32_t)));
        CU(cudaMemcpy(db->d_pair_off, db->h_pair_off.data(), (size_t)npairs * sizeof(int64_t), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(db->d_pair_len, db->h_pair_len.data(), (size_t)npairs * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(db->d_pair_idx, db->h_pair_idx.data(), (size_t)npairs * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        std::vector<int32_t> ident((size_t)npairs);
        std::iota(ident.begin(), ident.end(), 0);
        CU(cudaMemcpy(db->d_identity, ident.data(), (size_t)npairs * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    CU(cudaMalloc(&db->d_matrix, 32 * 32 * sizeof(int32_t)));
    CU(cudaMalloc(&db->d_counts, 16 * sizeof(int32_t)));
    CU(cudaStreamCreateWithFlags(&db->stream, cudaStreamNonBlocking));
    {
        // the side stream carries the longest targets: highest priority so its CTAs are placed first
        int lo_prio = 0, hi_prio = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
        CU(cudaStreamCreateWithPriority(&db->side, cudaStreamNonBlocking, hi_prio));
        CU(cudaStreamCreateWithPriority(&db->side2, cudaStreamNonBlocking, hi_prio));
    }
    CU(cudaEventCreateWithFlags(&db->ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&db->ev_join, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&db->ev_join2, cudaEventDisableTiming));
    CU(cudaEventCreate(&db->ev0));
    CU(cudaEventCreate(&db->ev1));
    for (int i = 0; i < 6; i++) CU(cudaEventCreate(&db->kev[i]));
    guard.d = nullptr;
    *out = db;
    return 0;
}

extern "C" int pyopal_cuda_db_create(int device, const unsigned char* residues, const int64_t* offsets,
                                     const int32_t* lengths, int64_t n, int alphabetLength,
                                     pyopal_cuda_db** out) {
    if (n > 0 && (!residues || !offsets || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    return db_build(device, [&](int32_t i) { return residues + offsets[i]; }, lengths, n, alphabetLength, out);
}

extern "C" int pyopal_cuda_db_create_from_pointers(int device, const unsigned char* const* sequences,
                                                   const int32_t* lengths, int64_t n, int alphabetLength,
                                                   pyopal_cuda_db** out) {
    if (n > 0 && (!sequences || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    return db_build(device, [&](int32_t i) { return sequences[i]; }, lengths, n, alphabetLength, out);
}

// ---------------------------------------------------------------------------
// Search planning

namespace {

struct Plan {
    int G = 32, logG = 5, K = 32, npass = 1, pitch = 0, padtop = 0, kstride = 32;
    bool psmem = true;
    size_t smem = 0;
};

Plan finish_plan(Plan p, int Lq, int A, size_t smem_limit) {
    const int T = p.G * p.K;
    p.npass = (Lq + T - 1) / T;
    p.kstride = ((p.K / 4) % 2 == 0) ? p.K + 4 : p.K;  // odd number of words per lane chunk (kernels.cuh)
    // Letter rows start 16 (mod 32) bytes apart, i.e. 4 (mod 8) words: the rows of the letters then start in 8
    // different bank quads while lanes p = 0..3 of a group (lane chunks an odd number of words apart) stay in
    // different residues mod 4, so only lanes with the same p in different groups can collide.  Measured on
    // config 2: a pitch that is a multiple of 128 bytes is 28 % slower, an odd number of words 3-4 % slower.
    p.pitch = (p.npass * p.G * p.kstride + 15) / 32 * 32 + 16;
    p.padtop = p.npass * T - Lq;
    p.smem = (size_t)(A + 1) * (size_t)p.pitch;
    p.psmem = p.smem <= smem_limit;
    if (!p.psmem) p.smem = 0;
    return p;
}

// Pick the tile (G threads x K rows) that minimises issue slots for this query
// length and the store's mean target length.  Per column step a thread issues
// about C0 instructions of bookkeeping plus C1 per row; a target of L columns
// takes L + G - 1 steps per pass.
//
// Critical path: a pair of L columns is walked by ONE group of G lanes, pass
// after pass, while about 448 / G groups per SM share the machine (14 resident
// warps).  Its solo time equals the whole search's when
//     L = L_even(G) = ncols * G / (448 * SMs)          (independent of K)
// so a small shard (many GPUs) or a long query needs wide groups.  Pairs
// longer than 0.6 * L_even go to the giant kernel when it is available (at
// most `giant_cap` of them); what remains above L_even stretches the search
// by len / L_even and is charged to the tile.
struct GiantPolicy {
    bool ok = false;        // giant kernel usable for this query
    int64_t cap = 0;        // most pairs it may take
    double col_cost = 0;    // SM time of one giant target column, in the tile cost model's units
    int64_t kg = 0;         // out: pairs [0, kg) go to the giant kernel
};

Plan make_plan(int Lq, int A, const pyopal_cuda_db* db, size_t smem_limit, GiantPolicy* gp) {
    const double C0 = 28.0, C1 = 6.5;
    const double mean_cols = db->mean_cols;
    Plan best;
    double best_cost = 1e300;
    int64_t best_kg = 0;
    for (int logG = 2; logG <= 5; logG++) {
        const int G = 1 << logG;
        // calibrated on 1/2, 1/4 and 1/8 shards of config 2 (scratch/shard_sweep.py): groups of 8+ lanes reach the
        // critical path about 25 % earlier than 448 / G groups per SM suggest; 4-lane groups do not
        static const double solo_env = getenv("PYOPAL_CUDA_SOLO") ? atof(getenv("PYOPAL_CUDA_SOLO")) : 0.0;
        const double solo = solo_env > 0 ? solo_env : (G == 4 ? 448.0 : 560.0);
        const double L_even = (double)db->ncols * G / (solo * db->sm_count);
        int64_t kg = 0;
        double giant_cols = 0;
        if (gp && gp->ok) {
            const double gthr = std::max(1024.0, 0.6 * L_even);
            while (kg < gp->cap && kg < db->npairs && (double)db->h_pair_len[(size_t)(2 * kg)] > gthr) {
                giant_cols += (double)db->h_pair_len[(size_t)(2 * kg)] + (double)db->h_pair_len[(size_t)(2 * kg + 1)];
                kg++;
            }
        }
        // the giants hold whole SMs at a fraction of the bulk kernel's efficiency
        const double giant_cost = giant_cols * (gp ? gp->col_cost : 0.0) / std::max<double>(1.0, (double)db->npairs);
        const double longest_bulk = kg < db->npairs ? (double)db->h_pair_len[(size_t)(2 * kg)] : 0.0;
        const double stretch = std::max(1.0, longest_bulk / std::max(L_even, 1.0));
        for (int K = 4; K <= 40; K += 4) {
            const int T = G * K;
            const int npass = (Lq + T - 1) / T;
            // several passes also pay the boundary round trip through L2 every step
            const double cost = stretch * (double)npass * (mean_cols + G - 1) * (C0 + (npass > 1 ? 12.0 : 0.0) + C1 * K) * (double)G +
                                giant_cost;
            if (cost < best_cost - 1e-9) {
                best_cost = cost;
                best.G = G; best.logG = logG; best.K = K; best.npass = npass;
                best_kg = kg;
            }
        }
    }
    if (const char* ov = getenv("PYOPAL_CUDA_PLAN")) {  // "logG,K[,threads]": tuning override
        int lg = 0, k = 0;
        if (sscanf(ov, "%d,%d", &lg, &k) == 2 && lg >= 2 && lg <= 5 && k >= 4 && k <= 40 && k % 4 == 0) {
            best.G = 1 << lg; best.logG = lg; best.K = k; best.npass = (Lq + best.G * k - 1) / (best.G * k);
        }
    }
    if (gp) gp->kg = best_kg;
    best = finish_plan(best, Lq, A, smem_limit);
    return best;
}

Plan make_forced_plan(int Lq) {
    Plan p; p.G = 32; p.logG = 5; p.K = 32;
    p = finish_plan(p, Lq, 0, 0);
    p.psmem = false; p.smem = 0;
    return p;
}
// Latency tile for the longest targets: a whole warp per pair and as few rows
// per thread as one pass allows (the per-step critical path is ~12 cycles per
// row); with several passes, deep tiles amortise the per-step latency better.
Plan make_tail_plan(int Lq, int A, size_t smem_limit) {
    Plan p; p.G = 32; p.logG = 5;
    p.K = std::min(40, std::max(4, (Lq + 127) / 128 * 4));
    return finish_plan(p, Lq, A, smem_limit);
}

struct Rung {
    bool fits8 = false;       // folded scores fit the int8 profile
    bool ok16 = false;        // the 16-bit rung may be used at all
    int64_t max_cols16 = 0;   // NW/HW/OV: longest pair (columns) the 16-bit rung is exact for
    int neg16 = 0, ovf_thresh = 0, pad_score = 0;
    int bias16 = 0;           // 16-bit rung: stored value = true value + bias16 (kernels.cuh)
    bool tdadd = false;       // every folded profile entry >= 0: Hdiag + S may be a 32-bit add
};

Rung make_rung(const int* M, int A, int go, int ge, int mode, int Lp) {
    Rung r;
    int smax = INT_MIN, smin = INT_MAX;
    for (int i = 0; i < A * A; i++) { smax = std::max(smax, M[i]); smin = std::min(smin, M[i]); }
    // padding rows / letter: 0 keeps a free top border at 0; -max(open, ext)
    // reproduces NW's penalised border (kernels.cuh, profile_entry)
    const int pad_raw = (mode == MODE_NW) ? -std::max(go, ge) : 0;
    r.pad_score = pad_raw + go;
    const int lo = std::min(smin, pad_raw) + go, hi = std::max(smax, pad_raw) + go;
    r.fits8 = (lo >= -128 && hi <= 127);
    r.tdadd = (lo >= 0);
    // Bias of the u16 halves.  Local alignment never goes below -(open + ext), so a small bias
    // leaves almost the whole 16-bit range to the score; the other modes are symmetric.
    r.bias16 = (mode == MODE_SW) ? 2 * go + ge + 256 : 32768;
    // stored as max(open, ext): neither (E - ext) nor (F - open) wraps below 0 (kernels.cuh, sweep 2)
    r.neg16 = std::max(go, ge) - r.bias16;
    r.ovf_thresh = 65535 - 128 - 1 - r.bias16;     // SW: a running max >= this is rescored at 32 bit
    const int64_t slack = 512;  // covers S + open <= 127 twice (M - open and F - open are formed below H - open)
    // the 16-bit kernel's short F chain needs ext <= open
    r.ok16 = r.fits8 && (3LL * go + 2LL * ge + slack < 16000) && ge <= go;
    if (r.ok16 && mode != MODE_SW) {
        // lower bound of any live value: -(3*open + (Lp + cols + 2)*ext + 128) must stay above neg16
        int64_t cap = INT32_MAX;
        if (ge > 0) cap = (32768LL - ge - slack - 3LL * go) / ge - Lp - 2;
        // upper bound: min(Lp, cols) * smax + open + 128 must stay below 32767
        if (smax > 0) {
            const int64_t lim = (32767LL - slack - go) / smax;
            if (Lp > lim) cap = std::min(cap, lim);
        }
        r.max_cols16 = std::max<int64_t>(cap, 0);
    } else {
        r.max_cols16 = INT32_MAX;
    }
    return r;
}

// closed forms for empty query / empty target (oracle/opal_oracle.c dp_end, m == 0 || n == 0)
int degenerate_score(int mode, int m, int n, int go, int ge) {
    if (mode == MODE_NW) {
        if (m == 0 && n == 0) return 0;
        return -(go + (std::max(m, n) - 1) * ge);
    }
    if (mode == MODE_HW) return (m == 0) ? 0 : -(go + (m - 1) * ge);
    return 0;
}

__global__ void scatter_fixes_kernel(const int32_t* fixes, int nfix, int32_t* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nfix) out[fixes[2 * i]] = fixes[2 * i + 1];
}

search_kernel_t pick_kernel(int K, int W, bool local, bool psmem, bool tdadd = true) {
    if (W == 16) return local ? get_search_kernel_w16_l1(K, psmem, tdadd) : get_search_kernel_w16_l0(K, psmem, tdadd);
    return local ? get_search_kernel_w32_l1(K, psmem) : get_search_kernel_w32_l0(K, psmem);
}

int launch_search(pyopal_cuda_db* db, search_kernel_t fn, const KParams& P, size_t smem, int threads,
                  int64_t max_tasks, cudaStream_t st) {
    CU(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem));
    if (per_sm < 1) return fail(PYOPAL_CUDA_ERR_INTERNAL, "kernel does not fit on an SM");
    int64_t grid = (int64_t)db->sm_count * per_sm;
    const int64_t need = (max_tasks + (threads / 32) - 1) / (threads / 32);
    grid = std::max<int64_t>(1, std::min(grid, need));
    fn<<<(unsigned)grid, threads, smem, st>>>(P);
    CU(cudaGetLastError());
    db->stats[0]++;
    return 0;
}

void apply_plan(KParams& P, const Plan& pl) {
    P.G = pl.G; P.logG = pl.logG; P.npass = pl.npass; P.pitch = pl.pitch; P.padtop = pl.padtop;
    P.krows = pl.K; P.kstride = pl.kstride;
}

// Enqueue a SCORE search of [start, end) on `st`; scores land in d_out (device).
//
// Launches (all optional but the bulk):
//   tgiants: targets longer than one slab of tgiant.cu when the query is short enough for it, 32 bit, one CTA
//            per target, on the second high-priority stream
//   giants : the other "long" pairs chosen with the tile (make_plan), giant.cu, 32 bit, high-priority side stream
//   wide32 : 32-bit rung for pairs the 16-bit rung cannot hold exactly (NW/HW/OV static bound, or gap costs /
//            scores outside its range: then every pair), latency tile, side stream
//   tail   : 16-bit rung for the longest pairs, one warp per pair, only when no giant kernel can take the query
//   bulk   : 16-bit rung, throughput tile (make_plan), persistent 224/480-thread CTAs, on `st`
//   rescue : 32-bit rescore of SW targets whose 16-bit running max overflowed; its work list is produced on the
//            device by bulk/tail.
int enqueue_score_search(pyopal_cuda_db* db, const unsigned char* query, int Lq, int go, int ge,
                         const int* M, int A, int mode, int64_t start, int64_t end, int32_t* d_out,
                         cudaStream_t st) {
    for (int i = 0; i < 12; i++) db->stats[i] = 0;
    db->kev_valid[0] = db->kev_valid[1] = db->kev_valid[2] = false;
    const int64_t nout = end - start;
    if (nout <= 0) return 0;
    const bool local = (mode == MODE_SW);

    // host-side fix-ups: zero-length targets, or everything when the query is empty
    std::vector<int32_t> fixes;
    if (Lq == 0) {
        for (int64_t i = start; i < end; i++) {
            fixes.push_back((int32_t)(i - start));
            fixes.push_back(degenerate_score(mode, 0, db->h_lengths[(size_t)i], go, ge));
        }
    } else {
        for (int32_t z : db->h_zero_idx)
            if (z >= start && z < end) {
                fixes.push_back((int32_t)(z - start));
                fixes.push_back(degenerate_score(mode, Lq, 0, go, ge));
            }
    }
    auto apply_fixes = [&]() -> int {
        if (fixes.empty()) return 0;
        const int nfix = (int)(fixes.size() / 2);
        if (db->fixes.reserve(fixes.size())) return PYOPAL_CUDA_ERR_NOMEM;
        CU(cudaMemcpyAsync(db->fixes.p, fixes.data(), fixes.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        scatter_fixes_kernel<<<(nfix + 255) / 256, 256, 0, st>>>(db->fixes.p, nfix, d_out);
        CU(cudaGetLastError());
        return 0;
    };
    if (Lq == 0 || db->npairs == 0) return apply_fixes();

    // ---- plans -------------------------------------------------------------
    const size_t SMEM_LIMIT = (size_t)200 * 1024;
    // giants: the longest targets get a team of warps each (giant.cu), 32 bit; needs the int8
    // profile of the whole query in shared memory
    GiantPolicy gpol;
    int gK = 4, gwarps = 1, gteams = 1, gpitch = 0;
    size_t gsmem = 0;
    // gap_extend > gap_open: NW's padding rows need the border fix of the 32-bit kernel (kernels.cuh, BFIX); the
    // giant kernels do not carry it, so everything stays on search_kernel then
    const bool bfix = (mode == MODE_NW && ge > go);
    if (make_rung(M, A, go, ge, mode, 0).fits8 && Lq <= 8192 && !bfix && !getenv("PYOPAL_CUDA_NO_GIANT")) {
        gK = (Lq <= 2048) ? 4 : (Lq <= 4096) ? 8 : 16;            // rows per lane
        gwarps = ((Lq + gK - 1) / gK + 31) / 32;                  // warps per team (<= 16): the whole query in one pass
        gteams = (gwarps == 1) ? 16 : std::min(15, 16 / gwarps);  // teams per 512-thread CTA (15 named barriers)
        gpitch = (gwarps * 32 * (gK + 4) + 15) / 16 * 16;         // gK = 4, 8, 16 words-even: chunks of gK + 4 bytes
        gsmem = giant_smem_bytes(A, gpitch, gwarps, gteams);
        gpol.ok = gsmem <= SMEM_LIMIT + 24 * 1024;
        gpol.cap = (int64_t)db->sm_count * 8;
        // one SM clock ~ 56 cost units (calibrated on config 2); a team advances one column in about
        // 36 clocks per row held by a lane, 1.6 x that when the wavefront crosses warps
        const int teams_now = (gwarps == 1) ? 16 : std::max(1, 16 / gwarps);
        static const double gmul = getenv("PYOPAL_CUDA_GMUL") ? atof(getenv("PYOPAL_CUDA_GMUL")) : 1.0;
        gpol.col_cost = gmul * 56.0 * 36.0 * gK * (gwarps > 1 ? 1.6 : 1.0) / teams_now;
    }
    const bool old_giants_ok = gpol.ok;
    // transposed giants (tgiant.cu): the target over the lanes, the query streamed; needs the short F chain
    bool use_t = false;
    int tK = 32, tNW = 4, tpitch = 0, tsplit_override = -1;
    size_t tsmem = 0;
    int smem_optin = 0;
    CU(cudaDeviceGetAttribute(&smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, db->device));
    if (make_rung(M, A, go, ge, mode, 0).fits8 && ge <= go && !getenv("PYOPAL_CUDA_NO_TGIANT") &&
        !getenv("PYOPAL_CUDA_NO_GIANT")) {
        if (const char* ov = getenv("PYOPAL_CUDA_TGIANT")) {  // "warps,K": tuning override
            int a = 0, b = 0, c = -1;
            const int got = sscanf(ov, "%d,%d,%d", &a, &b, &c);
            if (got >= 2 && a >= 1 && a <= 8 && (b == 16 || b == 32)) { tNW = a; tK = b; }
            if (got == 3 && c >= 0) tsplit_override = c;
        }
        tpitch = tNW * 32 * (tK + 4);
        tsmem = tgiant_smem_bytes(A, tpitch, tNW, Lq);
        // one warp per scheduler issues a 32-row step in ~820 clocks: faster than giant.cu up to ~1500 rows
        use_t = tsmem + 1024 <= (size_t)smem_optin && (Lq <= 1536 || getenv("PYOPAL_CUDA_TGIANT"));
        if (use_t && !old_giants_ok) {
            gpol.ok = true;
            gpol.cap = (int64_t)db->sm_count * 8;
            // a revolution of Lq + 32 * warps - 1 steps covers 32 * warps * K target columns; a step costs
            // about 25 clocks per row held by a lane (measured); 56 cost units per SM clock
            const double step_clk = 25.0 * tK + 20.0;
            gpol.col_cost = 56.0 * step_clk * (double)(Lq + tNW * 32 - 1) / (double)(tNW * 32 * tK);
        }
    }
    Plan bulk = make_plan(Lq, A, db, SMEM_LIMIT, &gpol);
    Plan tail = make_tail_plan(Lq, A, SMEM_LIMIT);
    int smax = INT_MIN, smin = INT_MAX;
    for (int i = 0; i < A * A; i++) { smax = std::max(smax, M[i]); smin = std::min(smin, M[i]); }
    const int Lp = std::max(bulk.npass * bulk.G * bulk.K, tail.npass * tail.G * tail.K);
    Rung rung = make_rung(M, A, go, ge, mode, Lp);
    const bool wide = !rung.fits8;
    if (wide || bfix || !bulk.psmem || !tail.psmem) {
        // profile in global memory: one tile shape (32 x 32) for every launch
        bulk = make_forced_plan(Lq);
        tail = bulk;
    }
    {
        // 32-bit overflow cannot happen below this bound (scores are exact otherwise)
        int64_t mag = std::max<int64_t>({(int64_t)go, (int64_t)ge, 1, std::llabs((long long)smax), std::llabs((long long)smin)});
        const int64_t longest = db->h_pair_len[0];
        if (mag * ((int64_t)Lp + longest + 8) >= (1LL << 29))
            return fail(OPAL_ERR_OVERFLOW, "scores may exceed the 32-bit range");
    }
    const bool same_plan = (bulk.G == tail.G && bulk.K == tail.K && bulk.npass == tail.npass) ||
                           getenv("PYOPAL_CUDA_NO_TAIL") != nullptr;
    if (getenv("PYOPAL_CUDA_NO_TAIL")) tail = bulk;

    // ---- split the (length-sorted) pairs ------------------------------------
    // [0, n32p): 32-bit static; [n32p, n32p + ntailp): 16-bit tail; rest: 16-bit bulk
    int64_t n32p = 0;
    if (!rung.ok16) n32p = db->npairs;
    else if (!local) {
        int64_t lo = 0, hi = db->npairs;
        while (lo < hi) {
            int64_t mid = (lo + hi) / 2;
            if (db->h_pair_len[(size_t)(2 * mid)] > rung.max_cols16) lo = mid + 1; else hi = mid;
        }
        n32p = lo;
    }
    int64_t tail_end = n32p;
    // a pair leaves the bulk kernel when walking it alone in the bulk tile would
    // take more than a quarter of the whole search at full machine throughput
    double thr = 256.0;
    {
        const double C0 = 28.0, C1 = 6.5;
        const double total_instr = (double)db->ncols * bulk.npass * bulk.G * (C0 + C1 * bulk.K);
        const double ideal_cycles = total_instr / ((double)db->sm_count * 64.0);
        const double step_latency = 12.0 * bulk.K + 80.0;
        thr = std::max(256.0, ideal_cycles / 4.0 / (bulk.npass * step_latency));
    }
    // ---- giants: pairs [0, kg), chosen with the tile (make_plan) ----
    const bool giants_ok = gpol.ok && bulk.psmem && !wide;
    const int64_t kg = giants_ok ? gpol.kg : 0;
    n32p = std::max(n32p, kg);
    tail_end = n32p;
    // the one-warp-per-pair tail kernel only remains for queries the giant kernel cannot take
    if (!same_plan && !giants_ok) {
        int64_t lo = n32p, hi = db->npairs;
        while (lo < hi) {
            int64_t mid = (lo + hi) / 2;
            if ((double)db->h_pair_len[(size_t)(2 * mid)] > thr) lo = mid + 1; else hi = mid;
        }
        tail_end = lo;
    }
    const bool whole = (start == 0 && end >= db->n);
    auto in_range = [&](int32_t idx) { return idx >= start && idx < end; };
    auto pair_in_range = [&](int64_t k) {
        return in_range(db->h_pair_idx[(size_t)(2 * k)]) ||
               (db->h_pair_idx[(size_t)(2 * k + 1)] >= 0 && in_range(db->h_pair_idx[(size_t)(2 * k + 1)]));
    };
    std::vector<int32_t> wl32, wl16, wlg, wlt;
    // giants longer than one slab of the transposed kernel go to tgiant.cu, the others to giant.cu
    const int tsplit = !use_t ? INT32_MAX : tsplit_override >= 0 ? tsplit_override : tNW * 32 * tK;
    for (int64_t k = 0; k < kg; k++) {
        for (int h = 0; h < 2; h++) {
            const int32_t ix = db->h_pair_idx[(size_t)(2 * k + h)];
            if (ix < 0 || !in_range(ix)) continue;
            const int32_t e = (int32_t)((k << 1) | h);
            if (db->h_pair_len[(size_t)(2 * k + h)] > tsplit || !old_giants_ok) wlt.push_back(e);
            else wlg.push_back(e);
        }
    }
    for (int64_t k = kg; k < n32p; k++) {
        if (in_range(db->h_pair_idx[(size_t)(2 * k)])) wl32.push_back((int32_t)(k << 1));
        if (db->h_pair_idx[(size_t)(2 * k + 1)] >= 0 && in_range(db->h_pair_idx[(size_t)(2 * k + 1)]))
            wl32.push_back((int32_t)((k << 1) | 1));
    }
    const int32_t* d_wl_tail = nullptr;
    const int32_t* d_wl_bulk = nullptr;
    int64_t ntail = 0, nbulk = 0;
    if (whole) {
        d_wl_tail = db->d_identity + n32p; ntail = tail_end - n32p;
        d_wl_bulk = db->d_identity + tail_end; nbulk = db->npairs - tail_end;
    } else {
        for (int64_t k = n32p; k < tail_end; k++) if (pair_in_range(k)) wl16.push_back((int32_t)k);
        ntail = (int64_t)wl16.size();
        for (int64_t k = tail_end; k < db->npairs; k++) if (pair_in_range(k)) wl16.push_back((int32_t)k);
        nbulk = (int64_t)wl16.size() - ntail;
        if (db->wl16.reserve(wl16.size() + 1)) return PYOPAL_CUDA_ERR_NOMEM;
        if (!wl16.empty())
            CU(cudaMemcpyAsync(db->wl16.p, wl16.data(), wl16.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        d_wl_tail = db->wl16.p; d_wl_bulk = db->wl16.p + ntail;
    }
    const int64_t n32 = (int64_t)wl32.size();
    const int64_t ng = (int64_t)wlg.size(), nt = (int64_t)wlt.size();
    wl32.insert(wl32.end(), wlg.begin(), wlg.end());  // one upload: [static 32-bit | giants | transposed giants]
    wl32.insert(wl32.end(), wlt.begin(), wlt.end());
    if (db->wl32.reserve(wl32.size() + 1)) return PYOPAL_CUDA_ERR_NOMEM;
    if (!wl32.empty()) CU(cudaMemcpyAsync(db->wl32.p, wl32.data(), wl32.size() * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    const int64_t n16 = ntail + nbulk;
    if (same_plan) { d_wl_bulk = d_wl_tail; nbulk = n16; ntail = 0; }
    if (local && n16 && db->ovf.reserve((size_t)(2 * n16))) return PYOPAL_CUDA_ERR_NOMEM;

    // ---- query, matrix, counters, buffers ------------------------------------
    if (db->query.reserve((size_t)Lq)) return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaMemcpyAsync(db->query.p, query, (size_t)Lq, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(db->d_matrix, M, (size_t)A * A * sizeof(int32_t), cudaMemcpyHostToDevice, st));
    int32_t counts[16] = {0};
    counts[0] = (int32_t)nbulk; counts[1] = (int32_t)n32; counts[2] = 0; counts[3] = (int32_t)ntail;
    counts[8] = (int32_t)ng; counts[10] = (int32_t)nt;
    CU(cudaMemcpyAsync(db->d_counts, counts, sizeof(counts), cudaMemcpyHostToDevice, st));
    if (bulk.npass > 1 || tail.npass > 1) {
        if (db->bndH.reserve((size_t)db->ncols + 16) || db->bndF.reserve((size_t)db->ncols + 16)) return PYOPAL_CUDA_ERR_NOMEM;
        if (n32 > 0 || local)
            if (db->bnd32H.reserve((size_t)db->ncols * 2 + 16) || db->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                return PYOPAL_CUDA_ERR_NOMEM;
    }

    KParams P;
    memset(&P, 0, sizeof(P));
    P.stream = db->d_stream; P.pair_off = db->d_pair_off; P.pair_len = db->d_pair_len; P.pair_idx = db->d_pair_idx;
    P.query = db->query.p; P.matrix = db->d_matrix;
    P.Lq = Lq; P.A = A; P.go = go; P.ge = ge;
    P.mode = mode; P.pad_score = rung.pad_score; P.ovf_thresh = rung.ovf_thresh;
    P.one = 1u; P.bias = 0;
    P.out = d_out; P.out_start = start; P.out_end = end;
    P.ovf_list = db->ovf.p; P.ovf_count = db->d_counts + 2;
    P.bndH = db->bndH.p; P.bndF = db->bndF.p;

    search_kernel_t k32 = nullptr;
    if (wide) {
        apply_plan(P, tail);
        if (db->wprofile.reserve((size_t)(A + 1) * tail.pitch)) return PYOPAL_CUDA_ERR_NOMEM;
        P.wprofile = db->wprofile.p;
        build_profile_kernel<<<64, 256, 0, st>>>(P, nullptr, db->wprofile.p);
        CU(cudaGetLastError());
        k32 = bfix ? get_search_kernel_bfix(true) : get_search_kernel_wide(local);
    } else {
        if (!tail.psmem) {
            apply_plan(P, tail);
            if (db->gprofile.reserve((size_t)(A + 1) * tail.pitch)) return PYOPAL_CUDA_ERR_NOMEM;
            P.gprofile = db->gprofile.p;
            build_profile_kernel<<<64, 256, 0, st>>>(P, db->gprofile.p, nullptr);
            CU(cudaGetLastError());
        }
        k32 = bfix ? get_search_kernel_bfix(false) : pick_kernel(tail.K, 32, local, tail.psmem);
    }
    if (!k32) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 32-bit kernel for this tile");

    // ---- side stream: longest targets first, concurrently with the bulk -------
    const bool side_work = (ng > 0) || (nt > 0) || (n32 > 0) || (ntail > 0 && !same_plan);
    cudaStream_t side = db->side;
    if (side_work) {
        CU(cudaEventRecord(db->ev_fork, st));
        CU(cudaStreamWaitEvent(side, db->ev_fork, 0));
        CU(cudaEventRecord(db->kev[0], side));
    }
    if (nt > 0) {
        search_kernel_t kt = get_tgiant_kernel(tK, local);
        if (!kt) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no transposed giant kernel for this tile");
        KParams Q = P;
        Q.G = tNW; Q.logG = 0; Q.npass = 1; Q.pitch = tpitch; Q.padtop = 0;
        Q.krows = tK; Q.kstride = tK + 4;
        // padding rows above target position 0 reproduce the border H[i][-1] (kernels.cuh, profile_entry)
        Q.pad_score = ((mode == MODE_NW || mode == MODE_HW) ? -std::max(go, ge) : 0) + go;
        Q.worklist = db->wl32.p + n32 + ng; Q.nwork = db->d_counts + 10; Q.counter = db->d_counts + 11;
        Q.neg = -(1 << 29);
        // the whole SM's shared memory: no throughput CTA shares the SM with a giant (giant.cu)
        const size_t smem_all = std::max<size_t>(tsmem, (size_t)smem_optin - 1024);
        CU(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_all));
        // its own stream: the few longest targets start at once, beside giant.cu's medium ones
        CU(cudaStreamWaitEvent(db->side2, db->ev_fork, 0));
        kt<<<(unsigned)std::min<int64_t>(nt, db->sm_count), tNW * 32, smem_all, db->side2>>>(Q);
        CU(cudaGetLastError());
        CU(cudaEventRecord(db->ev_join2, db->side2));
        db->stats[0]++;
    }
    if (ng > 0) {
        search_kernel_t kgiant = get_giant_kernel(gK, local);
        if (!kgiant) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no giant kernel for this tile");
        KParams Q = P;
        // a few teams per CTA keep the column chains short; more CTAs than that would only idle SMs
        const int teams = (int)std::max<int64_t>(1, std::min<int64_t>(gteams, (ng + 7) / 8));
        Q.G = gwarps; Q.logG = 0; Q.npass = teams; Q.pitch = gpitch; Q.padtop = gwarps * 32 * gK - Lq;
        Q.krows = gK; Q.kstride = (gK == 4) ? 4 : gK + 4;
        Q.worklist = db->wl32.p + n32; Q.nwork = db->d_counts + 8; Q.counter = db->d_counts + 9;
        Q.neg = -(1 << 29);
        // ask for the whole SM's shared memory: no throughput CTA may share the SM with the giants (giant.cu)
        const size_t gsmem_all = std::max<size_t>(gsmem, (size_t)smem_optin - 1024);
        CU(cudaFuncSetAttribute(kgiant, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem_all));
        // with the longest targets on tgiant.cu, what is left here is many medium targets: a quarter of the
        // SMs is enough for their latency and leaves the rest to the bulk kernel from the start
        const int64_t gcap = use_t ? std::max<int64_t>(8, db->sm_count / 4) : db->sm_count;
        kgiant<<<(unsigned)std::min<int64_t>((ng + teams - 1) / teams, gcap), 512, gsmem_all, side>>>(Q);
        CU(cudaGetLastError());
        db->stats[0]++;
    }
    if (n32 > 0) {
        KParams Q = P;
        apply_plan(Q, tail);
        Q.worklist = db->wl32.p; Q.nwork = db->d_counts + 1; Q.counter = db->d_counts + 5;
        Q.neg = -(1 << 29);
        Q.bndH = db->bnd32H.p; Q.bndF = db->bnd32F.p;
        int rc = launch_search(db, k32, Q, tail.smem, 32, n32, side);
        if (rc) return rc;
    }
    if (ntail > 0 && !same_plan) {
        search_kernel_t kt = pick_kernel(tail.K, 16, local, tail.psmem, rung.tdadd);
        if (!kt) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 16-bit kernel for the tail tile");
        KParams Q = P;
        apply_plan(Q, tail);
        Q.worklist = d_wl_tail; Q.nwork = db->d_counts + 3; Q.counter = db->d_counts + 7;
        Q.neg = rung.neg16; Q.bias = rung.bias16;
        int rc = launch_search(db, kt, Q, tail.smem, 32, ntail, side);
        if (rc) return rc;
    }
    if (side_work) {
        if (nt > 0) CU(cudaStreamWaitEvent(side, db->ev_join2, 0));  // joins only after everything is launched
        CU(cudaEventRecord(db->kev[1], side));
        db->kev_valid[0] = true;
        CU(cudaEventRecord(db->ev_join, side));
    }

    // ---- bulk -------------------------------------------------------------------
    {
        const int32_t* wl = d_wl_bulk;
        const int64_t nw = nbulk;
        if (nw > 0) {
            search_kernel_t kb = pick_kernel(bulk.K, 16, local, bulk.psmem, rung.tdadd);
            if (!kb) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 16-bit kernel for this tile");
            KParams Q = P;
            apply_plan(Q, bulk);
            Q.worklist = wl; Q.nwork = db->d_counts + 0; Q.counter = db->d_counts + 4;
            Q.neg = rung.neg16; Q.bias = rung.bias16;
            const int gpw = 32 >> bulk.logG;
            // one CTA per SM when the profile is large: use the whole register file with 16 warps
            // (with side-stream CTAs resident, 7-warp CTAs leave them registers: 2 x 224 x 128 < 64K)
            int threads = (bulk.smem > (size_t)110 * 1024) ? (side_work ? 480 : 512) : (side_work ? 224 : 256);
            if (const char* ov = getenv("PYOPAL_CUDA_PLAN")) {
                int lg = 0, k = 0, t = 0;
                if (sscanf(ov, "%d,%d,%d", &lg, &k, &t) == 3 && t >= 32 && t <= 512 && t % 32 == 0) threads = t;
            }
            CU(cudaEventRecord(db->kev[2], st));
            if (getenv("PYOPAL_CUDA_DEBUG"))
                fprintf(stderr, "[pyopal_cuda] Lq=%d bulk: G=%d K=%d npass=%d pairs=%lld | giants=%lld tgiants=%lld static32=%lld threads=%d smem=%zu\n",
                        Lq, bulk.G, bulk.K, bulk.npass, (long long)nw, (long long)ng, (long long)nt, (long long)n32, threads,
                        bulk.smem);
            int rc = launch_search(db, kb, Q, bulk.smem, threads, (nw + gpw - 1) / gpw, st);
            if (rc) return rc;
            CU(cudaEventRecord(db->kev[3], st));
            db->kev_valid[1] = true;
        }
    }
    if (side_work) CU(cudaStreamWaitEvent(st, db->ev_join, 0));

    // ---- rescue: 32-bit rescore of what overflowed (list and size live on the device)
    if (local && n16 > 0) {
        KParams R = P;
        apply_plan(R, tail);
        R.worklist = db->ovf.p; R.nwork = db->d_counts + 2; R.counter = db->d_counts + 6;
        R.neg = -(1 << 29);
        R.bndH = db->bnd32H.p; R.bndF = db->bnd32F.p;
        CU(cudaEventRecord(db->kev[4], st));
        int rc = launch_search(db, k32, R, tail.smem, 32, std::min<int64_t>(2 * n16, (int64_t)db->sm_count * 4), st);
        if (rc) return rc;
        CU(cudaEventRecord(db->kev[5], st));
        db->kev_valid[2] = true;
    }
    db->stats[1] = n16 * 2; db->stats[2] = n32 + ng + nt;
    db->stats[5] = bulk.G; db->stats[6] = bulk.K; db->stats[7] = bulk.npass; db->stats[11] = bulk.psmem ? 1 : 0;
    return apply_fixes();
}

int check_search_args(pyopal_cuda_db* db, const unsigned char* query, int Lq, int go, int ge, const int* M,
                      int A, int mode, int64_t* start, int64_t* end) {
    if (!db || !M || (Lq > 0 && !query)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    if (mode < MODE_NW || mode > MODE_SW) return fail(OPAL_ERR_INVALID_MODE, "invalid alignment mode");
    if (A != db->A) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "alphabet length differs from the store's");
    if (Lq < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative query length");
    if (go < 0 || ge < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "gap penalties must be non-negative");
    for (int i = 0; i < Lq; i++)
        if (query[i] >= A) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "query residue code outside the alphabet");
    if (*start < 0) *start = 0;
    if (*end > db->n) *end = db->n;
    if (*end < *start) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "slice end is lower than start");
    return 0;
}

}  // namespace

// ---------------------------------------------------------------------------
// Public search entry points

extern "C" int pyopal_cuda_search_scores_device(pyopal_cuda_db* db, const unsigned char* query, int queryLength,
                                                int gapOpen, int gapExt, const int* scoreMatrix,
                                                int alphabetLength, int mode, int64_t start, int64_t end,
                                                int32_t* d_scores, void* stream) {
    int rc = check_search_args(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, &start, &end);
    if (rc) return rc;
    if (end > start && !d_scores) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    std::lock_guard<std::mutex> lock(db->mu);
    CU(cudaSetDevice(db->device));
    return enqueue_score_search(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start,
                                end, d_scores, (cudaStream_t)stream);
}

extern "C" int pyopal_cuda_search(pyopal_cuda_db* db, const unsigned char* query, int queryLength, int gapOpen,
                                  int gapExt, const int* scoreMatrix, int alphabetLength, int searchType, int mode,
                                  int overflowMethod, int64_t start, int64_t end, int32_t* scores, int32_t* endQ,
                                  int32_t* endT, int32_t* startQ, int32_t* startT, unsigned char* ops,
                                  const int64_t* ops_offsets, int32_t* ops_len) {
    int rc = check_search_args(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, &start, &end);
    if (rc) return rc;
    if (searchType < OPAL_SEARCH_SCORE || searchType > OPAL_SEARCH_ALIGNMENT)
        return fail(OPAL_ERR_INVALID_MODE, "invalid search type");
    if (overflowMethod != OPAL_OVERFLOW_SIMPLE && overflowMethod != OPAL_OVERFLOW_BUCKETS)
        return fail(OPAL_ERR_INVALID_MODE, "invalid overflow method");
    const int64_t nout = end - start;
    if (nout <= 0) return 0;
    if (!scores) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    std::lock_guard<std::mutex> lock(db->mu);
    CU(cudaSetDevice(db->device));
    if (searchType != OPAL_SEARCH_SCORE)
        return pyopal_end_full_search(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength,
                                      searchType, mode, start, end, scores, endQ, endT, startQ, startT, ops,
                                      ops_offsets, ops_len);
    if (db->out.reserve((size_t)nout)) return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaEventRecord(db->ev0, db->stream));
    rc = enqueue_score_search(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end,
                              db->out.p, db->stream);
    if (rc) return rc;
    CU(cudaMemcpyAsync(scores, db->out.p, (size_t)nout * sizeof(int32_t), cudaMemcpyDeviceToHost, db->stream));
    CU(cudaEventRecord(db->ev1, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, db->ev0, db->ev1));
    db->stats[4] = (int64_t)(ms * 1000.0f);
    int32_t novf = 0;
    CU(cudaMemcpy(&novf, db->d_counts + 2, sizeof(int32_t), cudaMemcpyDeviceToHost));
    db->stats[3] = novf;
    return 0;
}

extern "C" int pyopal_cuda_search_batch(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                                        const int32_t* q_lengths, int nq, int gapOpen, int gapExt,
                                        const int* scoreMatrix, int alphabetLength, int mode, int32_t* scores) {
    if (!db || nq < 0 || (nq > 0 && (!queries || !q_offsets || !q_lengths || !scores)))
        return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    const int64_t n = db->n;
    if (n == 0) return 0;
    std::lock_guard<std::mutex> lock(db->mu);
    CU(cudaSetDevice(db->device));
    // double-buffer the score rows so the D2H of query i overlaps the kernels of query i+1
    int32_t* d_rows[2] = {nullptr, nullptr};
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t done[2] = {nullptr, nullptr}, copied[2] = {nullptr, nullptr};
    int rc = 0;
    auto cleanup = [&]() {
        for (int b = 0; b < 2; b++) {
            if (d_rows[b]) cudaFree(d_rows[b]);
            if (done[b]) cudaEventDestroy(done[b]);
            if (copied[b]) cudaEventDestroy(copied[b]);
        }
        if (copy_stream) cudaStreamDestroy(copy_stream);
    };
#define CUB(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(PYOPAL_CUDA_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); } } while (0)
    CUB(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
    for (int b = 0; b < 2; b++) {
        CUB(cudaMalloc(&d_rows[b], (size_t)n * sizeof(int32_t)));
        CUB(cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming));
        CUB(cudaEventCreateWithFlags(&copied[b], cudaEventDisableTiming));
    }
    for (int i = 0; i < nq && rc == 0; i++) {
        const int b = i & 1;
        int64_t s = 0, e = n;
        const unsigned char* q = queries + q_offsets[i];
        rc = check_search_args(db, q, q_lengths[i], gapOpen, gapExt, scoreMatrix, alphabetLength, mode, &s, &e);
        if (rc) break;
        if (i >= 2) CUB(cudaStreamWaitEvent(db->stream, copied[b], 0));
        rc = enqueue_score_search(db, q, q_lengths[i], gapOpen, gapExt, scoreMatrix, alphabetLength, mode, 0, n,
                                  d_rows[b], db->stream);
        if (rc) break;
        CUB(cudaEventRecord(done[b], db->stream));
        CUB(cudaStreamWaitEvent(copy_stream, done[b], 0));
        CUB(cudaMemcpyAsync(scores + (size_t)i * (size_t)n, d_rows[b], (size_t)n * sizeof(int32_t),
                            cudaMemcpyDeviceToHost, copy_stream));
        CUB(cudaEventRecord(copied[b], copy_stream));
        // the per-search workspace (query, counters, lists) is reused by the next query
        CUB(cudaStreamSynchronize(db->stream));
    }
    if (rc == 0) CUB(cudaStreamSynchronize(copy_stream));
#undef CUB
    cleanup();
    return rc;
}

extern "C" int pyopal_cuda_topk_device(pyopal_cuda_db* db, const int32_t* d_scores, int64_t n, int k,
                                       int32_t* d_scores_out, int32_t* d_indices_out, void* stream) {
    if (!db || (n > 0 && k > 0 && (!d_scores || !d_scores_out || !d_indices_out))) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    if (k < 0 || n < 0 || k > n || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "top-k needs 0 <= k <= n < 2^31");
    std::lock_guard<std::mutex> lock(db->mu);
    CU(cudaSetDevice(db->device));
    if (db->tk_ws.reserve(topk_workspace_words())) return PYOPAL_CUDA_ERR_NOMEM;
    return topk_device(d_scores, n, k, d_scores_out, d_indices_out, db->tk_ws.p, db->sm_count, (cudaStream_t)stream);
}

extern "C" int pyopal_cuda_search_topk(pyopal_cuda_db* db, const unsigned char* query, int queryLength, int gapOpen,
                                       int gapExt, const int* scoreMatrix, int alphabetLength, int mode, int64_t start,
                                       int64_t end, int k, int32_t* scores, int32_t* indices, int32_t* n_hits) {
    int rc = check_search_args(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, &start, &end);
    if (rc) return rc;
    if (k < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative k");
    const int64_t nout = end - start;
    const int kk = (int)std::min<int64_t>(k, nout);
    if (n_hits) *n_hits = kk > 0 ? kk : 0;
    if (kk <= 0) return 0;
    if (!scores || !indices) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    std::lock_guard<std::mutex> lock(db->mu);
    CU(cudaSetDevice(db->device));
    if (db->out.reserve((size_t)nout) || db->tk_ws.reserve(topk_workspace_words()) || db->tk_vals.reserve((size_t)kk) ||
        db->tk_idx.reserve((size_t)kk))
        return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaEventRecord(db->ev0, db->stream));
    rc = enqueue_score_search(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end,
                              db->out.p, db->stream);
    if (rc) return rc;
    rc = topk_device(db->out.p, nout, kk, db->tk_vals.p, db->tk_idx.p, db->tk_ws.p, db->sm_count, db->stream);
    if (rc) return rc;
    db->stats[0] += 9;
    std::vector<int32_t> v((size_t)kk), ix((size_t)kk);
    CU(cudaMemcpyAsync(v.data(), db->tk_vals.p, (size_t)kk * sizeof(int32_t), cudaMemcpyDeviceToHost, db->stream));
    CU(cudaMemcpyAsync(ix.data(), db->tk_idx.p, (size_t)kk * sizeof(int32_t), cudaMemcpyDeviceToHost, db->stream));
    CU(cudaEventRecord(db->ev1, db->stream));
    CU(cudaStreamSynchronize(db->stream));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, db->ev0, db->ev1));
    db->stats[4] = (int64_t)(ms * 1000.0f);
    // the k hits: score descending, ties by ascending database index
    std::vector<int32_t> order((size_t)kk);
    std::iota(order.begin(), order.end(), 0);
    std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) {
        return v[(size_t)a] != v[(size_t)b] ? v[(size_t)a] > v[(size_t)b] : ix[(size_t)a] < ix[(size_t)b];
    });
    for (int i = 0; i < kk; i++) {
        scores[i] = v[(size_t)order[(size_t)i]];
        indices[i] = (int32_t)(ix[(size_t)order[(size_t)i]] + start);
    }
    return 0;
}

// Generic path with the reference's own signature (opal.pxd:40-54): stage the
// database, search, unpack into the caller's OpalSearchResult structs.
extern "C" int opalSearchDatabase(unsigned char query[], int queryLength, unsigned char* db[], int dbLength,
                                  int dbSeqLengths[], int gapOpen, int gapExt, int* scoreMatrix,
                                  int alphabetLength, OpalSearchResult* results[], const int searchType, int mode,
                                  int overflowMethod) {
    if (dbLength < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative database length");
    if (dbLength == 0) return 0;
    if (!db || !dbSeqLengths || !results) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    int device = 0;
    if (pyopal_cuda_device_count() <= 0) return fail(OPAL_ERR_NO_SIMD_SUPPORT, "no CUDA device available");
    cudaGetDevice(&device);
    pyopal_cuda_db* h = nullptr;
    int rc = pyopal_cuda_db_create_from_pointers(device, (const unsigned char* const*)db, dbSeqLengths, dbLength,
                                                 alphabetLength, &h);
    if (rc) return rc;
    const size_t n = (size_t)dbLength;
    std::vector<int32_t> sc(n), eq(n, -1), et(n, -1), sq(n, -1), st(n, -1), ol(n, 0);
    std::vector<int64_t> oo;
    std::vector<unsigned char> ops;
    if (searchType == OPAL_SEARCH_ALIGNMENT) {
        oo.resize(n);
        int64_t tot = 0;
        for (size_t k = 0; k < n; k++) { oo[k] = tot; tot += (int64_t)queryLength + dbSeqLengths[k]; }
        ops.resize((size_t)tot + 1);
    }
    rc = pyopal_cuda_search(h, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, searchType, mode,
                            overflowMethod, 0, dbLength, sc.data(), eq.data(), et.data(), sq.data(), st.data(),
                            ops.empty() ? nullptr : ops.data(), oo.empty() ? nullptr : oo.data(), ol.data());
    pyopal_cuda_db_destroy(h);
    if (rc) return rc;
    for (size_t k = 0; k < n; k++) {
        OpalSearchResult* r = results[k];
        if (!r) continue;
        r->scoreSet = 1;
        r->score = sc[k];
        if (searchType >= OPAL_SEARCH_SCORE_END) { r->endLocationQuery = eq[k]; r->endLocationTarget = et[k]; }
        if (searchType == OPAL_SEARCH_ALIGNMENT) {
            r->startLocationQuery = sq[k];
            r->startLocationTarget = st[k];
            r->alignmentLength = ol[k];
            r->alignment = nullptr;
            if (ol[k] > 0) {
                r->alignment = (unsigned char*)malloc((size_t)ol[k]);
                if (!r->alignment) return fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed");
                memcpy(r->alignment, ops.data() + oo[k], (size_t)ol[k]);
            }
        }
    }
    return 0;
}
