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

int pyopal_cuda_device_pci_bus_id(int device, char* out, int cap) {
    if (!out || cap < 16) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "buffer too small");
    CU(cudaDeviceGetPCIBusId(out, cap, device));
    return 0;
}

void pyopal_cuda_db_destroy(pyopal_cuda_db* db) {
    if (!db) return;
    DeviceGuard guard(db->device);
    cudaDeviceSynchronize();
    cudaGetLastError();
    cudaFree(db->d_stream);
    cudaFree(db->d_pair_off);
    cudaFree(db->d_pair_len);
    cudaFree(db->d_pair_idx);
    cudaFree(db->d_identity);
    for (auto& ws : db->pool) ws->destroy();
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
    pyopal_cuda_db* m = const_cast<pyopal_cuda_db*>(db);
    std::lock_guard<std::mutex> lock(m->mu);
    // resolve the per-kernel events of the last search if they have completed
    if (Workspace* ws = m->last_ws) {
        DeviceGuard guard(db->device);
        for (int k = 0; k < 3; k++) {
            m->last_stats[8 + k] = 0;
            if (ws->kev_valid[k] && cudaEventQuery(ws->kev[2 * k + 1]) == cudaSuccess) {
                float ms = 0;
                if (cudaEventElapsedTime(&ms, ws->kev[2 * k], ws->kev[2 * k + 1]) == cudaSuccess)
                    m->last_stats[8 + k] = (int64_t)(ms * 1000.0f + 0.5f);
            }
        }
        cudaGetLastError();
    }
    for (int i = 0; i < n && i < 12; i++) stats[i] = m->last_stats[i];
    return 0;
}

int pyopal_host_batch_units(const int32_t* q_lengths, int nq, int pair_queries, int32_t* units) {
    if (nq < 0 || (nq > 0 && (!q_lengths || !units))) return -1;
    std::vector<std::pair<int, int>> u;
    pyopal::batch_units(q_lengths, nq, pair_queries != 0, u);
    for (size_t i = 0; i < u.size(); i++) { units[2 * i] = u[i].first; units[2 * i + 1] = u[i].second; }
    return (int)u.size();
}

void pyopal_host_offsets_from_lengths(const int32_t* lengths, int64_t n, int64_t* offsets, int64_t* total) {
    int64_t t = 0;
    for (int64_t i = 0; i < n; i++) {
        if (offsets) offsets[i] = t;
        t += lengths[i] > 0 ? lengths[i] : 0;
    }
    if (total) *total = t;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Tuning switches, workspaces

void Tuning::load() {
    if (const char* ov = getenv("PYOPAL_CUDA_PLAN")) {  // "logG,K[,threads]"
        int lg = 0, k = 0, t = 0;
        const int got = sscanf(ov, "%d,%d,%d", &lg, &k, &t);
        if (got >= 2 && lg >= 2 && lg <= 5 && k >= 4 && k <= 40 && k % 4 == 0) {
            has_plan = true; plan_logG = lg; plan_K = k;
            if (got == 3 && t >= 32 && t <= 512 && t % 32 == 0) plan_threads = t;
        }
    }
    no_giant = getenv("PYOPAL_CUDA_NO_GIANT") != nullptr;
    no_tgiant = getenv("PYOPAL_CUDA_NO_TGIANT") != nullptr;
    no_tail = getenv("PYOPAL_CUDA_NO_TAIL") != nullptr;
    no_end16 = getenv("PYOPAL_CUDA_NO_END16") != nullptr;
    no_mid = getenv("PYOPAL_CUDA_NO_MID") != nullptr;
    no_stream = getenv("PYOPAL_CUDA_NO_STREAM") != nullptr;
    no_giant16 = getenv("PYOPAL_CUDA_NO_GIANT16") != nullptr;
    no_dual = getenv("PYOPAL_CUDA_NO_DUAL") != nullptr;
    if (const char* v = getenv("PYOPAL_CUDA_DUAL")) force_dual = atoi(v) > 0;
    if (const char* v = getenv("PYOPAL_CUDA_STREAM")) force_stream = atoi(v) > 0;
    debug = getenv("PYOPAL_CUDA_DEBUG") != nullptr;
    if (const char* ov = getenv("PYOPAL_CUDA_TGIANT")) {  // "warps,K[,split]"
        int a = 0, b = 0, c = -1;
        const int got = sscanf(ov, "%d,%d,%d", &a, &b, &c);
        has_tgiant = true;
        if (got >= 2 && a >= 1 && a <= 8 && (b == 16 || b == 32)) { tg_warps = a; tg_K = b; }
        if (got == 3 && c >= 0) tg_split = c;
    }
    if (const char* v = getenv("PYOPAL_CUDA_SOLO")) solo = atof(v);
    if (const char* v = getenv("PYOPAL_CUDA_GMUL")) gmul = atof(v);
    if (const char* v = getenv("PYOPAL_CUDA_MID")) mid = atof(v);
    if (const char* v = getenv("PYOPAL_CUDA_LANES")) lanes = atoi(v) >= 2 ? 2 : 1;
    if (const char* v = getenv("PYOPAL_CUDA_PAIR_K")) pair_k = atoi(v);
    if (const char* v = getenv("PYOPAL_CUDA_PAIR_C")) pair_c = atoi(v);
}

int Workspace::init(int dev) {
    device = dev;
    CU(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    // the side streams carry the longest targets: highest priority so their CTAs are placed first
    int lo_prio = 0, hi_prio = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
    CU(cudaStreamCreateWithPriority(&side, cudaStreamNonBlocking, hi_prio));
    CU(cudaStreamCreateWithPriority(&side2, cudaStreamNonBlocking, hi_prio));
    CU(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_join, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_join2, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_free, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_aux0, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ev_aux1, cudaEventDisableTiming));
    CU(cudaEventCreate(&ev0));
    CU(cudaEventCreate(&ev1));
    for (int i = 0; i < 6; i++) CU(cudaEventCreate(&kev[i]));
    for (int i = 0; i < NPARAM; i++) CU(cudaEventCreateWithFlags(&par_done[i], cudaEventDisableTiming));
    return 0;
}

void Workspace::destroy() {
    ovf.release(); ovf2.release(); out.release(); out2.release(); gprofile.release(); wprofile.release(); end_out.release();
    bndH.release(); bndF.release(); bnd32H.release(); bnd32F.release();
    pd_i32.release(); pd_tb.release(); pd_ops.release(); pd_jobs.release(); pd_team.release(); pd_query.release();
    tk_ws.release(); tk_vals.release(); tk_idx.release(); tk_vals2.release(); tk_idx2.release();
    d_par.release();
    for (int i = 0; i < NPARAM; i++) { h_par[i].release(); if (par_done[i]) cudaEventDestroy(par_done[i]); par_done[i] = nullptr; }
    for (cudaEvent_t* e : {&ev_fork, &ev_join, &ev_join2, &ev_free, &ev_aux0, &ev_aux1, &ev0, &ev1}) { if (*e) cudaEventDestroy(*e); *e = nullptr; }
    for (int i = 0; i < 6; i++) { if (kev[i]) cudaEventDestroy(kev[i]); kev[i] = nullptr; }
    for (cudaStream_t* s : {&side, &side2, &stream}) { if (*s) cudaStreamDestroy(*s); *s = nullptr; }
}

Workspace* pyopal_cuda_db::acquire() {
    std::unique_lock<std::mutex> lock(mu);
    for (;;) {
        if (!free_ws.empty()) {
            Workspace* ws = free_ws.back();
            free_ws.pop_back();
            return ws;
        }
        if ((int)pool.size() < max_ws) {
            std::unique_ptr<Workspace> ws(new (std::nothrow) Workspace());
            if (!ws) { fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed"); return nullptr; }
            DeviceGuard guard(device);
            if (ws->init(device)) { ws->destroy(); return nullptr; }
            pool.push_back(std::move(ws));
            return pool.back().get();
        }
        cv.wait(lock);
    }
}

Workspace* pyopal_cuda_db::try_acquire() {
    std::unique_lock<std::mutex> lock(mu);
    if (!free_ws.empty()) {
        Workspace* ws = free_ws.back();
        free_ws.pop_back();
        return ws;
    }
    if ((int)pool.size() >= max_ws) return nullptr;
    std::unique_ptr<Workspace> ws(new (std::nothrow) Workspace());
    if (!ws) return nullptr;
    DeviceGuard guard(device);
    if (ws->init(device)) { ws->destroy(); return nullptr; }
    pool.push_back(std::move(ws));
    return pool.back().get();
}

void pyopal_cuda_db::release(Workspace* ws) {
    {
        std::lock_guard<std::mutex> lock(mu);
        for (int i = 0; i < 12; i++) last_stats[i] = ws->stats[i];
        last_ws = ws;
        free_ws.push_back(ws);
    }
    cv.notify_one();
}

void pyopal_cuda_db::publish(Workspace* ws) {
    std::lock_guard<std::mutex> lock(mu);
    for (int i = 0; i < 12; i++) last_stats[i] = ws->stats[i];
    last_ws = ws;
}

// ---------------------------------------------------------------------------
// Store construction: sort by length (descending, ties by index), pair neighbours, upload the raw residues once
// and interleave each pair's residues into one u16 stream ON THE DEVICE.

namespace {

// One warp per pair: column c of pair k = a[c] | b[c] << 8 (padding letter A past the end of the shorter target
// and up to the 16-byte boundary).  Also checks that every residue indexes the profile (< A).
__global__ void interleave_kernel(const unsigned char* __restrict__ raw, const int64_t* __restrict__ src_off,
                                  const int64_t* __restrict__ pair_off, const int32_t* __restrict__ pair_len,
                                  int64_t npairs, int A, uint16_t* __restrict__ stream, int* bad) {
    const int lane = threadIdx.x & 31;
    const int64_t warps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    unsigned any_bad = 0;
    for (int64_t k = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; k < npairs; k += warps) {
        const int la = pair_len[2 * k], lb = pair_len[2 * k + 1];
        const unsigned char* a = raw + src_off[2 * k];
        const unsigned char* b = lb > 0 ? raw + src_off[2 * k + 1] : nullptr;
        uint16_t* dst = stream + pair_off[k];
        const int padded = (la + PAD_ALIGN_COLS - 1) / PAD_ALIGN_COLS * PAD_ALIGN_COLS;
        for (int c = lane; c < padded; c += 32) {
            unsigned ra = (c < la) ? a[c] : (unsigned)A;
            unsigned rb = (c < lb) ? b[c] : (unsigned)A;
            any_bad |= (unsigned)(c < la && ra >= (unsigned)A) | (unsigned)(c < lb && rb >= (unsigned)A);
            dst[c] = (uint16_t)(ra | (rb << 8));
        }
    }
    if (any_bad) atomicOr(bad, 1);
}

__global__ void iota_kernel(int32_t* p, int64_t n) {
    for (int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; x < n; x += (int64_t)gridDim.x * blockDim.x) p[x] = (int32_t)x;
}

}  // namespace

int pyopal::db_build_from(int device, const unsigned char* residues, const int64_t* offsets,
                          const unsigned char* const* pointers, const int32_t* index, const int32_t* lengths, int64_t n,
                          int A, pyopal_cuda_db** out) {
    if (!out) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output handle");
    *out = nullptr;
    if (n < 0 || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "database size out of range");
    if (A < 1 || A > 32) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "alphabet length must be in 1..32");
    if (pyopal_cuda_device_count() <= 0) return fail(OPAL_ERR_NO_SIMD_SUPPORT, "no CUDA device available");
    DeviceGuard dguard(device);
    {
        int cur = -1;
        CU(cudaGetDevice(&cur));
        if (cur != device) CU(cudaSetDevice(device));
    }
    pyopal_cuda_db* db = new (std::nothrow) pyopal_cuda_db();
    if (!db) return fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed");
    struct Guard {
        pyopal_cuda_db* d;
        ~Guard() { if (d) pyopal_cuda_db_destroy(d); }
    } guard{db};
    db->device = device;
    db->A = A;
    db->n = n;
    db->tune.load();
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    db->sm_count = prop.multiProcessorCount;
    CU(cudaDeviceGetAttribute(&db->smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, device));

    // ---- sort by length, descending, ties by index: counting sort (lengths are small integers) ----
    db->h_lengths.assign(lengths, lengths + n);
    int32_t maxlen = 0;
    int64_t m = 0;
    for (int64_t i = 0; i < n; i++) {
        if (lengths[i] < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative sequence length");
        db->total_len += lengths[i];
        maxlen = std::max(maxlen, lengths[i]);
        if (lengths[i] == 0) db->h_zero_idx.push_back((int32_t)i);
        else m++;
    }
    std::vector<int32_t> order((size_t)m);
    if ((int64_t)maxlen <= std::max<int64_t>(4 * n, (int64_t)1 << 20)) {
        std::vector<int64_t> first((size_t)maxlen + 2, 0);
        for (int64_t i = 0; i < n; i++) first[(size_t)lengths[i]]++;
        int64_t pos = 0;
        for (int64_t l = maxlen; l >= 1; l--) { const int64_t c = first[(size_t)l]; first[(size_t)l] = pos; pos += c; }
        for (int64_t i = 0; i < n; i++)
            if (lengths[i] > 0) order[(size_t)first[(size_t)lengths[i]]++] = (int32_t)i;
    } else {
        int64_t w = 0;
        for (int64_t i = 0; i < n; i++) if (lengths[i] > 0) order[(size_t)w++] = (int32_t)i;
        std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return lengths[a] > lengths[b]; });
    }
    const int64_t npairs = (m + 1) / 2;
    db->npairs = npairs;
    db->h_pair_off.resize((size_t)npairs);
    db->h_pair_len.resize((size_t)npairs * 2);
    db->h_pair_idx.resize((size_t)npairs * 2);
    db->h_where.assign((size_t)n, -1);
    int64_t ncols = 0;
    for (int64_t k = 0; k < npairs; k++) {
        const int32_t a = order[(size_t)(2 * k)];
        const int32_t b = (2 * k + 1 < m) ? order[(size_t)(2 * k + 1)] : -1;
        db->h_pair_off[(size_t)k] = ncols;
        db->h_pair_len[(size_t)(2 * k)] = lengths[a];
        db->h_pair_len[(size_t)(2 * k + 1)] = b >= 0 ? lengths[b] : 0;
        db->h_pair_idx[(size_t)(2 * k)] = a;
        db->h_pair_idx[(size_t)(2 * k + 1)] = b;
        db->h_where[(size_t)a] = (int32_t)(2 * k);
        if (b >= 0) db->h_where[(size_t)b] = (int32_t)(2 * k + 1);
        ncols += ((int64_t)lengths[a] + PAD_ALIGN_COLS - 1) / PAD_ALIGN_COLS * PAD_ALIGN_COLS;
    }
    db->ncols = ncols;
    db->mean_cols = npairs ? (double)ncols / (double)npairs : 0.0;

    if (npairs > 0) {
        CU(cudaMalloc(&db->d_pair_off, (size_t)npairs * sizeof(int64_t)));
        CU(cudaMalloc(&db->d_pair_len, (size_t)npairs * 2 * sizeof(int32_t)));
        CU(cudaMalloc(&db->d_pair_idx, (size_t)npairs * 2 * sizeof(int32_t)));
        CU(cudaMalloc(&db->d_identity, (size_t)npairs * sizeof(int32_t)));
        CU(cudaMemcpy(db->d_pair_off, db->h_pair_off.data(), (size_t)npairs * sizeof(int64_t), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(db->d_pair_len, db->h_pair_len.data(), (size_t)npairs * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        CU(cudaMemcpy(db->d_pair_idx, db->h_pair_idx.data(), (size_t)npairs * 2 * sizeof(int32_t), cudaMemcpyHostToDevice));
        iota_kernel<<<256, 256>>>(db->d_identity, npairs);
        CU(cudaGetLastError());
    }

    // ---- raw residues to the device (one copy), interleave there ----
    if (ncols > 0) {
        CU(cudaMalloc(&db->d_stream, (size_t)ncols * sizeof(uint16_t)));
        struct Tmp {
            unsigned char* raw = nullptr;
            int64_t* off = nullptr;
            int* bad = nullptr;
            unsigned char* stage = nullptr;
            ~Tmp() { cudaFree(raw); cudaFree(off); cudaFree(bad); if (stage) cudaFreeHost(stage); }
        } tmp;
        std::vector<int64_t> src((size_t)npairs * 2, 0);
        int64_t raw_bytes = 0;
        auto source = [&](int32_t local) -> int64_t { return index ? (int64_t)index[local] : (int64_t)local; };
        if (pointers == nullptr) {
            // packed input: upload the byte range that covers this store's sequences as it lies in host memory
            int64_t lo = INT64_MAX, hi = 0;
            for (int64_t x = 0; x < 2 * npairs; x++) {
                const int32_t i = db->h_pair_idx[(size_t)x];
                if (i < 0) continue;
                const int64_t o = offsets[source(i)];
                lo = std::min(lo, o);
                hi = std::max(hi, o + lengths[i]);
            }
            for (int64_t x = 0; x < 2 * npairs; x++) {
                const int32_t i = db->h_pair_idx[(size_t)x];
                src[(size_t)x] = i < 0 ? 0 : offsets[source(i)] - lo;
            }
            raw_bytes = hi - lo;
            CU(cudaMalloc(&tmp.raw, (size_t)raw_bytes));
            CU(cudaMemcpy(tmp.raw, residues + lo, (size_t)raw_bytes, cudaMemcpyHostToDevice));
        } else {
            // per-sequence pointers (the reference's own layout, lib.pxd:90-92): gather into pinned chunks in
            // sorted order on several host threads, copy chunk by chunk
            int64_t t = 0;
            for (int64_t x = 0; x < 2 * npairs; x++) {
                src[(size_t)x] = t;
                t += db->h_pair_len[(size_t)x];
            }
            raw_bytes = t;
            CU(cudaMalloc(&tmp.raw, (size_t)std::max<int64_t>(raw_bytes, 1)));
            const int64_t CHUNK = (int64_t)64 << 20;
            const int64_t stage_bytes = std::max<int64_t>(std::min(CHUNK, raw_bytes), (int64_t)db->h_pair_len[0]);
            CU(cudaMallocHost(&tmp.stage, (size_t)std::max<int64_t>(stage_bytes, 1)));
            int64_t x0 = 0;
            while (x0 < 2 * npairs) {
                int64_t x1 = x0;
                const int64_t base = src[(size_t)x0];
                while (x1 < 2 * npairs && src[(size_t)x1] + db->h_pair_len[(size_t)x1] - base <= stage_bytes) x1++;
                const int64_t bytes = (x1 < 2 * npairs ? src[(size_t)x1] : raw_bytes) - base;
                auto fill = [&](int64_t a, int64_t b) {
                    for (int64_t x = a; x < b; x++) {
                        const int32_t i = db->h_pair_idx[(size_t)x];
                        if (i < 0) continue;
                        memcpy(tmp.stage + (src[(size_t)x] - base), pointers[source(i)], (size_t)db->h_pair_len[(size_t)x]);
                    }
                };
                unsigned nth = std::min<unsigned>(8u, std::max(1u, std::thread::hardware_concurrency()));
                if (bytes < ((int64_t)4 << 20)) nth = 1;
                if (nth <= 1) fill(x0, x1);
                else {
                    std::vector<std::thread> pool;
                    for (unsigned t2 = 0; t2 < nth; t2++)
                        pool.emplace_back(fill, x0 + (x1 - x0) * t2 / nth, x0 + (x1 - x0) * (t2 + 1) / nth);
                    for (auto& th : pool) th.join();
                }
                CU(cudaMemcpy(tmp.raw + base, tmp.stage, (size_t)bytes, cudaMemcpyHostToDevice));
                x0 = x1;
            }
        }
        CU(cudaMalloc(&tmp.off, (size_t)npairs * 2 * sizeof(int64_t)));
        CU(cudaMalloc(&tmp.bad, sizeof(int)));
        CU(cudaMemset(tmp.bad, 0, sizeof(int)));
        CU(cudaMemcpy(tmp.off, src.data(), (size_t)npairs * 2 * sizeof(int64_t), cudaMemcpyHostToDevice));
        const int64_t want_blocks = (npairs + 7) / 8;
        const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>(want_blocks, (int64_t)db->sm_count * 32));
        interleave_kernel<<<blocks, 256>>>(tmp.raw, tmp.off, db->d_pair_off, db->d_pair_len, npairs, A, db->d_stream, tmp.bad);
        CU(cudaGetLastError());
        int bad = 0;
        CU(cudaMemcpy(&bad, tmp.bad, sizeof(int), cudaMemcpyDeviceToHost));
        if (bad) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "residue code outside the alphabet");
    }
    CU(cudaDeviceSynchronize());
    guard.d = nullptr;
    *out = db;
    return 0;
}

extern "C" int pyopal_cuda_db_create(int device, const unsigned char* residues, const int64_t* offsets,
                                     const int32_t* lengths, int64_t n, int alphabetLength,
                                     pyopal_cuda_db** out) {
    if (n > 0 && (!residues || !offsets || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    return db_build_from(device, residues, offsets, nullptr, nullptr, lengths, n, alphabetLength, out);
}

extern "C" int pyopal_cuda_db_create_from_pointers(int device, const unsigned char* const* sequences,
                                                   const int32_t* lengths, int64_t n, int alphabetLength,
                                                   pyopal_cuda_db** out) {
    if (n > 0 && (!sequences || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    static const unsigned char none = 0;
    return db_build_from(device, &none, nullptr, sequences, nullptr, lengths, n, alphabetLength, out);
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

// Latency tile for the longer targets: a whole warp per pair and as few rows
// per thread as one pass allows (the per-step critical path is ~12 cycles per
// row); with several passes, deep tiles amortise the per-step latency better.
Plan make_tail_plan(int Lq, int A, size_t smem_limit) {
    Plan p; p.G = 32; p.logG = 5;
    p.K = std::min(40, std::max(4, (Lq + 127) / 128 * 4));
    return finish_plan(p, Lq, A, smem_limit);
}

// first pair (pairs are sorted by length, descending) whose longer target has at most `len` columns
int64_t first_pair_at_most(const pyopal_cuda_db* db, double len, int64_t from) {
    int64_t lo = from, hi = db->npairs;
    while (lo < hi) {
        const int64_t mid = (lo + hi) / 2;
        if ((double)db->h_pair_len[(size_t)(2 * mid)] > len) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Tiers.  Pairs are sorted by length; the longest [0, kg) go to the giant kernels, the next [kg, kmid) -- too
// long to be walked by a narrow group within the time the whole search takes -- to the same 16-bit kernel in
// its 32-lane tile (`mid`), the rest to the throughput tile.  A wide tile costs the same instructions per
// cell as a narrow one of the same depth; what it loses is the G - 1 fill steps per pass, which only matter
// for short pairs -- so the split lets small shards (8 GPUs) keep the narrow tile for the bulk of their cells.
Plan make_plan(int Lq, int A, const pyopal_cuda_db* db, size_t smem_limit, GiantPolicy* gp, bool allow_mid,
               int64_t* kmid_out) {
    const double C0 = 28.0, C1 = 6.5;
    const Tuning& tune = db->tune;
    Plan best;
    double best_cost = 1e300;
    int64_t best_kg = 0, best_kmid = 0;
    const Plan tailp = make_tail_plan(Lq, A, smem_limit);
    auto cols_of = [&](int64_t k0, int64_t k1) {
        return (double)((k1 < db->npairs ? db->h_pair_off[(size_t)k1] : db->ncols) - (k0 < db->npairs ? db->h_pair_off[(size_t)k0] : db->ncols));
    };
    for (int logG = 2; logG <= 5; logG++) {
        if (tune.has_plan && logG != tune.plan_logG) continue;  // a forced tile still gets the giants that go with it
        const int G = 1 << logG;
        // calibrated on 1/2, 1/4 and 1/8 shards of config 2 (scratch/shard_sweep.py): groups of 8+ lanes reach the
        // critical path about 25 % earlier than 448 / G groups per SM suggest; 4-lane groups do not
        auto l_even = [&](int g) {
            const double solo = tune.solo > 0 ? tune.solo : (g == 4 ? 448.0 : 560.0);
            return (double)db->ncols * g / (solo * db->sm_count);
        };
        const bool mid = allow_mid && G < 32;
        // giants: beyond what the widest tile of this launch can walk in time
        const double L_top = mid ? l_even(32) : l_even(G);
        int64_t kg = 0;
        double giant_cols = 0;
        if (gp && gp->ok) {
            const double gthr = std::max(1024.0, 0.6 * L_top);
            while (kg < gp->cap && kg < db->npairs && (double)db->h_pair_len[(size_t)(2 * kg)] > gthr) {
                giant_cols += (double)db->h_pair_len[(size_t)(2 * kg)] + (double)db->h_pair_len[(size_t)(2 * kg + 1)];
                kg++;
            }
        }
        // the giants hold whole SMs at a fraction of the bulk kernel's efficiency
        const double giant_cost = giant_cols * (gp ? gp->col_cost : 0.0);
        const double midf = tune.mid > 0 ? tune.mid : 0.25;
        const int64_t kmid = mid ? first_pair_at_most(db, std::max(256.0, midf * l_even(G)), kg) : kg;
        const double cols_mid = cols_of(kg, kmid), n_mid = (double)(kmid - kg);
        // without a mid tier the bulk term covers every pair, the giants' columns included: the planner constants
        // (solo, gmul, the giants' col_cost) were calibrated on shards of config 2 with exactly this expression
        const double cols_bulk = mid ? cols_of(kmid, db->npairs) : (double)db->ncols;
        const double n_bulk = mid ? (double)(db->npairs - kmid) : (double)db->npairs;
        const double longest_bulk = kmid < db->npairs ? (double)db->h_pair_len[(size_t)(2 * kmid)] : 0.0;
        const double stretch = std::max(1.0, longest_bulk / std::max(l_even(G), 1.0));
        const double mid_cost = (double)tailp.npass * (cols_mid + 31.0 * n_mid) *
                                (C0 + (tailp.npass > 1 ? 12.0 : 0.0) + C1 * tailp.K) * 32.0;
        for (int K = 4; K <= 40; K += 4) {
            if (tune.has_plan && K != tune.plan_K) continue;
            const int T = G * K;
            const int npass = (Lq + T - 1) / T;
            // several passes also pay the boundary round trip through L2 every step
            // measured on config 2 (round 2, scratch/run_gpu5.sh sweeps): at the 128-register cap the deepest
            // tiles schedule worse than K = 28 (5262 GCUPS at K = 28, 5107 / 5160 at 32 / 36, 4514 at 40)
            const double deep = (K == 40) ? 1.10 : (K >= 32) ? 1.06 : 1.0;
            const double cost = stretch * deep * (double)npass * (cols_bulk + (G - 1) * n_bulk) *
                                    (C0 + (npass > 1 ? 12.0 : 0.0) + C1 * K) * (double)G +
                                mid_cost + giant_cost;
            if (cost < best_cost - 1e-9) {
                best_cost = cost;
                best.G = G; best.logG = logG; best.K = K; best.npass = npass;
                best_kg = kg; best_kmid = kmid;
            }
        }
    }
    if (tune.has_plan) best_kmid = best_kg;  // "logG,K[,threads]": tuning override, no mid tier
    if (gp) gp->kg = best_kg;
    if (kmid_out) *kmid_out = best_kmid;
    best = finish_plan(best, Lq, A, smem_limit);
    return best;
}

Plan make_forced_plan(int Lq) {
    Plan p; p.G = 32; p.logG = 5; p.K = 32;
    p = finish_plan(p, Lq, 0, 0);
    p.psmem = false; p.smem = 0;
    return p;
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

search_kernel_t pick_kernel(int K, int W, bool local, bool psmem, bool tdadd = true, bool endm = false) {
    if (W == 16) {
        if (endm) return local ? get_search_kernel_w16_l1_end(K, psmem, tdadd) : get_search_kernel_w16_l0_end(K, psmem, tdadd);
        return local ? get_search_kernel_w16_l1(K, psmem, tdadd) : get_search_kernel_w16_l0(K, psmem, tdadd);
    }
    return local ? get_search_kernel_w32_l1(K, psmem) : get_search_kernel_w32_l0(K, psmem);
}

int launch_search(pyopal_cuda_db* db, Workspace* ws, search_kernel_t fn, const KParams& P, size_t smem, int threads,
                  int64_t max_tasks, cudaStream_t st, int64_t grid_cap = 0) {
    CU(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // Shared-memory carve-out.  Left to the driver, a launch whose CTAs need more shared memory than the
    // configuration the previous kernel left behind can run with ONE resident CTA per SM where the occupancy query
    // promises two: persistent CTAs never leave the SM idle, so it is never reconfigured (measured: the 2000-aa
    // query at 3.4 instead of 5.2 TCUPS inside a batch, a 2 x 59 KB launch at a quarter of its speed).  Asking for
    // everything costs the L1 that caches the residue stream (- 2 %), so ask for what the resident CTAs need.
    CU(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, threads, smem));
    if (per_sm < 1) return fail(PYOPAL_CUDA_ERR_INTERNAL, "kernel does not fit on an SM");
    {
        const size_t need = (size_t)per_sm * (smem + 1024) + 1024;   // 1 KB per CTA is reserved by the system
        const size_t all = (size_t)db->smem_optin + 1024;
        const int pct = (int)std::min<size_t>(100, (need * 100 + all - 1) / all);
        CU(cudaFuncSetAttribute(fn, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
    }
    int64_t grid = (int64_t)db->sm_count * per_sm;
    const int64_t need = (max_tasks + (threads / 32) - 1) / (threads / 32);
    grid = std::min(grid, need);
    if (grid_cap > 0) grid = std::min(grid, grid_cap);
    grid = std::max<int64_t>(1, grid);
    fn<<<(unsigned)grid, threads, smem, st>>>(P);
    CU(cudaGetLastError());
    ws->stats[0]++;
    return 0;
}

void apply_plan(KParams& P, const Plan& pl) {
    P.G = pl.G; P.logG = pl.logG; P.npass = pl.npass; P.pitch = pl.pitch; P.padtop = pl.padtop;
    P.krows = pl.K; P.kstride = pl.kstride;
}

// layout of the per-search parameter block (one pinned host copy -> one H2D)
constexpr size_t PAR_COUNTS = 0;                  // int32[16]: [0..3] list sizes, [4..7] task counters, [8..11] giants
constexpr size_t PAR_MATRIX = 64;                 // int32[32 * 32]
constexpr size_t PAR_LISTS = PAR_MATRIX + 4096;   // work lists, fix-ups, query

}  // namespace

// Enqueue a SCORE search of [start, end) on `st`; scores land in d_out (device).  With `endp` the 16-bit
// launches also locate the end cell (END mode); targets they cannot hold are returned in endp->spill.
//
// Launches (all optional but the bulk):
//   tgiants: targets longer than one slab of tgiant.cu when the query is short enough for it, 32 bit, one CTA
//            per target, on the second high-priority stream
//   giants : the other "long" pairs chosen with the tile (make_plan), giant.cu, 32 bit, high-priority side stream
//   wide32 : 32-bit rung for pairs the 16-bit rung cannot hold exactly (NW/HW/OV static bound, or gap costs /
//            scores outside its range: then every pair), latency tile, side stream
//   mid    : 16-bit rung in its 32-lane tile for pairs too long for the narrow groups of the bulk tile, side stream
//   bulk   : 16-bit rung, throughput tile (make_plan), persistent 224/480-thread CTAs, on `st`
//   rescue : 32-bit rescore of SW targets whose 16-bit running max overflowed; its work list is produced on the
//            device by bulk/mid.
int pyopal::enqueue_score_search_ex(pyopal_cuda_db* db, Workspace* ws, const unsigned char* query, int Lq, int go,
                                    int ge, const int* M, int A, int mode, int64_t start, int64_t end,
                                    int32_t* d_out, cudaStream_t st, EndPlan* endp, SearchOpts* opts) {
    // opts (enqueue_dual_search): plan with the length of the longer query of a pair, report the plan without
    // enqueuing anything, or enqueue everything BUT the 16-bit bulk
    const bool plan_only = opts && opts->plan_only, no_bulk = opts && opts->no_bulk;
    const int LqP = (opts && opts->plan_Lq > 0) ? opts->plan_Lq : Lq;
    if (!plan_only) {
        for (int i = 0; i < 12; i++) ws->stats[i] = 0;
        ws->kev_valid[0] = ws->kev_valid[1] = ws->kev_valid[2] = false;
        ws->d_ovf_count = nullptr;
    }
    const int64_t nout = end - start;
    if (nout <= 0) return 0;
    const bool local = (mode == MODE_SW);
    const bool endm = endp != nullptr;
    const Tuning& tune = db->tune;
    if (endm) { endp->spill.clear(); endp->d_ovf = nullptr; endp->d_ovf_count = nullptr; endp->packed = false; }
    // PYOPAL_CUDA_DEBUG: synchronise after every launch and name the one that failed
    auto dbg = [&](const char* what) {
        if (!tune.debug) return;
        const cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) fprintf(stderr, "[pyopal_cuda] after %s: %s\n", what, cudaGetErrorString(e));
    };

    // a workspace is reused across streams: wait for its previous search, mark the end of this one
    if (!plan_only) {
        if (ws->used) CU(cudaStreamWaitEvent(st, ws->ev_free, 0));
        ws->used = true;
    }
    struct Mark {
        Workspace* w; cudaStream_t s;
        ~Mark() { if (w) cudaEventRecord(w->ev_free, s); }
    } mark{plan_only ? nullptr : ws, st};

    // host-side fix-ups: zero-length targets, or everything when the query is empty (score mode; the end-mode
    // caller fills these from closed forms on the host)
    std::vector<int32_t> fixes;
    if (!endm) {
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
    }
    const bool device_dp = !(Lq == 0 || db->npairs == 0);

    // ---- plans -------------------------------------------------------------
    const size_t SMEM_LIMIT = (size_t)200 * 1024;
    GiantPolicy gpol;
    int gK = 4, gwarps = 1, gteams = 1, gpitch = 0;
    size_t gsmem = 0;
    bool use_t = false, old_giants_ok = false, bfix = false, wide = false, same_plan = true, stream = false;
    bool packed_giants = false;
    int tK = 32, tNW = 4, tpitch = 0, tsplit_override = -1;
    size_t tsmem = 0;
    const int smem_optin = db->smem_optin;
    Plan bulk, tail;
    Rung rung;
    int64_t kmid = 0, kg = 0, n32p = 0, tail_end = 0;
    std::vector<int32_t> wl32, wl16, wlg, wlt;
    int64_t ntail = 0, nbulk = 0, n32 = 0, ng = 0, nt = 0;
    bool whole = false;
    if (device_dp) {
        // giants: the longest targets get a team of warps each (giant.cu), 32 bit; needs the int8
        // profile of the whole query in shared memory.
        // gap_extend > gap_open: NW's padding rows need the border fix of the 32-bit kernel (kernels.cuh, BFIX); the
        // giant kernels do not carry it, so everything stays on search_kernel then
        bfix = (mode == MODE_NW && ge > go);
        const bool fits8 = make_rung(M, A, go, ge, mode, 0).fits8;
        if (fits8 && LqP <= 8192 && !bfix && !tune.no_giant && !endm) {
            gK = (LqP <= 2048) ? 4 : (LqP <= 4096) ? 8 : 16;          // rows per lane
            gwarps = ((LqP + gK - 1) / gK + 31) / 32;                 // warps per team (<= 16): the whole query in one pass
            gteams = (gwarps == 1) ? 16 : std::min(15, 16 / gwarps);  // teams per 512-thread CTA (15 named barriers)
            gpitch = (gwarps * 32 * (gK + 4) + 15) / 16 * 16;         // gK = 4, 8, 16 words-even: chunks of gK + 4 bytes
            gsmem = giant_smem_bytes(A, gpitch, gwarps, gteams);
            gpol.ok = gsmem <= SMEM_LIMIT + 24 * 1024;
            gpol.cap = (int64_t)db->sm_count * 8;
            // one SM clock ~ 56 cost units (calibrated on config 2); a team advances one column in about
            // 36 clocks per row held by a lane, 1.6 x that when the wavefront crosses warps
            const int teams_now = (gwarps == 1) ? 16 : std::max(1, 16 / gwarps);
            gpol.col_cost = tune.gmul * 56.0 * 36.0 * gK * (gwarps > 1 ? 1.6 : 1.0) / teams_now;
        }
        old_giants_ok = gpol.ok;
        // transposed giants (tgiant.cu): the target over the lanes, the query streamed; needs the short F chain
        if (fits8 && ge <= go && !tune.no_tgiant && !tune.no_giant && !endm) {
            if (tune.has_tgiant) { tNW = tune.tg_warps; tK = tune.tg_K; tsplit_override = tune.tg_split; }
            tpitch = tNW * 32 * (tK + 4);
            tsmem = tgiant_smem_bytes(A, tpitch, tNW, LqP);
            // one warp per scheduler issues a 32-row step in ~820 clocks: faster than giant.cu up to ~1500 rows
            use_t = tsmem + 1024 <= (size_t)smem_optin && (LqP <= 1536 || tune.has_tgiant);
            if (use_t && !old_giants_ok) {
                gpol.ok = true;
                gpol.cap = (int64_t)db->sm_count * 8;
                // a revolution of Lq + 32 * warps - 1 steps covers 32 * warps * K target columns; a step costs
                // about 25 clocks per row held by a lane (measured); 56 cost units per SM clock
                const double step_clk = 25.0 * tK + 20.0;
                gpol.col_cost = 56.0 * step_clk * (double)(LqP + tNW * 32 - 1) / (double)(tNW * 32 * tK);
            }
        }
        if (endm && fits8 && !tune.no_giant) {
            // end mode: the "giants" are walked by pairdp.cu's band teams; it takes what the 32-lane tile cannot
            gpol.ok = true;
            gpol.cap = (int64_t)db->sm_count * 8;
            gpol.col_cost = 56.0 * 40.0 * (double)std::max(1, (Lq + 127) / 128);
        }
        // whole-database score searches run the bulk as one continuous wavefront per group (stream.cuh)
        stream = !endm && !tune.no_stream && start == 0 && end >= db->n && fits8 && ge <= go && !bfix &&
                 db->ncols < ((int64_t)1 << 31);
        const bool allow_mid = tune.mid > 0 && !tune.no_tail && !tune.has_plan;  // opt-in (PYOPAL_CUDA_MID): see DESIGN.md 4.1
        bulk = make_plan(LqP, A, db, SMEM_LIMIT, &gpol, allow_mid, &kmid);
        // ... and only where it pays: wide groups (the fill steps it removes are (G - 1) per pair and pass) of deep
        // tiles (its longer step prologue is amortised over K rows).  Measured on a 1/8 shard of config 2: + 5.5 % at
        // G = 32, K = 32; - 12 % at G = 16, K = 20.  PYOPAL_CUDA_STREAM=1 forces it (tests).
        if (stream && !tune.force_stream && !(bulk.G >= 16 && bulk.K >= 28 && bulk.npass <= 2)) stream = false;
        tail = make_tail_plan(Lq, A, SMEM_LIMIT);
        int smax = INT_MIN, smin = INT_MAX;
        for (int i = 0; i < A * A; i++) { smax = std::max(smax, M[i]); smin = std::min(smin, M[i]); }
        const int Lp = std::max(bulk.npass * bulk.G * bulk.K, tail.npass * tail.G * tail.K);
        rung = make_rung(M, A, go, ge, mode, Lp);
        wide = !rung.fits8;
        if (wide || bfix || !bulk.psmem || !tail.psmem) {
            // profile in global memory: one tile shape (32 x 32) for every launch
            bulk = make_forced_plan(Lq);
            tail = bulk;
            stream = false;
        }
        {
            // 32-bit overflow cannot happen below this bound (scores are exact otherwise)
            int64_t mag = std::max<int64_t>({(int64_t)go, (int64_t)ge, 1, std::llabs((long long)smax), std::llabs((long long)smin)});
            const int64_t longest = db->h_pair_len[0];
            if (mag * ((int64_t)Lp + longest + 8) >= (1LL << 29))
                return fail(OPAL_ERR_OVERFLOW, "scores may exceed the 32-bit range");
        }
        same_plan = (bulk.G == tail.G && bulk.K == tail.K && bulk.npass == tail.npass) || tune.no_tail;
        if (tune.no_tail) tail = bulk;

        // ---- split the (length-sorted) pairs ------------------------------------
        // [0, kg): giants; [kg, n32p): 32-bit static; [n32p, tail_end): 16-bit mid tier; rest: 16-bit bulk
        if (!rung.ok16) { n32p = db->npairs; stream = false; }
        else if (!local) n32p = first_pair_at_most(db, (double)rung.max_cols16, 0);
        const bool giants_ok = gpol.ok && bulk.psmem && !wide;
        kg = giants_ok ? gpol.kg : 0;
        n32p = std::max(n32p, kg);
        tail_end = n32p;
        if (!same_plan) {
            if (giants_ok) {
                tail_end = std::max(n32p, kmid);
            } else {
                // no giant kernel for this query: a pair leaves the bulk kernel when walking it alone in the bulk tile
                // would take more than a quarter of the whole search at full machine throughput
                const double C0 = 28.0, C1 = 6.5;
                const double total_instr = (double)db->ncols * bulk.npass * bulk.G * (C0 + C1 * bulk.K);
                const double ideal_cycles = total_instr / ((double)db->sm_count * 64.0);
                const double step_latency = 12.0 * bulk.K + 80.0;
                const double thr = std::max(256.0, ideal_cycles / 4.0 / (bulk.npass * step_latency));
                tail_end = first_pair_at_most(db, thr, n32p);
            }
        }
        whole = (start == 0 && end >= db->n);
        auto in_range = [&](int32_t idx) { return idx >= start && idx < end; };
        auto pair_in_range = [&](int64_t k) {
            return in_range(db->h_pair_idx[(size_t)(2 * k)]) ||
                   (db->h_pair_idx[(size_t)(2 * k + 1)] >= 0 && in_range(db->h_pair_idx[(size_t)(2 * k + 1)]));
        };
        // giants longer than one slab of the transposed kernel go to tgiant.cu, the others to giant.cu
        const int tsplit = !use_t ? INT32_MAX : tsplit_override >= 0 ? tsplit_override : tNW * 32 * tK;
        // Smith-Waterman: giant.cu walks the two targets of a pair together at 16 bit (its list holds pair ids then)
        packed_giants = local && !endm && old_giants_ok && rung.ok16 && !tune.no_giant16;
        for (int64_t k = 0; k < kg; k++) {
            if (packed_giants && db->h_pair_len[(size_t)(2 * k)] <= tsplit) {
                if (pair_in_range(k)) wlg.push_back((int32_t)k);
                continue;
            }
            for (int h = 0; h < 2; h++) {
                const int32_t ix = db->h_pair_idx[(size_t)(2 * k + h)];
                if (ix < 0 || !in_range(ix)) continue;
                const int32_t e = (int32_t)((k << 1) | h);
                if (endm) endp->spill.push_back(e);
                else if (db->h_pair_len[(size_t)(2 * k + h)] > tsplit || !old_giants_ok || packed_giants) wlt.push_back(e);
                else wlg.push_back(e);
            }
        }
        for (int64_t k = kg; k < n32p; k++) {
            for (int h = 0; h < 2; h++) {
                const int32_t ix = db->h_pair_idx[(size_t)(2 * k + h)];
                if (ix < 0 || !in_range(ix)) continue;
                if (endm) endp->spill.push_back((int32_t)((k << 1) | h));
                else wl32.push_back((int32_t)((k << 1) | h));
            }
        }
        if (!whole) {
            for (int64_t k = n32p; k < tail_end; k++) if (pair_in_range(k)) wl16.push_back((int32_t)k);
            ntail = (int64_t)wl16.size();
            for (int64_t k = tail_end; k < db->npairs; k++) if (pair_in_range(k)) wl16.push_back((int32_t)k);
            nbulk = (int64_t)wl16.size() - ntail;
        } else {
            ntail = tail_end - n32p;
            nbulk = db->npairs - tail_end;
        }
        n32 = (int64_t)wl32.size();
        ng = (int64_t)wlg.size();
        nt = (int64_t)wlt.size();
        wl32.insert(wl32.end(), wlg.begin(), wlg.end());  // one list: [static 32-bit | giants | transposed giants]
        wl32.insert(wl32.end(), wlt.begin(), wlt.end());
    }
    const int64_t n16_all = ntail + nbulk;
    if (same_plan) { nbulk = n16_all; ntail = 0; }
    if (opts) {
        opts->bulk_G = bulk.G; opts->bulk_logG = bulk.logG; opts->bulk_K = bulk.K;
        opts->first_bulk = same_plan ? n32p : tail_end;
        // "clean": everything below the giants is one 16-bit bulk launch over a contiguous range of pairs
        opts->clean = device_dp && whole && local && !endm && rung.ok16 && !wide && !bfix && bulk.psmem && ntail == 0 &&
                      n32 == 0 && (same_plan ? n32p : tail_end) == kg;
    }
    if (plan_only) return 0;
    if (no_bulk) { nbulk = 0; ntail = 0; }
    const int64_t n16 = no_bulk ? 0 : n16_all;

    // ---- the parameter block: counters, matrix, lists, fix-ups, query: one pinned copy, one H2D ------------
    const size_t off_wl32 = PAR_LISTS;
    const size_t off_wl16 = off_wl32 + ((wl32.size() * 4 + 15) & ~(size_t)15);
    const size_t off_fix = off_wl16 + ((wl16.size() * 4 + 15) & ~(size_t)15);
    const size_t off_query = off_fix + ((fixes.size() * 4 + 15) & ~(size_t)15);
    const size_t par_bytes = off_query + (((size_t)Lq + 15) & ~(size_t)15) + 16;
    const int slot = ws->par_next;
    ws->par_next = (slot + 1) % Workspace::NPARAM;
    if (ws->par_used[slot]) CU(cudaEventSynchronize(ws->par_done[slot]));
    if (ws->h_par[slot].reserve(par_bytes) || ws->d_par.reserve(par_bytes)) return PYOPAL_CUDA_ERR_NOMEM;
    unsigned char* hp = ws->h_par[slot].p;
    unsigned char* dp = ws->d_par.p;
    {
        int32_t* counts = reinterpret_cast<int32_t*>(hp + PAR_COUNTS);
        memset(counts, 0, 64);
        counts[0] = (int32_t)nbulk; counts[1] = (int32_t)n32; counts[2] = 0; counts[3] = (int32_t)ntail;
        counts[8] = (int32_t)ng; counts[10] = (int32_t)nt;
        memcpy(hp + PAR_MATRIX, M, (size_t)A * A * sizeof(int32_t));
        if (!wl32.empty()) memcpy(hp + off_wl32, wl32.data(), wl32.size() * 4);
        if (!wl16.empty()) memcpy(hp + off_wl16, wl16.data(), wl16.size() * 4);
        if (!fixes.empty()) memcpy(hp + off_fix, fixes.data(), fixes.size() * 4);
        if (Lq > 0) memcpy(hp + off_query, query, (size_t)Lq);
    }
    CU(cudaMemcpyAsync(dp, hp, par_bytes, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(ws->par_done[slot], st));
    ws->par_used[slot] = true;
    dbg("parameter block");
    if (endm && endp->launch_spill && !endp->spill.empty()) {
        // the targets left to pairdp.cu go first, on their own high-priority stream: launched behind the persistent
        // CTAs of the 16-bit kernels they would only start when those drain
        int rc = endp->launch_spill();
        if (rc) return rc;
    }
    int32_t* d_counts = reinterpret_cast<int32_t*>(dp + PAR_COUNTS);
    const int32_t* d_wl32 = reinterpret_cast<const int32_t*>(dp + off_wl32);
    const int32_t* d_wl16 = reinterpret_cast<const int32_t*>(dp + off_wl16);

    auto apply_fixes = [&]() -> int {
        if (fixes.empty() || no_bulk) return 0;   // no_bulk: the dual search writes them, once per query
        const int nfix = (int)(fixes.size() / 2);
        scatter_fixes_kernel<<<(nfix + 255) / 256, 256, 0, st>>>(reinterpret_cast<const int32_t*>(dp + off_fix), nfix, d_out);
        CU(cudaGetLastError());
        return 0;
    };
    if (!device_dp) return apply_fixes();

    const int32_t* d_wl_tail = whole ? db->d_identity + n32p : d_wl16;
    const int32_t* d_wl_bulk = whole ? db->d_identity + tail_end : d_wl16 + (same_plan ? 0 : ntail);
    if (same_plan) d_wl_bulk = d_wl_tail;
    const int64_t n16x = n16 + (packed_giants ? ng : 0);   // pairs scored at 16 bit: bulk, mid and packed giants
    if (local && n16x && ws->ovf.reserve((size_t)(2 * n16x))) return PYOPAL_CUDA_ERR_NOMEM;
    if ((bulk.npass > 1 && !no_bulk) || tail.npass > 1) {
        if (ws->bndH.reserve((size_t)db->ncols + 16) || ws->bndF.reserve((size_t)db->ncols + 16)) return PYOPAL_CUDA_ERR_NOMEM;
        if (!endm && (n32 > 0 || local))
            if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                return PYOPAL_CUDA_ERR_NOMEM;
    }

    KParams P;
    memset(&P, 0, sizeof(P));
    P.stream = db->d_stream; P.pair_off = db->d_pair_off; P.pair_len = db->d_pair_len; P.pair_idx = db->d_pair_idx;
    P.query = dp + off_query; P.matrix = reinterpret_cast<const int32_t*>(dp + PAR_MATRIX);
    P.Lq = Lq; P.A = A; P.go = go; P.ge = ge;
    P.mode = mode; P.pad_score = rung.pad_score; P.ovf_thresh = rung.ovf_thresh;
    P.one = 1u; P.bias = 0;
    P.out = d_out; P.out_start = start; P.out_end = end;
    if (endm) { P.out_q = endp->d_endq; P.out_t = endp->d_endt; }
    P.ovf_list = ws->ovf.p; P.ovf_count = d_counts + 2;
    ws->d_ovf_count = (local && n16x > 0) ? d_counts + 2 : nullptr;
    P.bndH = ws->bndH.p; P.bndF = ws->bndF.p;
    if (endm) { endp->d_ovf = ws->ovf.p; endp->d_ovf_count = d_counts + 2; endp->packed = n16 > 0; }

    search_kernel_t k32 = nullptr;
    if (wide) {
        apply_plan(P, tail);
        if (ws->wprofile.reserve((size_t)(A + 1) * tail.pitch)) return PYOPAL_CUDA_ERR_NOMEM;
        P.wprofile = ws->wprofile.p;
        build_profile_kernel<<<64, 256, 0, st>>>(P, nullptr, ws->wprofile.p);
        CU(cudaGetLastError());
        k32 = bfix ? get_search_kernel_bfix(true) : get_search_kernel_wide(local);
    } else {
        if (!tail.psmem) {
            apply_plan(P, tail);
            if (ws->gprofile.reserve((size_t)(A + 1) * tail.pitch)) return PYOPAL_CUDA_ERR_NOMEM;
            P.gprofile = ws->gprofile.p;
            build_profile_kernel<<<64, 256, 0, st>>>(P, ws->gprofile.p, nullptr);
            CU(cudaGetLastError());
        }
        k32 = bfix ? get_search_kernel_bfix(false) : pick_kernel(tail.K, 32, local, tail.psmem);
    }
    if (!k32) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 32-bit kernel for this tile");

    // ---- side stream: longest targets first, concurrently with the bulk -------
    const bool mid_work = ntail > 0 && !same_plan;
    const bool side_work = (ng > 0) || (nt > 0) || (n32 > 0) || mid_work;
    cudaStream_t side = ws->side;
    if (side_work) {
        CU(cudaEventRecord(ws->ev_fork, st));
        CU(cudaStreamWaitEvent(side, ws->ev_fork, 0));
        CU(cudaEventRecord(ws->kev[0], side));
    }
    if (nt > 0) {
        search_kernel_t kt = get_tgiant_kernel(tK, local);
        if (!kt) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no transposed giant kernel for this tile");
        KParams Q = P;
        Q.G = tNW; Q.logG = 0; Q.npass = 1; Q.pitch = tpitch; Q.padtop = 0;
        Q.krows = tK; Q.kstride = tK + 4;
        // padding rows above target position 0 reproduce the border H[i][-1] (kernels.cuh, profile_entry)
        Q.pad_score = ((mode == MODE_NW || mode == MODE_HW) ? -std::max(go, ge) : 0) + go;
        Q.worklist = d_wl32 + n32 + ng; Q.nwork = d_counts + 10; Q.counter = d_counts + 11;
        Q.neg = -(1 << 29);
        // the whole SM's shared memory: no throughput CTA shares the SM with a giant (giant.cu)
        const size_t smem_all = std::max<size_t>(tsmem, (size_t)smem_optin - 1024);
        CU(cudaFuncSetAttribute(kt, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_all));
        // its own stream: the few longest targets start at once, beside giant.cu's medium ones
        CU(cudaStreamWaitEvent(ws->side2, ws->ev_fork, 0));
        kt<<<(unsigned)std::min<int64_t>(nt, db->sm_count), tNW * 32, smem_all, ws->side2>>>(Q);
        CU(cudaGetLastError());
        CU(cudaEventRecord(ws->ev_join2, ws->side2));
        ws->stats[0]++;
        dbg("tgiants");
    }
    if (ng > 0) {
        search_kernel_t kgiant = packed_giants ? get_giant_kernel_packed(gK, rung.tdadd) : get_giant_kernel(gK, local);
        if (!kgiant) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no giant kernel for this tile");
        KParams Q = P;
        // a few teams per CTA keep the column chains short; more CTAs than that would only idle SMs
        const int teams = (int)std::max<int64_t>(1, std::min<int64_t>(gteams, (ng + 7) / 8));
        Q.G = gwarps; Q.logG = 0; Q.npass = teams; Q.pitch = gpitch; Q.padtop = gwarps * 32 * gK - Lq;
        Q.krows = gK; Q.kstride = (gK == 4) ? 4 : gK + 4;
        Q.worklist = d_wl32 + n32; Q.nwork = d_counts + 8; Q.counter = d_counts + 9;
        Q.neg = -(1 << 29);
        if (packed_giants) { Q.neg = rung.neg16; Q.bias = rung.bias16; }
        // ask for the whole SM's shared memory: no throughput CTA may share the SM with the giants (giant.cu)
        const size_t gsmem_all = std::max<size_t>(gsmem, (size_t)smem_optin - 1024);
        CU(cudaFuncSetAttribute(kgiant, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem_all));
        // with the longest targets on tgiant.cu, what is left here is many medium targets: a quarter of the
        // SMs is enough for their latency and leaves the rest to the bulk kernel from the start
        const int64_t gcap = use_t ? std::max<int64_t>(8, db->sm_count / 4) : db->sm_count;
        kgiant<<<(unsigned)std::min<int64_t>((ng + teams - 1) / teams, gcap), 512, gsmem_all, side>>>(Q);
        CU(cudaGetLastError());
        ws->stats[0]++;
        dbg("giants");
    }
    if (n32 > 0) {
        KParams Q = P;
        apply_plan(Q, tail);
        Q.worklist = d_wl32; Q.nwork = d_counts + 1; Q.counter = d_counts + 5;
        Q.neg = -(1 << 29);
        Q.bndH = ws->bnd32H.p; Q.bndF = ws->bnd32F.p;
        int rc = launch_search(db, ws, k32, Q, tail.smem, 32, n32, side);
        if (rc) return rc;
        dbg("static32");
    }
    if (mid_work) {
        search_kernel_t kt = pick_kernel(tail.K, 16, local, tail.psmem, rung.tdadd, endm);
        if (!kt) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 16-bit kernel for the mid tile");
        KParams Q = P;
        apply_plan(Q, tail);
        Q.worklist = d_wl_tail; Q.nwork = d_counts + 3; Q.counter = d_counts + 7;
        Q.neg = rung.neg16; Q.bias = rung.bias16;
        // CTAs of four warps share one copy of the profile; as many of them as the mid tier's share of the
        // columns is of the machine's resident warps, so that both tiers finish together
        const double share = (double)((tail_end < db->npairs ? db->h_pair_off[(size_t)tail_end] : db->ncols) -
                                      (n32p < db->npairs ? db->h_pair_off[(size_t)n32p] : db->ncols)) /
                             std::max<double>(1.0, (double)db->ncols);
        const int mthreads = ntail >= 4 * (int64_t)db->sm_count ? 128 : 32;
        const int64_t cap = std::max<int64_t>(db->sm_count / 4,
                                              (int64_t)(1.15 * share * db->sm_count * 14.0 / (mthreads / 32) + 0.999));
        int rc = launch_search(db, ws, kt, Q, tail.smem, mthreads, ntail, side, cap);
        if (rc) return rc;
        dbg("mid");
    }
    if (side_work) {
        if (nt > 0) CU(cudaStreamWaitEvent(side, ws->ev_join2, 0));  // joins only after everything is launched
        CU(cudaEventRecord(ws->kev[1], side));
        ws->kev_valid[0] = true;
        CU(cudaEventRecord(ws->ev_join, side));
    }

    // ---- bulk -------------------------------------------------------------------
    if (nbulk > 0) {
        search_kernel_t kb = pick_kernel(bulk.K, 16, local, bulk.psmem, rung.tdadd, endm);
        if (stream && whole) {
            search_kernel_t ks = local ? get_stream_kernel_l1(bulk.K, rung.tdadd) : get_stream_kernel_l0(bulk.K, rung.tdadd);
            if (ks) kb = ks; else stream = false;
        } else stream = false;
        if (!kb) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 16-bit kernel for this tile");
        KParams Q = P;
        apply_plan(Q, bulk);
        Q.worklist = d_wl_bulk; Q.nwork = d_counts + 0; Q.counter = d_counts + 4;
        Q.first_pair = (int)(same_plan ? n32p : tail_end);
        Q.neg = rung.neg16; Q.bias = rung.bias16;
        const int gpw = 32 >> bulk.logG;
        // One CTA per SM that fills the register file (16 warps x 32 lanes x 128 registers): one copy of the profile
        // per SM and one shared-memory configuration whatever the query length (two 7- or 8-warp CTAs were 0-5 %
        // slower, and see launch_search on carve-outs); 15 warps when one-warp CTAs of the 32-bit rung or the mid
        // tier run beside it, so that they find registers on every SM.  The giants take whole SMs either way.
        int threads = (n32 > 0 || mid_work) ? 480 : 512;
        if (tune.plan_threads) threads = tune.plan_threads;
        CU(cudaEventRecord(ws->kev[2], st));
        if (tune.debug)
            fprintf(stderr, "[pyopal_cuda] Lq=%d %s%s bulk: G=%d K=%d npass=%d pairs=%lld | mid=%lld (K=%d npass=%d) giants=%lld tgiants=%lld static32=%lld spill=%lld threads=%d smem=%zu\n",
                    Lq, endm ? "END" : "score", stream ? " (stream)" : "", bulk.G, bulk.K, bulk.npass, (long long)nbulk, (long long)(mid_work ? ntail : 0), tail.K,
                    tail.npass, (long long)ng, (long long)nt, (long long)n32, (long long)(endm ? endp->spill.size() : 0), threads, bulk.smem);
        int rc = launch_search(db, ws, kb, Q, bulk.smem, threads, (nbulk + gpw - 1) / gpw, st);
        if (rc) return rc;
        CU(cudaEventRecord(ws->kev[3], st));
        ws->kev_valid[1] = true;
        dbg("bulk");
    }
    if (side_work) CU(cudaStreamWaitEvent(st, ws->ev_join, 0));

    // ---- rescue: 32-bit rescore of what overflowed (list and size live on the device)
    if (local && n16x > 0 && !endm) {
        KParams R = P;
        apply_plan(R, tail);
        R.worklist = ws->ovf.p; R.nwork = d_counts + 2; R.counter = d_counts + 6;
        R.neg = -(1 << 29);
        R.bndH = ws->bnd32H.p; R.bndF = ws->bnd32F.p;
        CU(cudaEventRecord(ws->kev[4], st));
        int rc = launch_search(db, ws, k32, R, tail.smem, 32, std::min<int64_t>(2 * n16x, (int64_t)db->sm_count * 4), st);
        if (rc) return rc;
        CU(cudaEventRecord(ws->kev[5], st));
        ws->kev_valid[2] = true;
        dbg("rescue");
    }
    ws->stats[1] = n16 * 2; ws->stats[2] = n32 + nt + (packed_giants ? 2 * ng : ng);   // targets on the 32-bit rung / the giant kernels
    ws->stats[5] = bulk.G; ws->stats[6] = bulk.K; ws->stats[7] = bulk.npass; ws->stats[11] = bulk.psmem ? 1 : 0;
    return apply_fixes();
}

// Two Smith-Waterman score searches of the WHOLE database with ONE pass of dual_kernel (dual.cuh) over everything
// below the giants: the two queries share every target column, so their folded scores sit pre-packed in one profile
// word and a cell pair costs 4.5 ALU-pipe instructions instead of 5.5.  The launch plan is the single-query one for
// the LONGER query (same tile, same giants); each query's giants run through the single-query path with the bulk
// switched off, on two more workspaces so that they overlap each other and the dual launch.  Returns
// DUAL_NOT_ELIGIBLE when the pair has to be searched one query after the other (nothing has been enqueued then):
// profile words that do not fit shared memory, a plan with more than giants + one 16-bit bulk launch, no spare
// workspace.
int pyopal::enqueue_dual_search(pyopal_cuda_db* db, Workspace* ws, const unsigned char* qa, int La, const unsigned char* qb,
                                int Lb, int go, int ge, const int* M, int A, int32_t* d_out_a, int32_t* d_out_b,
                                cudaStream_t st) {
    const Tuning& tune = db->tune;
    if (tune.no_dual || La <= 0 || Lb <= 0 || db->npairs == 0 || ge > go) return DUAL_NOT_ELIGIBLE;
    if (tune.has_tgiant || tune.force_stream) return DUAL_NOT_ELIGIBLE;   // a test asked for another kernel
    const int Lp = std::max(La, Lb);
    const size_t SMEM_LIMIT = (size_t)200 * 1024;
    // the single-query plan of the longer query: tile, giants
    SearchOpts po;
    po.plan_only = true; po.plan_Lq = Lp;
    if (enqueue_score_search_ex(db, ws, qa, La, go, ge, M, A, MODE_SW, 0, db->n, d_out_a, st, nullptr, &po) || !po.clean)
        return DUAL_NOT_ELIGIBLE;
    Plan pl;
    pl.G = po.bulk_G; pl.logG = po.bulk_logG; pl.K = po.bulk_K;
    const int64_t first = po.first_bulk;
    if (first >= db->npairs) return DUAL_NOT_ELIGIBLE;   // nothing but giants
    // small databases are latency-bound by their giants: two queries per pass only pays when the bulk dominates
    if (first > 0 && !tune.force_dual && db->ncols < (int64_t)4 << 20) return DUAL_NOT_ELIGIBLE;
    const int T = pl.G * pl.K;
    const int npass = (Lp + T - 1) / T;
    Rung rung = make_rung(M, A, go, ge, MODE_SW, npass * T);
    if (!rung.ok16) return DUAL_NOT_ELIGIBLE;
    // Profile of 32-bit words, read with LDS.128: the hardware serves a quarter warp (8 lanes x 16 bytes) per phase,
    // conflict-free when the eight 16-byte units fall into the eight different bank groups.  A lane's chunk is K words
    // padded to an ODD number of units, so the lanes p = 0..7 of a group sit in eight different groups; every letter
    // row starts at a multiple of 128 bytes, so the letter (different in every lane) does not move them.  With G = 4 a
    // quarter warp holds two groups, always in the same pass: half a row of skew per letter (64 bytes) separates them
    // whenever their letters differ in parity.
    const int kstride = ((pl.K / 4) % 2 == 1) ? pl.K : pl.K + 4;
    int pitch = (npass * pl.G * kstride + 31) / 32 * 32;
    if (pl.G == 4) pitch += 16;
    const size_t smem = (size_t)(A + 1) * (size_t)pitch * sizeof(uint32_t);
    if (smem > SMEM_LIMIT) return DUAL_NOT_ELIGIBLE;
    search_kernel_t kd = get_dual_kernel(pl.K, rung.tdadd);
    if (!kd) return DUAL_NOT_ELIGIBLE;
    Plan tail_a = make_tail_plan(La, A, SMEM_LIMIT), tail_b = make_tail_plan(Lb, A, SMEM_LIMIT);
    if (!tail_a.psmem || !tail_b.psmem) return DUAL_NOT_ELIGIBLE;
    {
        int smax = INT_MIN, smin = INT_MAX;
        for (int i = 0; i < A * A; i++) { smax = std::max(smax, M[i]); smin = std::min(smin, M[i]); }
        const int64_t mag = std::max<int64_t>({(int64_t)go, (int64_t)ge, 1, std::llabs((long long)smax), std::llabs((long long)smin)});
        if (mag * ((int64_t)std::max(tail_a.npass * tail_a.G * tail_a.K, tail_b.npass * tail_b.G * tail_b.K) + db->h_pair_len[0] + 8) >= (1LL << 29))
            return DUAL_NOT_ELIGIBLE;   // the single-query path reports OPAL_ERR_OVERFLOW
    }

    // the giants of either query: two more workspaces (never waited for: a pool that is busy means "not eligible")
    struct TryLease {
        pyopal_cuda_db* d; Workspace* w;
        ~TryLease() { if (w) d->release(w); }
    } ga{db, nullptr}, gb{db, nullptr};
    if (first > 0) {
        ga.w = db->try_acquire();
        gb.w = db->try_acquire();
        if (!ga.w || !gb.w) return DUAL_NOT_ELIGIBLE;
    }

    for (int i = 0; i < 12; i++) ws->stats[i] = 0;
    ws->kev_valid[0] = ws->kev_valid[1] = ws->kev_valid[2] = false;
    ws->d_ovf_count = nullptr;
    if (ws->used) CU(cudaStreamWaitEvent(st, ws->ev_free, 0));
    ws->used = true;
    struct Mark {
        Workspace* w; cudaStream_t s;
        ~Mark() { cudaEventRecord(w->ev_free, s); }
    } mark{ws, st};

    int64_t giant_launches = 0;
    if (first > 0) {
        // fork: the giants start where `st` stands (the result rows are free), on their own streams
        CU(cudaEventRecord(ws->ev_aux0, st));
        SearchOpts ob;
        ob.no_bulk = true; ob.plan_Lq = Lp;
        struct Half { Workspace* w; const unsigned char* q; int L; int32_t* out; };
        const Half halves[2] = {{ga.w, qa, La, d_out_a}, {gb.w, qb, Lb, d_out_b}};
        for (const Half& h : halves) {
            CU(cudaStreamWaitEvent(h.w->stream, ws->ev_aux0, 0));
            int rc = enqueue_score_search_ex(db, h.w, h.q, h.L, go, ge, M, A, MODE_SW, 0, db->n, h.out, h.w->stream, nullptr, &ob);
            if (rc) return rc;
            giant_launches += h.w->stats[0];
            CU(cudaEventRecord(h.w->ev_aux1, h.w->stream));
        }
    }

    // zero-length targets score 0 against either query
    std::vector<int32_t> fixes;
    for (int32_t z : db->h_zero_idx) { fixes.push_back(z); fixes.push_back(0); }

    // ---- parameter block: counters, matrix, fix-ups, the two queries ----------------------------------------
    const size_t off_fix = PAR_LISTS;
    const size_t off_qa = off_fix + ((fixes.size() * 4 + 15) & ~(size_t)15);
    const size_t off_qb = off_qa + (((size_t)La + 15) & ~(size_t)15) + 16;
    const size_t par_bytes = off_qb + (((size_t)Lb + 15) & ~(size_t)15) + 16;
    const int slot = ws->par_next;
    ws->par_next = (slot + 1) % Workspace::NPARAM;
    if (ws->par_used[slot]) CU(cudaEventSynchronize(ws->par_done[slot]));
    if (ws->h_par[slot].reserve(par_bytes) || ws->d_par.reserve(par_bytes)) return PYOPAL_CUDA_ERR_NOMEM;
    unsigned char* hp = ws->h_par[slot].p;
    unsigned char* dp = ws->d_par.p;
    const int64_t ntargets = 2 * (db->npairs - first);
    {
        int32_t* counts = reinterpret_cast<int32_t*>(hp + PAR_COUNTS);
        memset(counts, 0, 64);
        counts[0] = (int32_t)ntargets;   // [4] task counter; [2] / [12] overflow counts of A / B; [6] / [13] rescue counters
        memcpy(hp + PAR_MATRIX, M, (size_t)A * A * sizeof(int32_t));
        if (!fixes.empty()) memcpy(hp + off_fix, fixes.data(), fixes.size() * 4);
        memcpy(hp + off_qa, qa, (size_t)La);
        memcpy(hp + off_qb, qb, (size_t)Lb);
    }
    const size_t n_ovf = (size_t)ntargets;
    if (ws->ovf.reserve(n_ovf) || ws->ovf2.reserve(n_ovf)) return PYOPAL_CUDA_ERR_NOMEM;
    if (npass > 1 || tail_a.npass > 1 || tail_b.npass > 1)   // pass boundaries, one entry per target and column
        if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
            return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaMemcpyAsync(dp, hp, par_bytes, cudaMemcpyHostToDevice, st));
    CU(cudaEventRecord(ws->par_done[slot], st));
    ws->par_used[slot] = true;
    int32_t* d_counts = reinterpret_cast<int32_t*>(dp + PAR_COUNTS);

    KParams P;
    memset(&P, 0, sizeof(P));
    P.stream = db->d_stream; P.pair_off = db->d_pair_off; P.pair_len = db->d_pair_len; P.pair_idx = db->d_pair_idx;
    P.matrix = reinterpret_cast<const int32_t*>(dp + PAR_MATRIX);
    P.A = A; P.go = go; P.ge = ge; P.mode = MODE_SW;
    P.pad_score = rung.pad_score; P.ovf_thresh = rung.ovf_thresh;
    P.one = 1u;
    P.out_start = 0; P.out_end = db->n;

    // ---- the dual launch ---------------------------------------------------------------------------------------
    {
        KParams Q = P;
        Q.query = dp + off_qa; Q.Lq = La; Q.padtop = npass * T - La;
        Q.query2 = dp + off_qb; Q.Lq2 = Lb; Q.padtop2 = npass * T - Lb;
        Q.G = pl.G; Q.logG = pl.logG; Q.npass = npass; Q.pitch = pitch; Q.krows = pl.K; Q.kstride = kstride;
        Q.neg = rung.neg16; Q.bias = rung.bias16;
        Q.first_pair = (int)first;
        Q.nwork = d_counts + 0; Q.counter = d_counts + 4;
        Q.out = d_out_a; Q.out2 = d_out_b;
        Q.ovf_list = ws->ovf.p; Q.ovf_count = d_counts + 2;
        Q.ovf_list2 = ws->ovf2.p; Q.ovf_count2 = d_counts + 12;
        Q.bndH = ws->bnd32H.p; Q.bndF = ws->bnd32F.p;
        const int gpw = 32 >> pl.logG;
        if (tune.debug)
            fprintf(stderr, "[pyopal_cuda] Lq=%d+%d score (dual) bulk: G=%d K=%d npass=%d targets=%lld giant pairs=%lld threads=512 smem=%zu\n",
                    La, Lb, pl.G, pl.K, npass, (long long)ntargets, (long long)first, smem);
        CU(cudaEventRecord(ws->kev[2], st));
        int rc = launch_search(db, ws, kd, Q, smem, 512, (ntargets + gpw - 1) / gpw, st);
        if (rc) return rc;
        CU(cudaEventRecord(ws->kev[3], st));
        ws->kev_valid[1] = true;
    }
    // ---- rescue of either query's 16-bit overflows at 32 bit (lists and sizes live on the device) --------------
    struct Side { const Plan* tail; size_t off_q; int Lq; int32_t* out; int32_t* list; int count, counter; };
    const Side sides[2] = {{&tail_a, off_qa, La, d_out_a, ws->ovf.p, 2, 6}, {&tail_b, off_qb, Lb, d_out_b, ws->ovf2.p, 12, 13}};
    for (const Side& sd : sides) {
        search_kernel_t k32 = pick_kernel(sd.tail->K, 32, true, true);
        if (!k32) return fail(PYOPAL_CUDA_ERR_INTERNAL, "no 32-bit kernel for this tile");
        KParams R = P;
        apply_plan(R, *sd.tail);
        R.query = dp + sd.off_q; R.Lq = sd.Lq;
        R.worklist = sd.list; R.nwork = d_counts + sd.count; R.counter = d_counts + sd.counter;
        R.neg = -(1 << 29); R.bias = 0;
        R.out = sd.out;
        R.bndH = ws->bnd32H.p; R.bndF = ws->bnd32F.p;
        int rc = launch_search(db, ws, k32, R, sd.tail->smem, 32, std::min<int64_t>(ntargets, (int64_t)db->sm_count * 4), st);
        if (rc) return rc;
    }
    if (!fixes.empty()) {
        const int nfix = (int)(fixes.size() / 2);
        for (int32_t* out : {d_out_a, d_out_b}) {
            scatter_fixes_kernel<<<(nfix + 255) / 256, 256, 0, st>>>(reinterpret_cast<const int32_t*>(dp + off_fix), nfix, out);
            CU(cudaGetLastError());
        }
        ws->stats[0] += 2;
    }
    if (first > 0) {   // join: `st` is the stream the caller orders everything else after
        CU(cudaStreamWaitEvent(st, ga.w->ev_aux1, 0));
        CU(cudaStreamWaitEvent(st, gb.w->ev_aux1, 0));
        ws->stats[0] += giant_launches;
    }
    ws->d_ovf_count = d_counts + 2;
    ws->stats[1] = ntargets * 2;
    ws->stats[5] = pl.G; ws->stats[6] = pl.K; ws->stats[7] = npass; ws->stats[11] = 2;   // 2: profile in shared memory, two queries per pass
    return 0;
}

int pyopal::enqueue_score_search(pyopal_cuda_db* db, Workspace* ws, const unsigned char* query, int Lq, int go, int ge,
                                 const int* M, int A, int mode, int64_t start, int64_t end, int32_t* d_out,
                                 cudaStream_t st) {
    return enqueue_score_search_ex(db, ws, query, Lq, go, ge, M, A, mode, start, end, d_out, st, nullptr);
}

int pyopal::check_search_args(pyopal_cuda_db* db, const unsigned char* query, int Lq, int go, int ge, const int* M,
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


// ---------------------------------------------------------------------------
// Public search entry points

extern "C" int pyopal_cuda_search_scores_device(pyopal_cuda_db* db, const unsigned char* query, int queryLength,
                                                int gapOpen, int gapExt, const int* scoreMatrix,
                                                int alphabetLength, int mode, int64_t start, int64_t end,
                                                int32_t* d_scores, void* stream) {
    int rc = check_search_args(db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, &start, &end);
    if (rc) return rc;
    if (end > start && !d_scores) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    DeviceGuard guard(db->device);
    Lease lease(db);
    if (!lease.ws) return PYOPAL_CUDA_ERR_NOMEM;
    // the workspace goes back to the pool when this returns; its `ev_free` event orders the next search that
    // picks it up (on whatever stream) after this one
    return enqueue_score_search(db, lease.ws, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode,
                                start, end, d_scores, (cudaStream_t)stream);
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
    DeviceGuard guard(db->device);
    Lease lease(db);
    Workspace* ws = lease.ws;
    if (!ws) return PYOPAL_CUDA_ERR_NOMEM;
    if (searchType != OPAL_SEARCH_SCORE)
        return pyopal_end_full_search(db, ws, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength,
                                      searchType, mode, start, end, scores, endQ, endT, startQ, startT, ops,
                                      ops_offsets, ops_len);
    if (ws->out.reserve((size_t)nout)) return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaEventRecord(ws->ev0, ws->stream));
    rc = enqueue_score_search(db, ws, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end,
                              ws->out.p, ws->stream);
    if (rc) return rc;
    CU(cudaMemcpyAsync(scores, ws->out.p, (size_t)nout * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    CU(cudaEventRecord(ws->ev1, ws->stream));
    CU(cudaStreamSynchronize(ws->stream));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, ws->ev0, ws->ev1));
    ws->stats[4] = (int64_t)(ms * 1000.0f);
    if (ws->d_ovf_count) {  // targets escalated 16 -> 32 bit (statistics only)
        int32_t novf = 0;
        CU(cudaMemcpy(&novf, ws->d_ovf_count, sizeof(int32_t), cudaMemcpyDeviceToHost));
        ws->stats[3] = novf;
    }
    return 0;
}

// opal.pxd:56-67: Smith-Waterman SCORE search over `char` sequences (declared by the interface PyOpal binds,
// never called by it).  Same staging as opalSearchDatabase.
extern "C" int opalSearchDatabaseCharSW(char query[], int queryLength, char** db, int dbLength, int dbSeqLengths[],
                                        int gapOpen, int gapExt, int* scoreMatrix, int alphabetLength,
                                        OpalSearchResult* results[]) {
    return opalSearchDatabase(reinterpret_cast<unsigned char*>(query), queryLength,
                              reinterpret_cast<unsigned char**>(db), dbLength, dbSeqLengths, gapOpen, gapExt,
                              scoreMatrix, alphabetLength, results, OPAL_SEARCH_SCORE, OPAL_MODE_SW,
                              OPAL_OVERFLOW_BUCKETS);
}

// Batched SCORE search, stream-ordered: query i runs on workspace i mod 2, so the kernels of query i + 1 are
// already queued behind those of query i (its persistent CTAs take over the SMs as the previous search's drain),
// score rows are copied out on the workspace's own stream, and the host only waits once, at the end.  With
// `k` > 0 only the k best hits of every query leave the device (topk.cu).
// Queries of a batch in the order they are searched: by length, neighbours paired.  A pair of Smith-Waterman
// queries goes down as ONE dual search when the database allows it (enqueue_dual_search), else one after the other.
void pyopal::plan_batch_units(const pyopal_cuda_db* db, const int32_t* q_lengths, int nq, int mode,
                              std::vector<std::pair<int, int>>& units) {
    batch_units(q_lengths, nq, mode == MODE_SW && !db->tune.no_dual, units);
}

// Pure host code (every rank of a multi-GPU handle must derive the same units: the collectives have to match).
// A dual pass costs 2 x max(La, Lb) x 4.5 ALU instructions per pair column (it walks the two targets of a pair one
// after the other), two single searches (La + Lb) x 5.5: pairing pays when the shorter query is at least 0.64 of the
// longer one, so neighbours in length are paired from 0.7 up and everything else is searched alone.
void pyopal::batch_units(const int32_t* q_lengths, int nq, bool pair_queries, std::vector<std::pair<int, int>>& units) {
    units.clear();
    std::vector<int> order((size_t)nq);
    std::iota(order.begin(), order.end(), 0);
    if (!pair_queries || nq < 2) {
        for (int i = 0; i < nq; i++) units.emplace_back(i, -1);
        return;
    }
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return q_lengths[a] > q_lengths[b]; });
    for (int k = 0; k < nq;) {
        const int a = order[(size_t)k];
        const int b = k + 1 < nq ? order[(size_t)k + 1] : -1;
        if (b >= 0 && q_lengths[b] > 0 && 10 * (int64_t)q_lengths[b] >= 7 * (int64_t)q_lengths[a]) {
            units.emplace_back(a, b);
            k += 2;
        } else {
            units.emplace_back(a, -1);
            k += 1;
        }
    }
}

static int search_batch_impl(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                             const int32_t* q_lengths, int nq, int go, int ge, const int* M, int A, int mode,
                             int32_t* scores, int k, int32_t* top_scores, int32_t* top_idx) {
    if (!db || nq < 0 || (nq > 0 && (!queries || !q_offsets || !q_lengths)))
        return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    const int64_t n = db->n;
    if (n == 0 || nq == 0) return 0;
    DeviceGuard guard(db->device);
    // one lane: the searches are stream-ordered back to back (no host round trip between them).  Two lanes let the
    // next search start while the last CTAs of the previous one drain, but its high-priority giants then queue
    // behind the resident persistent CTAs and end up last: measured slower, so opt-in.
    const int NB = db->tune.lanes;
    Lease l0(db);
    std::unique_ptr<Lease> l1(NB > 1 ? new Lease(db) : nullptr);
    Workspace* wsv[2] = {l0.ws, NB > 1 ? l1->ws : l0.ws};
    if (!wsv[0] || !wsv[1]) return PYOPAL_CUDA_ERR_NOMEM;
    const int kk = (int)std::min<int64_t>(k, n);
    std::vector<int32_t> tv, ti;
    if (kk > 0) { tv.resize((size_t)nq * kk); ti.resize((size_t)nq * kk); }
    for (int i = 0; i < nq; i++) {
        int64_t s0 = 0, e0 = n;
        int rc0 = check_search_args(db, queries + q_offsets[i], q_lengths[i], go, ge, M, A, mode, &s0, &e0);
        if (rc0) return rc0;
    }
    std::vector<std::pair<int, int>> units;
    plan_batch_units(db, q_lengths, nq, mode, units);
    // two sets of result buffers alternate between consecutive UNITS (one or two queries: two rows per set)
    // whatever the number of lanes
    DevBuf<int32_t>* rowb[2] = {&wsv[0]->out, NB > 1 ? &wsv[1]->out : &wsv[0]->out2};
    DevBuf<int32_t>* tvb[2] = {&wsv[0]->tk_vals, NB > 1 ? &wsv[1]->tk_vals : &wsv[0]->tk_vals2};
    DevBuf<int32_t>* tib[2] = {&wsv[0]->tk_idx, NB > 1 ? &wsv[1]->tk_idx : &wsv[0]->tk_idx2};
    for (int b = 0; b < 2; b++) {
        if (rowb[b]->reserve((size_t)(2 * n))) return PYOPAL_CUDA_ERR_NOMEM;
        if (kk > 0 && (wsv[b]->tk_ws.reserve(topk_workspace_words()) || tvb[b]->reserve((size_t)(2 * kk)) || tib[b]->reserve((size_t)(2 * kk))))
            return PYOPAL_CUDA_ERR_NOMEM;
    }
    int rc = 0;
    int64_t launches = 0;
    // results of unit u are copied out AFTER unit u + 1 has been enqueued: a copy into pageable host memory
    // blocks the host until the search before it is done, and by then the next one is already queued on the device
    auto copy_out = [&](int u) -> int {
        Workspace* ws = wsv[u % 2];
        const int b = u % 2;
        const int qi[2] = {units[(size_t)u].first, units[(size_t)u].second};
        for (int s = 0; s < 2; s++) {
            if (qi[s] < 0) continue;
            if (kk > 0) {
                CU(cudaMemcpyAsync(tv.data() + (size_t)qi[s] * kk, tvb[b]->p + (size_t)s * kk, (size_t)kk * 4, cudaMemcpyDeviceToHost, ws->stream));
                CU(cudaMemcpyAsync(ti.data() + (size_t)qi[s] * kk, tib[b]->p + (size_t)s * kk, (size_t)kk * 4, cudaMemcpyDeviceToHost, ws->stream));
            } else if (scores) {
                CU(cudaMemcpyAsync(scores + (size_t)qi[s] * (size_t)n, rowb[b]->p + (size_t)s * (size_t)n, (size_t)n * sizeof(int32_t),
                                   cudaMemcpyDeviceToHost, ws->stream));
            }
        }
        CU(cudaEventRecord(ws->ev_free, ws->stream));
        return 0;
    };
    const int nu = (int)units.size();
    for (int u = 0; u < nu && rc == 0; u++) {
        Workspace* ws = wsv[u % 2];
        const int b = u % 2;
        const int qa = units[(size_t)u].first, qb = units[(size_t)u].second;
        int32_t* rows[2] = {rowb[b]->p, rowb[b]->p + n};
        // this buffer set is overwritten: its previous copy-out was issued (and, into pageable memory, completed)
        // before this point, on a stream this search is ordered after
        bool done = false;
        if (qb >= 0) {
            rc = enqueue_dual_search(db, ws, queries + q_offsets[qa], q_lengths[qa], queries + q_offsets[qb], q_lengths[qb],
                                     go, ge, M, A, rows[0], rows[1], ws->stream);
            if (rc == 0) { done = true; launches += ws->stats[0]; }
            else if (rc != DUAL_NOT_ELIGIBLE) break;
            rc = 0;
        }
        if (!done) {
            const int qs[2] = {qa, qb};
            for (int s = 0; s < 2 && rc == 0; s++) {
                if (qs[s] < 0) continue;
                rc = enqueue_score_search(db, ws, queries + q_offsets[qs[s]], q_lengths[qs[s]], go, ge, M, A, mode, 0, n, rows[s], ws->stream);
                launches += ws->stats[0];
            }
            if (rc) break;
        }
        if (kk > 0) {
            for (int s = 0; s < 2 && rc == 0; s++) {
                if ((s == 0 ? qa : qb) < 0) continue;
                rc = topk_device(rows[s], n, kk, tvb[b]->p + (size_t)s * kk, tib[b]->p + (size_t)s * kk, wsv[b]->tk_ws.p, db->sm_count, ws->stream);
                launches += 9;
            }
            if (rc) break;
        }
        if (u > 0) rc = copy_out(u - 1);
    }
    if (rc == 0 && nu > 0) rc = copy_out(nu - 1);
    for (int b = 0; b < 2; b++) {
        cudaError_t e = cudaStreamSynchronize(wsv[b]->stream);
        if (e != cudaSuccess && rc == 0) rc = fail(PYOPAL_CUDA_ERR_CUDA, std::string("batch: ") + cudaGetErrorString(e));
    }
    wsv[0]->stats[0] = launches;
    wsv[1]->stats[0] = launches;
    if (rc) return rc;
    if (kk > 0) {
        // order the k hits of every query: score descending, ties by ascending database index
        std::vector<int32_t> order((size_t)kk);
        for (int i = 0; i < nq; i++) {
            const int32_t* v = tv.data() + (size_t)i * kk;
            const int32_t* ix = ti.data() + (size_t)i * kk;
            std::iota(order.begin(), order.end(), 0);
            std::sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return v[a] != v[b] ? v[a] > v[b] : ix[a] < ix[b]; });
            for (int j = 0; j < kk; j++) {
                top_scores[(size_t)i * k + j] = v[order[(size_t)j]];
                top_idx[(size_t)i * k + j] = ix[order[(size_t)j]];
            }
        }
    }
    return 0;
}

extern "C" int pyopal_cuda_search_batch(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                                        const int32_t* q_lengths, int nq, int gapOpen, int gapExt,
                                        const int* scoreMatrix, int alphabetLength, int mode, int32_t* scores) {
    if (nq > 0 && db && db->n > 0 && !scores) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    return search_batch_impl(db, queries, q_offsets, q_lengths, nq, gapOpen, gapExt, scoreMatrix, alphabetLength, mode,
                             scores, 0, nullptr, nullptr);
}

extern "C" int pyopal_cuda_search_batch_topk(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                                             const int32_t* q_lengths, int nq, int gapOpen, int gapExt,
                                             const int* scoreMatrix, int alphabetLength, int mode, int k,
                                             int32_t* scores, int32_t* indices) {
    if (k <= 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "k must be positive");
    if (nq > 0 && (!scores || !indices)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    if (db && k > db->n) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "k exceeds the database size");
    return search_batch_impl(db, queries, q_offsets, q_lengths, nq, gapOpen, gapExt, scoreMatrix, alphabetLength, mode,
                             nullptr, k, scores, indices);
}

extern "C" int pyopal_cuda_topk_device(pyopal_cuda_db* db, const int32_t* d_scores, int64_t n, int k,
                                       int32_t* d_scores_out, int32_t* d_indices_out, void* stream) {
    if (!db || (n > 0 && k > 0 && (!d_scores || !d_scores_out || !d_indices_out))) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    if (k < 0 || n < 0 || k > n || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "top-k needs 0 <= k <= n < 2^31");
    DeviceGuard guard(db->device);
    Lease lease(db);
    Workspace* ws = lease.ws;
    if (!ws) return PYOPAL_CUDA_ERR_NOMEM;
    cudaStream_t st = (cudaStream_t)stream;
    if (ws->tk_ws.reserve(topk_workspace_words())) return PYOPAL_CUDA_ERR_NOMEM;
    if (ws->used) CU(cudaStreamWaitEvent(st, ws->ev_free, 0));
    ws->used = true;
    int rc = topk_device(d_scores, n, k, d_scores_out, d_indices_out, ws->tk_ws.p, db->sm_count, st);
    CU(cudaEventRecord(ws->ev_free, st));
    return rc;
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
    DeviceGuard guard(db->device);
    Lease lease(db);
    Workspace* ws = lease.ws;
    if (!ws) return PYOPAL_CUDA_ERR_NOMEM;
    if (ws->out.reserve((size_t)nout) || ws->tk_ws.reserve(topk_workspace_words()) || ws->tk_vals.reserve((size_t)kk) ||
        ws->tk_idx.reserve((size_t)kk))
        return PYOPAL_CUDA_ERR_NOMEM;
    CU(cudaEventRecord(ws->ev0, ws->stream));
    rc = enqueue_score_search(db, ws, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end,
                              ws->out.p, ws->stream);
    if (rc) return rc;
    rc = topk_device(ws->out.p, nout, kk, ws->tk_vals.p, ws->tk_idx.p, ws->tk_ws.p, db->sm_count, ws->stream);
    if (rc) return rc;
    ws->stats[0] += 9;
    std::vector<int32_t> v((size_t)kk), ix((size_t)kk);
    CU(cudaMemcpyAsync(v.data(), ws->tk_vals.p, (size_t)kk * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    CU(cudaMemcpyAsync(ix.data(), ws->tk_idx.p, (size_t)kk * sizeof(int32_t), cudaMemcpyDeviceToHost, ws->stream));
    CU(cudaEventRecord(ws->ev1, ws->stream));
    CU(cudaEventRecord(ws->ev_free, ws->stream));
    CU(cudaStreamSynchronize(ws->stream));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, ws->ev0, ws->ev1));
    ws->stats[4] = (int64_t)(ms * 1000.0f);
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
    // the calling thread's current device, like any CUDA library entered without a handle; the Python layer selects
    // its default device (PYOPAL_CUDA_DEVICE / LOCAL_RANK) before this call
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
