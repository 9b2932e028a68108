// pyopal_b200/csrc/multi.cu -- one database over several B200s: the multi-device handle of the C-ABI.
//
// Replaces the thread fan-out of /root/reference/src/pyopal/_align.py:145-172 (a ThreadPool over contiguous
// chunks of equal sequence COUNT).  Here the database is dealt to the GPUs by cumulative CELL count: targets
// sorted by length go to the ranks in snake order, so every shard holds the same length mix, the longest targets
// spread out and the residue totals differ by at most one target.  Targets are independent, so a search is
//     per rank:  shard search (api.cu) -> int32 scores in shard order, padded to n_pad
//     exchange:  ncclAllGather of the padded shards over NVLink / NVSwitch        (the only collective)
//     per rank:  un-permute the gathered buffer to database order, copy out once
// Two layouts share the code: ONE process driving every visible GPU (ncclCommInitAll, one host thread per device;
// what Aligner.align / pyopal.align use), and one process PER GPU (ncclCommInitRank with a unique id the caller
// distributes; what `torchrun bench.py` uses).  NCCL is bound at run time (dlopen of libnccl.so.2), so a
// single-GPU install needs no NCCL at all.  No PyTorch anywhere.
#include "store.h"

#include <dlfcn.h>

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>
#include <thread>

using namespace pyopal;

// ---------------------------------------------------------------------------
// NCCL, bound at run time.  The handful of entry points used here have had the same C ABI since NCCL 2.0.
namespace {

typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { NCCL_INT32 = 2 };  // ncclDataType_t::ncclInt32

struct Nccl {
    void* so = nullptr;
    int (*GetVersion)(int*) = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllGather)(const void*, void*, size_t, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    std::string path;
};
Nccl g_nccl;
std::mutex g_nccl_mu;

int nccl_load(const char* explicit_path) {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.so) return 0;
    std::vector<std::string> names;
    // a copy the process has already mapped wins (RTLD_NOLOAD): one NCCL per process, whoever came first
    if (explicit_path && *explicit_path) names.push_back(explicit_path);
    if (const char* e = getenv("PYOPAL_NCCL_LIB")) names.push_back(e);
    names.push_back("libnccl.so.2");
    names.push_back("libnccl.so");
    std::string errs;
    bool first = true;
    for (const std::string& nm : names) {
        // no RTLD_GLOBAL: this library looks its symbols up with dlsym and must not shadow anybody else's
        void* so = first ? dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD) : dlopen(nm.c_str(), RTLD_NOW | RTLD_LOCAL);
        if (first) { first = false; if (!so) so = dlopen(nm.c_str(), RTLD_NOW | RTLD_LOCAL); }
        if (!so) { errs += std::string(dlerror() ? dlerror() : "?") + "; "; continue; }
        Nccl n;
        n.so = so;
        n.path = nm;
#define SYM(field, name) *(void**)(&n.field) = dlsym(so, name)
        SYM(GetVersion, "ncclGetVersion");
        SYM(GetUniqueId, "ncclGetUniqueId");
        SYM(CommInitRank, "ncclCommInitRank");
        SYM(CommInitAll, "ncclCommInitAll");
        SYM(CommDestroy, "ncclCommDestroy");
        SYM(AllGather, "ncclAllGather");
        SYM(GroupStart, "ncclGroupStart");
        SYM(GroupEnd, "ncclGroupEnd");
        SYM(GetErrorString, "ncclGetErrorString");
#undef SYM
        if (n.GetUniqueId && n.CommInitRank && n.CommInitAll && n.CommDestroy && n.AllGather && n.GroupStart && n.GroupEnd) {
            g_nccl = n;
            return 0;
        }
        errs += nm + ": missing symbols; ";
        dlclose(so);
    }
    return fail(PYOPAL_CUDA_ERR_NCCL, "NCCL library not found (" + errs + ")");
}

#define NC(call)                                                                                      \
    do {                                                                                              \
        int r_ = (call);                                                                              \
        if (r_ != 0)                                                                                  \
            return ::pyopal::fail(PYOPAL_CUDA_ERR_NCCL, std::string(#call) + ": " +                   \
                                  (g_nccl.GetErrorString ? g_nccl.GetErrorString(r_) : "NCCL error")); \
    } while (0)

// database order <- gathered shard order.  `rows` score rows per rank travel together (a dual search produces two):
// gathered[(r * rows + s) * n_pad + k] is row s of rank r's slot k; row s lands at out + s * out_stride.
__global__ void unpermute_kernel(const int32_t* __restrict__ gathered, const int32_t* __restrict__ dest, int64_t total,
                                 int64_t n_pad, int rows, int64_t start, int64_t end, int32_t* __restrict__ out,
                                 int64_t out_stride) {
    for (int64_t j = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; j < total; j += (int64_t)gridDim.x * blockDim.x) {
        const int64_t d = dest[j];
        if (d < start || d >= end) continue;
        const int64_t r = j / n_pad, k = j - r * n_pad;
        for (int s = 0; s < rows; s++) out[s * out_stride + d - start] = gathered[(r * rows + s) * n_pad + k];
    }
}

// a persistent host thread per local rank: searches on several GPUs are planned and enqueued concurrently
struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    std::function<int()> job;
    bool has_job = false, done = false, quit = false;
    int rc = 0;
    std::string err;
    void loop() {
        for (;;) {
            std::function<int()> f;
            {
                std::unique_lock<std::mutex> lock(mu);
                cv.wait(lock, [&] { return has_job || quit; });
                if (quit) return;
                f = job;
                has_job = false;
            }
            int r = f();
            std::string e = r ? pyopal_cuda_last_error() : "";
            {
                std::lock_guard<std::mutex> lock(mu);
                rc = r; err = e; done = true;
            }
            cv.notify_all();
        }
    }
};

constexpr int NLANE = 2;  // searches in flight per rank

struct Lane {
    Workspace* ws = nullptr;                     // leased from the shard for the life of the handle
    cudaStream_t gstream = nullptr;              // high priority: gather, un-permute, top-k, copy-out
    cudaEvent_t ev_search = nullptr, ev_gather = nullptr;
    bool gathered_once = false;
    // two score rows per lane: a unit of a batch is one search or a dual search of two queries (api.cu)
    int32_t* d_local = nullptr;                  // [2 * n_pad] this rank's results, shard order
    int32_t* d_gathered = nullptr;               // [world * 2 * n_pad]
    int32_t* d_final = nullptr;                  // [2 * n_all] database order
    DevBuf<int32_t> tk_vals, tk_idx;
    DevBuf<uint32_t> tk_ws;
};

struct Local {
    int device = 0, rank = 0;
    pyopal_cuda_db* db = nullptr;
    ncclComm_t comm = nullptr;
    std::vector<int32_t> mine;                   // database indices of this shard, ascending
    int32_t* d_dest = nullptr;                   // [world * n_pad] database index of every gathered slot, -1 = padding
    Lane lane[NLANE];
    cudaEvent_t t0 = nullptr, t1 = nullptr;
    std::unique_ptr<Worker> worker;
};

}  // namespace

struct pyopal_cuda_mdb {
    int world = 1, nlocal = 1, A = 0;
    int64_t n_all = 0, n_pad = 0, total_len = 0;
    bool identity = false;                       // world 1 and shard order == database order: no gather, no un-permute
    std::vector<std::unique_ptr<Local>> local;
    std::mutex mu;                               // collectives are issued in one order: one search call at a time
    int64_t last_launches = 0;
    double last_device_ms = 0;
};

namespace {

int run_all(pyopal_cuda_mdb* m, const std::function<int(int)>& fn) {
    if (m->nlocal == 1) return fn(0);
    for (int L = 0; L < m->nlocal; L++) {
        Worker* w = m->local[(size_t)L]->worker.get();
        {
            std::lock_guard<std::mutex> lock(w->mu);
            w->job = [fn, L]() { return fn(L); };
            w->has_job = true;
            w->done = false;
        }
        w->cv.notify_all();
    }
    int rc = 0;
    std::string err;
    for (int L = 0; L < m->nlocal; L++) {
        Worker* w = m->local[(size_t)L]->worker.get();
        std::unique_lock<std::mutex> lock(w->mu);
        w->cv.wait(lock, [&] { return w->done; });
        if (w->rc && !rc) { rc = w->rc; err = w->err; }
    }
    if (rc) fail(rc, err);
    return rc;
}

void mdb_free(pyopal_cuda_mdb* m) {
    if (!m) return;
    for (auto& lp : m->local) {
        Local* l = lp.get();
        if (l->worker) {
            { std::lock_guard<std::mutex> lock(l->worker->mu); l->worker->quit = true; }
            l->worker->cv.notify_all();
            if (l->worker->th.joinable()) l->worker->th.join();
        }
        DeviceGuard guard(l->device);
        cudaDeviceSynchronize();
        cudaGetLastError();
        if (l->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(l->comm);
        for (Lane& ln : l->lane) {
            if (ln.ws && l->db) l->db->release(ln.ws);
            if (ln.gstream) cudaStreamDestroy(ln.gstream);
            if (ln.ev_search) cudaEventDestroy(ln.ev_search);
            if (ln.ev_gather) cudaEventDestroy(ln.ev_gather);
            cudaFree(ln.d_local); cudaFree(ln.d_gathered); cudaFree(ln.d_final);
            ln.tk_vals.release(); ln.tk_idx.release(); ln.tk_ws.release();
        }
        cudaFree(l->d_dest);
        if (l->t0) cudaEventDestroy(l->t0);
        if (l->t1) cudaEventDestroy(l->t1);
        if (l->db) pyopal_cuda_db_destroy(l->db);
    }
    delete m;
}

// lanes, buffers and the destination table of one local rank; `dest_all` = database index of every slot of the
// gathered buffer (world * n_pad entries, -1 for padding), nullptr when the handle is an identity
int local_setup(pyopal_cuda_mdb* m, Local* l, const int32_t* dest_all) {
    DeviceGuard guard(l->device);
    int lo_prio = 0, hi_prio = 0;
    CU(cudaDeviceGetStreamPriorityRange(&lo_prio, &hi_prio));
    CU(cudaEventCreate(&l->t0));
    CU(cudaEventCreate(&l->t1));
    const size_t npad = (size_t)std::max<int64_t>(m->n_pad, 1);
    for (Lane& ln : l->lane) {
        ln.ws = l->db->acquire();
        if (!ln.ws) return PYOPAL_CUDA_ERR_NOMEM;
        CU(cudaStreamCreateWithPriority(&ln.gstream, cudaStreamNonBlocking, hi_prio));
        CU(cudaEventCreateWithFlags(&ln.ev_search, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ln.ev_gather, cudaEventDisableTiming));
        CU(cudaMalloc(&ln.d_local, 2 * npad * sizeof(int32_t)));
        CU(cudaMemset(ln.d_local, 0, 2 * npad * sizeof(int32_t)));
        if (!m->identity) {
            CU(cudaMalloc(&ln.d_gathered, 2 * npad * (size_t)m->world * sizeof(int32_t)));
            CU(cudaMalloc(&ln.d_final, 2 * (size_t)std::max<int64_t>(m->n_all, 1) * sizeof(int32_t)));
        }
    }
    if (!m->identity) {
        CU(cudaMalloc(&l->d_dest, npad * (size_t)m->world * sizeof(int32_t)));
        CU(cudaMemcpy(l->d_dest, dest_all, npad * (size_t)m->world * sizeof(int32_t), cudaMemcpyHostToDevice));
    }
    if (m->nlocal > 1) {
        l->worker.reset(new Worker());
        Worker* w = l->worker.get();
        w->th = std::thread([w]() { w->loop(); });
    }
    return 0;
}

}  // namespace

// ---------------------------------------------------------------------------
extern "C" {

int pyopal_cuda_nccl_load(const char* path) { return nccl_load(path); }

int pyopal_cuda_nccl_loaded(void) {
    std::lock_guard<std::mutex> lock(g_nccl_mu);
    if (g_nccl.so) return 1;
    void* so = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);   // mapped by somebody else already?
    if (!so) return 0;
    dlclose(so);
    return 0;   // not bound yet; nccl_load(NULL) will pick the mapped copy up
}

int pyopal_cuda_nccl_version(void) {
    if (nccl_load(nullptr)) return 0;
    int v = 0;
    if (g_nccl.GetVersion) g_nccl.GetVersion(&v);
    return v;
}

int pyopal_cuda_nccl_unique_id(unsigned char id[128]) {
    if (!id) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    int rc = nccl_load(nullptr);
    if (rc) return rc;
    ncclUniqueId u;
    NC(g_nccl.GetUniqueId(&u));
    memcpy(id, u.internal, 128);
    return 0;
}

// Deal `n` targets to `world` ranks by cumulative cells: position in the length-sorted order (descending, ties
// by index) -> snake over the ranks.  Pure host code (also what the CPU tests of the N > 1 path exercise).
int pyopal_cuda_deal(const int32_t* lengths, int64_t n, int world, int32_t* owner) {
    if (n < 0 || world < 1 || (n > 0 && (!lengths || !owner))) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "bad argument");
    std::vector<int32_t> order((size_t)n);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return lengths[a] > lengths[b]; });
    for (int64_t pos = 0; pos < n; pos++) {
        const int64_t rnd = pos / world, k = pos % world;
        owner[order[(size_t)pos]] = (int32_t)((rnd % 2 == 0) ? k : world - 1 - k);
    }
    return 0;
}

void pyopal_cuda_mdb_destroy(pyopal_cuda_mdb* m) { mdb_free(m); }
int64_t pyopal_cuda_mdb_size(const pyopal_cuda_mdb* m) { return m ? m->n_all : 0; }
int pyopal_cuda_mdb_world(const pyopal_cuda_mdb* m) { return m ? m->world : 0; }
int pyopal_cuda_mdb_local_ranks(const pyopal_cuda_mdb* m) { return m ? m->nlocal : 0; }
int64_t pyopal_cuda_mdb_local_residues(const pyopal_cuda_mdb* m, int local_rank) {
    if (!m || local_rank < 0 || local_rank >= m->nlocal) return 0;
    return m->local[(size_t)local_rank]->db->total_len;
}
int64_t pyopal_cuda_mdb_local_size(const pyopal_cuda_mdb* m, int local_rank) {
    if (!m || local_rank < 0 || local_rank >= m->nlocal) return 0;
    return m->local[(size_t)local_rank]->db->n;
}
pyopal_cuda_db* pyopal_cuda_mdb_shard(pyopal_cuda_mdb* m, int local_rank) {
    if (!m || local_rank < 0 || local_rank >= m->nlocal) return nullptr;
    return m->local[(size_t)local_rank]->db;
}

// ONE process, `ndev` devices: deal the packed database, build every shard (one host thread per device),
// ncclCommInitAll.
int pyopal_cuda_mdb_create(const int* devices, int ndev, const unsigned char* residues, const int64_t* offsets,
                           const int32_t* lengths, int64_t n, int alphabetLength, pyopal_cuda_mdb** out) {
    if (!out) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output handle");
    *out = nullptr;
    if (ndev < 1 || !devices) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "no devices");
    if (n < 0 || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "database size out of range");
    if (n > 0 && (!residues || !offsets || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    const int navail = pyopal_cuda_device_count();
    if (navail <= 0) return fail(OPAL_ERR_NO_SIMD_SUPPORT, "no CUDA device available");
    for (int i = 0; i < ndev; i++)
        if (devices[i] < 0 || devices[i] >= navail) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "device index out of range");
    pyopal_cuda_mdb* m = new (std::nothrow) pyopal_cuda_mdb();
    if (!m) return fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed");
    struct Guard { pyopal_cuda_mdb* p; ~Guard() { if (p) mdb_free(p); } } guard{m};
    m->world = m->nlocal = ndev;
    m->A = alphabetLength;
    m->n_all = n;
    m->identity = (ndev == 1);
    std::vector<int32_t> owner((size_t)n, 0);
    if (ndev > 1) {
        int rc = pyopal_cuda_deal(lengths, n, ndev, owner.data());
        if (rc) return rc;
    }
    int64_t biggest = 0;
    for (int r = 0; r < ndev; r++) {
        std::unique_ptr<Local> l(new Local());
        l->device = devices[r];
        l->rank = r;
        m->local.push_back(std::move(l));
    }
    for (int64_t i = 0; i < n; i++) m->local[(size_t)owner[(size_t)i]]->mine.push_back((int32_t)i);
    for (auto& l : m->local) biggest = std::max<int64_t>(biggest, (int64_t)l->mine.size());
    m->n_pad = biggest;
    for (int64_t i = 0; i < n; i++) m->total_len += lengths[i];
    // build the shards concurrently, straight from the caller's packed buffer (no host-side repacking)
    std::vector<int> rcs((size_t)ndev, 0);
    std::vector<std::string> errs((size_t)ndev);
    {
        std::vector<std::thread> th;
        for (int r = 0; r < ndev; r++) {
            th.emplace_back([&, r]() {
                Local* l = m->local[(size_t)r].get();
                std::vector<int32_t> sl(l->mine.size());
                for (size_t k = 0; k < sl.size(); k++) sl[k] = lengths[l->mine[k]];
                rcs[(size_t)r] = db_build_from(l->device, residues ? residues : (const unsigned char*)"", offsets, nullptr,
                                               ndev > 1 ? l->mine.data() : nullptr, sl.data(), (int64_t)sl.size(),
                                               alphabetLength, &l->db);
                if (rcs[(size_t)r]) errs[(size_t)r] = pyopal_cuda_last_error();
            });
        }
        for (auto& t : th) t.join();
    }
    for (int r = 0; r < ndev; r++)
        if (rcs[(size_t)r]) return fail(rcs[(size_t)r], errs[(size_t)r]);
    std::vector<int32_t> dest;
    if (!m->identity) {
        dest.assign((size_t)(m->n_pad * ndev), -1);
        for (int r = 0; r < ndev; r++)
            memcpy(dest.data() + (size_t)r * (size_t)m->n_pad, m->local[(size_t)r]->mine.data(),
                   m->local[(size_t)r]->mine.size() * sizeof(int32_t));
        int rc = nccl_load(nullptr);
        if (rc) return rc;
        std::vector<ncclComm_t> comms((size_t)ndev, nullptr);
        NC(g_nccl.CommInitAll(comms.data(), ndev, devices));
        for (int r = 0; r < ndev; r++) m->local[(size_t)r]->comm = comms[(size_t)r];
    }
    for (auto& l : m->local) {
        int rc = local_setup(m, l.get(), dest.empty() ? nullptr : dest.data());
        if (rc) return rc;
    }
    guard.p = nullptr;
    *out = m;
    return 0;
}

// One process PER GPU: this process is `rank` of `world`, owns the n_local targets it passes (packed) whose
// database indices are `global_index` (ascending; NULL with world 1 = identity), and joins the communicator
// identified by `nccl_id` (pyopal_cuda_nccl_unique_id on rank 0, distributed by the caller).  Collective: every
// rank must call it, and later every search, in the same order.
static int create_rank_impl(int device, int rank, int world, const unsigned char nccl_id[128],
                            const unsigned char* residues, const int64_t* offsets, const int32_t* src_index,
                            const int32_t* lengths, int64_t n_local, const int32_t* global_index, int64_t n_all,
                            int alphabetLength, pyopal_cuda_mdb** out) {
    if (!out) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output handle");
    *out = nullptr;
    if (world < 1 || rank < 0 || rank >= world) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "bad rank / world");
    if (n_local < 0 || n_all < n_local || n_all > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "bad sizes");
    if (world > 1 && (!nccl_id || (n_local > 0 && !global_index))) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    pyopal_cuda_mdb* m = new (std::nothrow) pyopal_cuda_mdb();
    if (!m) return fail(PYOPAL_CUDA_ERR_NOMEM, "host allocation failed");
    struct Guard { pyopal_cuda_mdb* p; ~Guard() { if (p) mdb_free(p); } } guard{m};
    m->world = world;
    m->nlocal = 1;
    m->A = alphabetLength;
    m->n_all = n_all;
    m->identity = (world == 1 && global_index == nullptr);
    std::unique_ptr<Local> lp(new Local());
    Local* l = lp.get();
    l->device = device;
    l->rank = rank;
    m->local.push_back(std::move(lp));
    l->mine.resize((size_t)n_local);
    for (int64_t k = 0; k < n_local; k++) {
        l->mine[(size_t)k] = global_index ? global_index[k] : (int32_t)k;
        if (l->mine[(size_t)k] < 0 || l->mine[(size_t)k] >= n_all || (k > 0 && l->mine[(size_t)k] <= l->mine[(size_t)(k - 1)]))
            return fail(PYOPAL_CUDA_ERR_ARGUMENT, "global_index must be ascending and within the database");
    }
    int rc = db_build_from(device, residues ? residues : (const unsigned char*)"", offsets, nullptr, src_index, lengths,
                           n_local, alphabetLength, &l->db);
    if (rc) return rc;
    m->total_len = l->db->total_len;
    m->n_pad = n_local;
    std::vector<int32_t> dest;
    if (!m->identity) {
        DeviceGuard dg(device);
        if (world > 1) {
            rc = nccl_load(nullptr);
            if (rc) return rc;
            ncclUniqueId u;
            memcpy(u.internal, nccl_id, 128);
            NC(g_nccl.CommInitRank(&l->comm, world, u, rank));
        }
        // agree on the padded shard size, then exchange the index tables (both through the communicator itself)
        int32_t* d_tmp = nullptr;
        CU(cudaMalloc(&d_tmp, (size_t)(world + 1) * sizeof(int32_t)));
        const int32_t mine_n = (int32_t)n_local;
        CU(cudaMemcpy(d_tmp + world, &mine_n, sizeof(int32_t), cudaMemcpyHostToDevice));
        if (world > 1) NC(g_nccl.AllGather(d_tmp + world, d_tmp, 1, NCCL_INT32, l->comm, (cudaStream_t)0));
        else CU(cudaMemcpy(d_tmp, d_tmp + world, sizeof(int32_t), cudaMemcpyDeviceToDevice));
        std::vector<int32_t> counts((size_t)world);
        CU(cudaMemcpy(counts.data(), d_tmp, (size_t)world * sizeof(int32_t), cudaMemcpyDeviceToHost));
        cudaFree(d_tmp);
        int64_t sum = 0;
        for (int32_t c : counts) { m->n_pad = std::max<int64_t>(m->n_pad, c); sum += c; }
        if (sum != n_all) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "the ranks' shards do not add up to the database size");
        const size_t npad = (size_t)std::max<int64_t>(m->n_pad, 1);
        int32_t *d_mine = nullptr, *d_all = nullptr;
        CU(cudaMalloc(&d_mine, npad * sizeof(int32_t)));
        CU(cudaMalloc(&d_all, npad * (size_t)world * sizeof(int32_t)));
        CU(cudaMemset(d_mine, 0xff, npad * sizeof(int32_t)));
        CU(cudaMemcpy(d_mine, l->mine.data(), (size_t)n_local * sizeof(int32_t), cudaMemcpyHostToDevice));
        if (world > 1) NC(g_nccl.AllGather(d_mine, d_all, npad, NCCL_INT32, l->comm, (cudaStream_t)0));
        else CU(cudaMemcpy(d_all, d_mine, npad * sizeof(int32_t), cudaMemcpyDeviceToDevice));
        dest.resize(npad * (size_t)world);
        CU(cudaMemcpy(dest.data(), d_all, dest.size() * sizeof(int32_t), cudaMemcpyDeviceToHost));
        cudaFree(d_mine); cudaFree(d_all);
        // the shards must partition the database
        std::vector<unsigned char> seen((size_t)n_all, 0);
        for (int32_t d : dest) {
            if (d < 0) continue;
            if (d >= n_all || seen[(size_t)d]) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "shards do not partition the database");
            seen[(size_t)d] = 1;
        }
    }
    rc = local_setup(m, l, dest.empty() ? nullptr : dest.data());
    if (rc) return rc;
    guard.p = nullptr;
    *out = m;
    return 0;
}

int pyopal_cuda_mdb_create_rank(int device, int rank, int world, const unsigned char nccl_id[128],
                                const unsigned char* residues, const int64_t* offsets, const int32_t* lengths,
                                int64_t n_local, const int32_t* global_index, int64_t n_all, int alphabetLength,
                                pyopal_cuda_mdb** out) {
    if (n_local > 0 && (!residues || !offsets || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    return create_rank_impl(device, rank, world, nccl_id, residues, offsets, nullptr, lengths, n_local, global_index, n_all,
                            alphabetLength, out);
}

// The same for a database every rank holds in full on the host (packed): the cell-balanced deal is computed
// here, identically on every rank, and this rank uploads only its own shard.
int pyopal_cuda_mdb_create_rank_dealt(int device, int rank, int world, const unsigned char nccl_id[128],
                                      const unsigned char* residues, const int64_t* offsets, const int32_t* lengths,
                                      int64_t n, int alphabetLength, pyopal_cuda_mdb** out) {
    if (n < 0 || n > INT32_MAX) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "database size out of range");
    if (n > 0 && (!residues || !offsets || !lengths)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    if (world < 1 || rank < 0 || rank >= world) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "bad rank / world");
    if (world == 1)
        return create_rank_impl(device, 0, 1, nccl_id, residues, offsets, nullptr, lengths, n, nullptr, n, alphabetLength, out);
    std::vector<int32_t> owner((size_t)n);
    int rc = pyopal_cuda_deal(lengths, n, world, owner.data());
    if (rc) return rc;
    std::vector<int32_t> mine, sl;
    for (int64_t i = 0; i < n; i++)
        if (owner[(size_t)i] == rank) { mine.push_back((int32_t)i); sl.push_back(lengths[i]); }
    return create_rank_impl(device, rank, world, nccl_id, residues, offsets, mine.data(), sl.data(), (int64_t)mine.size(),
                            mine.data(), n, alphabetLength, out);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Search.  Query i of a call runs on lane i mod 2 of every rank: its shard search, the all-gather and the
// un-permutation are stream-ordered, and nothing waits on the host until the results are copied out -- which
// happens one query late, when the next search is already queued on the device.
namespace {

struct BatchArgs {
    const unsigned char* queries; const int64_t* q_offsets; const int32_t* q_lengths; int nq;
    int go, ge; const int* M; int A, mode;
    int64_t start, end;
    int32_t* scores;          // host [nq * (end - start)] or NULL
    int k; int32_t* top_scores; int32_t* top_idx;   // host [nq * k] or NULL
};

int rank_batch(pyopal_cuda_mdb* m, int L, const BatchArgs& a, bool output, int64_t* launches) {
    Local* l = m->local[(size_t)L].get();
    DeviceGuard guard(l->device);
    const int64_t nout = a.end - a.start;
    const int64_t npad = std::max<int64_t>(m->n_pad, 1);
    // this shard's part of the slice [start, end): mine[] is ascending
    const int64_t lo = std::lower_bound(l->mine.begin(), l->mine.end(), (int32_t)std::min<int64_t>(a.start, INT32_MAX)) - l->mine.begin();
    const int64_t hi = std::lower_bound(l->mine.begin(), l->mine.end(), (int32_t)std::min<int64_t>(a.end, INT32_MAX)) - l->mine.begin();
    const bool whole_shard = lo == 0 && hi == l->db->n;
    const int kk = (int)std::min<int64_t>(a.k, nout);
    std::vector<int32_t> tv, ti;
    if (kk > 0 && output) { tv.resize((size_t)a.nq * kk); ti.resize((size_t)a.nq * kk); }
    // Units (one query, or two searched together: api.cu plan_batch_units) alternate between the two buffer sets;
    // with one lane (the default) both run on lane 0's streams, i.e. the searches are stream-ordered back to back,
    // with two lanes each set has its own streams.  Every rank derives the same units (the collectives must match).
    std::vector<std::pair<int, int>> units;
    plan_batch_units(l->db, a.q_lengths, a.nq, a.mode, units);
    const int nlane = l->db->tune.lanes >= 2 ? NLANE : 1;
    auto copy_out = [&](int u) -> int {
        Lane& ln = l->lane[u % NLANE];
        cudaStream_t gs = l->lane[u % nlane].gstream;
        if (!output) return 0;
        const int qi[2] = {units[(size_t)u].first, units[(size_t)u].second};
        for (int s = 0; s < 2; s++) {
            if (qi[s] < 0) continue;
            const int32_t* src = m->identity ? ln.d_local + (size_t)s * npad + lo : ln.d_final + (size_t)s * (size_t)m->n_all;
            if (kk > 0) {
                CU(cudaMemcpyAsync(tv.data() + (size_t)qi[s] * kk, ln.tk_vals.p + (size_t)s * kk, (size_t)kk * 4, cudaMemcpyDeviceToHost, gs));
                CU(cudaMemcpyAsync(ti.data() + (size_t)qi[s] * kk, ln.tk_idx.p + (size_t)s * kk, (size_t)kk * 4, cudaMemcpyDeviceToHost, gs));
            } else if (a.scores) {
                CU(cudaMemcpyAsync(a.scores + (size_t)qi[s] * (size_t)nout, src, (size_t)nout * sizeof(int32_t), cudaMemcpyDeviceToHost, gs));
            }
        }
        CU(cudaEventRecord(ln.ev_gather, gs));
        return 0;
    };
    const int nu = (int)units.size();
    for (int u = 0; u < nu; u++) {
        Lane& ln = l->lane[u % NLANE];                 // buffers of this unit
        Workspace* ws = l->lane[u % nlane].ws;         // streams / workspace it runs on
        cudaStream_t gs = l->lane[u % nlane].gstream;
        const int qa = units[(size_t)u].first, qb = units[(size_t)u].second;
        const int rows = qb >= 0 ? 2 : 1;
        // d_local / d_final of this lane are about to be overwritten: its previous gather and copy-out must be done
        if (ln.gathered_once) CU(cudaStreamWaitEvent(ws->stream, ln.ev_gather, 0));
        if (hi > lo) {
            bool done = false;
            if (qb >= 0 && whole_shard) {
                int rc = enqueue_dual_search(l->db, ws, a.queries + a.q_offsets[qa], a.q_lengths[qa], a.queries + a.q_offsets[qb],
                                             a.q_lengths[qb], a.go, a.ge, a.M, a.A, ln.d_local, ln.d_local + npad, ws->stream);
                if (rc == 0) { done = true; *launches += ws->stats[0]; l->db->publish(ws); }
                else if (rc != DUAL_NOT_ELIGIBLE) return rc;
            }
            if (!done) {
                const int qs[2] = {qa, qb};
                for (int s = 0; s < 2; s++) {
                    if (qs[s] < 0) continue;
                    int rc = enqueue_score_search(l->db, ws, a.queries + a.q_offsets[qs[s]], a.q_lengths[qs[s]], a.go, a.ge, a.M, a.A,
                                                  a.mode, lo, hi, ln.d_local + (size_t)s * npad + lo, ws->stream);
                    if (rc) return rc;
                    *launches += ws->stats[0];
                    l->db->publish(ws);
                }
            }
        }
        CU(cudaEventRecord(ln.ev_search, ws->stream));
        CU(cudaStreamWaitEvent(gs, ln.ev_search, 0));
        if (!m->identity) {
            // rows of one rank are contiguous ([s][n_pad]), so both rows of a dual search travel in ONE all-gather
            if (m->world > 1) {
                NC(g_nccl.AllGather(ln.d_local, ln.d_gathered, (size_t)(rows * npad), NCCL_INT32, l->comm, gs));
            } else {
                CU(cudaMemcpyAsync(ln.d_gathered, ln.d_local, (size_t)(rows * npad) * sizeof(int32_t), cudaMemcpyDeviceToDevice, gs));
            }
            if (output) {  // only the rank that hands the result out puts it in database order
                const int64_t total = npad * m->world;
                const int blocks = (int)std::max<int64_t>(1, std::min<int64_t>((total + 255) / 256, 148 * 8));
                unpermute_kernel<<<blocks, 256, 0, gs>>>(ln.d_gathered, l->d_dest, total, npad, rows, a.start, a.end, ln.d_final, m->n_all);
                CU(cudaGetLastError());
            }
            *launches += (m->world > 1 ? 1 : 0) + (output ? 1 : 0);
        }
        if (kk > 0 && output) {
            if (ln.tk_ws.reserve(topk_workspace_words()) || ln.tk_vals.reserve((size_t)(2 * kk)) || ln.tk_idx.reserve((size_t)(2 * kk)))
                return PYOPAL_CUDA_ERR_NOMEM;
            for (int s = 0; s < rows; s++) {
                const int32_t* final_row = m->identity ? ln.d_local + (size_t)s * npad + lo : ln.d_final + (size_t)s * (size_t)m->n_all;
                int rc = topk_device(final_row, nout, kk, ln.tk_vals.p + (size_t)s * kk, ln.tk_idx.p + (size_t)s * kk, ln.tk_ws.p,
                                     l->db->sm_count, gs);
                if (rc) return rc;
                *launches += 9;
            }
        }
        CU(cudaEventRecord(ln.ev_gather, gs));
        ln.gathered_once = true;
        if (u > 0) { int rc = copy_out(u - 1); if (rc) return rc; }
    }
    if (nu > 0) { int rc = copy_out(nu - 1); if (rc) return rc; }
    for (Lane& ln : l->lane) CU(cudaStreamSynchronize(ln.gstream));
    if (kk > 0 && output) {
        std::vector<int32_t> order((size_t)kk);
        for (int i = 0; i < a.nq; i++) {
            const int32_t* v = tv.data() + (size_t)i * kk;
            const int32_t* ix = ti.data() + (size_t)i * kk;
            std::iota(order.begin(), order.end(), 0);
            std::sort(order.begin(), order.end(), [&](int32_t x, int32_t y) { return v[x] != v[y] ? v[x] > v[y] : ix[x] < ix[y]; });
            for (int j = 0; j < kk; j++) {
                a.top_scores[(size_t)i * a.k + j] = v[order[(size_t)j]];
                a.top_idx[(size_t)i * a.k + j] = (int32_t)(ix[order[(size_t)j]] + a.start);
            }
        }
    }
    return 0;
}

int mdb_batch(pyopal_cuda_mdb* m, BatchArgs a) {
    if (!m) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null handle");
    if (a.nq < 0 || (a.nq > 0 && (!a.queries || !a.q_offsets || !a.q_lengths || !a.M)))
        return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    if (a.mode < MODE_NW || a.mode > MODE_SW) return fail(OPAL_ERR_INVALID_MODE, "invalid alignment mode");
    if (a.A != m->A) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "alphabet length differs from the store's");
    if (a.go < 0 || a.ge < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "gap penalties must be non-negative");
    if (a.start < 0) a.start = 0;
    if (a.end > m->n_all) a.end = m->n_all;
    if (a.end < a.start) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "slice end is lower than start");
    for (int i = 0; i < a.nq; i++) {
        if (a.q_lengths[i] < 0) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "negative query length");
        const unsigned char* q = a.queries + a.q_offsets[i];
        for (int x = 0; x < a.q_lengths[i]; x++)
            if (q[x] >= a.A) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "query residue code outside the alphabet");
    }
    if (a.end == a.start || a.nq == 0) return 0;
    std::lock_guard<std::mutex> lock(m->mu);
    std::vector<int64_t> launches((size_t)m->nlocal, 0);
    int rc = run_all(m, [&](int L) { return rank_batch(m, L, a, L == 0, &launches[(size_t)L]); });
    m->last_launches = 0;
    for (int64_t x : launches) m->last_launches += x;
    return rc;
}

}  // namespace

extern "C" {

int pyopal_cuda_mdb_search_batch(pyopal_cuda_mdb* m, const unsigned char* queries, const int64_t* q_offsets,
                                 const int32_t* q_lengths, int nq, int gapOpen, int gapExt, const int* scoreMatrix,
                                 int alphabetLength, int mode, int64_t start, int64_t end, int32_t* scores) {
    BatchArgs a{queries, q_offsets, q_lengths, nq, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end, scores, 0, nullptr, nullptr};
    return mdb_batch(m, a);
}

int pyopal_cuda_mdb_search_batch_topk(pyopal_cuda_mdb* m, const unsigned char* queries, const int64_t* q_offsets,
                                      const int32_t* q_lengths, int nq, int gapOpen, int gapExt, const int* scoreMatrix,
                                      int alphabetLength, int mode, int k, int32_t* scores, int32_t* indices) {
    if (k <= 0 || !scores || !indices) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "top-k needs k > 0 and output buffers");
    if (m && k > m->n_all) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "k exceeds the database size");
    BatchArgs a{queries, q_offsets, q_lengths, nq, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, 0, m ? m->n_all : 0, nullptr, k, scores, indices};
    return mdb_batch(m, a);
}

// One query, any search type, host outputs in database order (the multi-GPU form of pyopal_cuda_search).
// SCORE: shard searches + NCCL all-gather.  END / ALIGNMENT (single-process handles): every shard is searched by
// its own host thread and the per-target results are scattered into the caller's arrays -- end cells and
// alignments are per target, nothing is exchanged between the devices.
int pyopal_cuda_mdb_search(pyopal_cuda_mdb* m, const unsigned char* query, int queryLength, int gapOpen, int gapExt,
                           const int* scoreMatrix, int alphabetLength, int searchType, int mode, int overflowMethod,
                           int64_t start, int64_t end, int32_t* scores, int32_t* endQ, int32_t* endT, int32_t* startQ,
                           int32_t* startT, unsigned char* ops, const int64_t* ops_offsets, int32_t* ops_len) {
    if (!m) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null handle");
    if (searchType < OPAL_SEARCH_SCORE || searchType > OPAL_SEARCH_ALIGNMENT) return fail(OPAL_ERR_INVALID_MODE, "invalid search type");
    if (overflowMethod != OPAL_OVERFLOW_SIMPLE && overflowMethod != OPAL_OVERFLOW_BUCKETS) return fail(OPAL_ERR_INVALID_MODE, "invalid overflow method");
    if (searchType == OPAL_SEARCH_SCORE) {
        const int64_t off = 0;
        const int32_t ql = queryLength;
        BatchArgs a{query ? query : (const unsigned char*)"", &off, &ql, 1, gapOpen, gapExt, scoreMatrix, alphabetLength, mode, start, end, scores, 0, nullptr, nullptr};
        if (queryLength > 0 && !query) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
        if (!scores && end > start) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
        return mdb_batch(m, a);
    }
    if (m->world != m->nlocal)
        return fail(PYOPAL_CUDA_ERR_ARGUMENT, "end / full mode on a one-process-per-GPU handle: re-align the hits on one rank");
    if (start < 0) start = 0;
    if (end > m->n_all) end = m->n_all;
    if (end < start) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "slice end is lower than start");
    if (end == start) return 0;
    if (!scores || !endQ || !endT) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null output");
    const bool full = searchType == OPAL_SEARCH_ALIGNMENT;
    if (full && (!startQ || !startT || !ops || !ops_offsets || !ops_len)) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null alignment output");
    std::lock_guard<std::mutex> lock(m->mu);
    std::vector<int64_t> launches((size_t)m->nlocal, 0);
    int rc = run_all(m, [&](int L) -> int {
        Local* l = m->local[(size_t)L].get();
        const int64_t lo = std::lower_bound(l->mine.begin(), l->mine.end(), (int32_t)start) - l->mine.begin();
        const int64_t hi = std::lower_bound(l->mine.begin(), l->mine.end(), (int32_t)std::min<int64_t>(end, INT32_MAX)) - l->mine.begin();
        const size_t cnt = (size_t)(hi - lo);
        if (cnt == 0) return 0;
        std::vector<int32_t> sc(cnt), eq(cnt, -1), et(cnt, -1), sq(cnt, -1), st(cnt, -1), ol(cnt, 0);
        std::vector<int64_t> oo;
        std::vector<unsigned char> lops;
        if (full) {
            oo.resize(cnt);
            int64_t tot = 0;
            for (size_t k = 0; k < cnt; k++) { oo[k] = tot; tot += (int64_t)queryLength + l->db->h_lengths[(size_t)lo + k]; }
            lops.resize((size_t)tot + 1);
        }
        int r = pyopal_cuda_search(l->db, query, queryLength, gapOpen, gapExt, scoreMatrix, alphabetLength, searchType, mode,
                                   overflowMethod, lo, hi, sc.data(), eq.data(), et.data(), sq.data(), st.data(),
                                   full ? lops.data() : nullptr, full ? oo.data() : nullptr, ol.data());
        if (r) return r;
        int64_t st12[12];
        pyopal_cuda_last_stats(l->db, st12, 12);
        launches[(size_t)L] = st12[0];
        for (size_t k = 0; k < cnt; k++) {
            const int64_t g = (int64_t)l->mine[(size_t)lo + k] - start;
            scores[g] = sc[k]; endQ[g] = eq[k]; endT[g] = et[k];
            if (full) {
                startQ[g] = sq[k]; startT[g] = st[k]; ops_len[g] = ol[k];
                if (ol[k] > 0) memcpy(ops + ops_offsets[g], lops.data() + oo[k], (size_t)ol[k]);
            }
        }
        return 0;
    });
    m->last_launches = 0;
    for (int64_t x : launches) m->last_launches += x;
    return rc;
}

// Device-side clock around a sequence of searches: begin() marks every local rank's streams, end() returns the
// longest elapsed time over the local ranks (CUDA events, after everything enqueued so far has completed).
int pyopal_cuda_mdb_time_begin(pyopal_cuda_mdb* m) {
    if (!m) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null handle");
    for (auto& lp : m->local) {
        DeviceGuard guard(lp->device);
        CU(cudaDeviceSynchronize());
        CU(cudaEventRecord(lp->t0, lp->lane[0].ws->stream));
    }
    return 0;
}

int pyopal_cuda_mdb_time_end(pyopal_cuda_mdb* m, double* ms) {
    if (!m || !ms) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null argument");
    double worst = 0;
    for (auto& lp : m->local) {
        DeviceGuard guard(lp->device);
        for (int b = 1; b < NLANE; b++) {
            CU(cudaEventRecord(lp->lane[b].ev_gather, lp->lane[b].gstream));
            CU(cudaStreamWaitEvent(lp->lane[0].gstream, lp->lane[b].ev_gather, 0));
        }
        CU(cudaEventRecord(lp->t1, lp->lane[0].gstream));
        CU(cudaEventSynchronize(lp->t1));
        float e = 0;
        CU(cudaEventElapsedTime(&e, lp->t0, lp->t1));
        worst = std::max(worst, (double)e);
    }
    *ms = worst;
    return 0;
}

int64_t pyopal_cuda_mdb_last_launches(const pyopal_cuda_mdb* m) { return m ? m->last_launches : 0; }

}  // extern "C"
