// pyopal_b200/csrc/topk.cu -- the k best scores of a search, selected on the device.
//
// The reference hands back one Python object per target (platform/pyx.in:65-72)
// and the caller sorts them (docs/examples/abc.ipynb cells 23-24: sort, keep the
// top hits, `extract`, re-align in "full" mode).  With the DP at milliseconds,
// moving and sorting 5e5 .. 5e7 scores on the host is what a user would wait on,
// so the selection runs where the scores are: an exact radix select
// (11 + 11 + 10 bits, three histogram passes) finds the k-th largest score t,
// then every score > t and the LOWEST-INDEX scores == t that fill up to k are
// written out in index order (deterministic: ties go to the lower database
// index).  Only k (score, index) pairs leave the device; the
// caller sorts those k.  HBM-bound: five reads of the score array (4 B each)
// = 20 B per target, nothing else scales with n.
#include "store.h"

namespace pyopal {

namespace {

constexpr int TK_THREADS = 256;
constexpr int TK_BINS = 2048;
constexpr int TK_MAX_BLOCKS = 2048;

// order-preserving map int32 -> uint32
__device__ __forceinline__ uint32_t tk_key(int32_t s) { return (uint32_t)s ^ 0x80000000u; }

// state[0] = key prefix found so far, [1] = how many are still wanted among the current bin,
// [2] = threshold key t, [3] = equals to take, [4] = number of keys > t
__global__ void tk_hist_kernel(const int32_t* __restrict__ s, int64_t n, const uint32_t* __restrict__ state, int pass,
                               uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[TK_BINS];
    for (int x = threadIdx.x; x < TK_BINS; x += blockDim.x) sh[x] = 0;
    __syncthreads();
    const uint32_t prefix = state[0];
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const uint32_t u = tk_key(s[i]);
        if (pass == 0) atomicAdd(&sh[u >> 21], 1u);
        else if (pass == 1) { if ((u >> 21) == prefix) atomicAdd(&sh[(u >> 10) & 2047u], 1u); }
        else { if ((u >> 10) == prefix) atomicAdd(&sh[u & 1023u], 1u); }
    }
    __syncthreads();
    for (int x = threadIdx.x; x < TK_BINS; x += blockDim.x)
        if (sh[x]) atomicAdd(&hist[x], sh[x]);
}

// one warp: walk the bins from the top until the wanted count is covered; zero the histogram for the next pass
__global__ void tk_scan_kernel(uint32_t* hist, uint32_t* state, int pass, int k) {
    const int lane = threadIdx.x;
    const int bits = (pass == 2) ? 10 : 11;
    const int nb = 1 << bits;
    const uint32_t want = (pass == 0) ? (uint32_t)k : state[1];
    uint32_t acc = 0;
    int found = -1;
    uint32_t left = 0;
    for (int base = nb - 32; base >= 0 && found < 0; base -= 32) {
        const int b = base + (31 - lane);  // lane 0 looks at the highest bin of the chunk
        const uint32_t c = hist[b];
        uint32_t incl = c;                 // inclusive prefix over lanes 0..lane (descending bins)
        for (int d = 1; d < 32; d <<= 1) {
            const uint32_t v = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += v;
        }
        const unsigned hit = __ballot_sync(0xffffffffu, acc + incl >= want);
        if (hit) {
            const int l = __ffs(hit) - 1;
            found = base + (31 - l);
            const uint32_t before = __shfl_sync(0xffffffffu, incl - c, l);
            left = want - (acc + before);
        } else {
            acc += __shfl_sync(0xffffffffu, incl, 31);
        }
    }
    __syncwarp();
    for (int x = lane; x < TK_BINS; x += 32) hist[x] = 0;
    if (lane == 0) {
        if (found < 0) { found = 0; left = 0; }  // fewer than k items: everything is taken by the caller's clamp
        const uint32_t prefix = (pass == 0) ? (uint32_t)found : ((state[0] << bits) | (uint32_t)found);
        state[0] = prefix;
        state[1] = left;
        if (pass == 2) { state[2] = prefix; state[3] = left; }
    }
}

__global__ void tk_count_kernel(const int32_t* __restrict__ s, int64_t n, const uint32_t* __restrict__ state,
                                int64_t chunk, uint32_t* __restrict__ cnt) {
    const uint32_t t = state[2];
    const int64_t lo = (int64_t)blockIdx.x * chunk, hi = min(n, lo + chunk);
    uint32_t gt = 0, eq = 0;
    for (int64_t i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        const uint32_t u = tk_key(s[i]);
        gt += (u > t);
        eq += (u == t);
    }
    __shared__ uint32_t sg[TK_THREADS / 32], se[TK_THREADS / 32];
    for (int d = 16; d > 0; d >>= 1) {
        gt += __shfl_xor_sync(0xffffffffu, gt, d);
        eq += __shfl_xor_sync(0xffffffffu, eq, d);
    }
    if ((threadIdx.x & 31) == 0) { sg[threadIdx.x >> 5] = gt; se[threadIdx.x >> 5] = eq; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (int w = 0; w < TK_THREADS / 32; w++) { a += sg[w]; b += se[w]; }
        cnt[2 * blockIdx.x] = a;
        cnt[2 * blockIdx.x + 1] = b;
    }
}

// exclusive scan of the per-block counts (one block; nblocks <= TK_MAX_BLOCKS)
__global__ void tk_offsets_kernel(const uint32_t* __restrict__ cnt, int nblocks, uint32_t* __restrict__ off,
                                  uint32_t* state) {
    __shared__ uint32_t sa[TK_MAX_BLOCKS], sb[TK_MAX_BLOCKS];
    for (int x = threadIdx.x; x < nblocks; x += blockDim.x) { sa[x] = cnt[2 * x]; sb[x] = cnt[2 * x + 1]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t a = 0, b = 0;
        for (int x = 0; x < nblocks; x++) {
            const uint32_t ca = sa[x], cb = sb[x];
            sa[x] = a; sb[x] = b;
            a += ca; b += cb;
        }
        state[4] = a;
    }
    __syncthreads();
    for (int x = threadIdx.x; x < nblocks; x += blockDim.x) { off[2 * x] = sa[x]; off[2 * x + 1] = sb[x]; }
}

// every key > t, then the first state[3] keys == t in index order
__global__ void tk_write_kernel(const int32_t* __restrict__ s, int64_t n, const uint32_t* __restrict__ state,
                                int64_t chunk, const uint32_t* __restrict__ off, int k, int32_t* __restrict__ vals,
                                int32_t* __restrict__ idx) {
    const uint32_t t = state[2], need_eq = state[3], total_gt = state[4];
    const int64_t lo = (int64_t)blockIdx.x * chunk, hi = min(n, lo + chunk);
    __shared__ uint32_t wg[TK_THREADS / 32], we[TK_THREADS / 32];
    __shared__ uint32_t run_gt, run_eq;
    if (threadIdx.x == 0) { run_gt = off[2 * blockIdx.x]; run_eq = off[2 * blockIdx.x + 1]; }
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1u;
    for (int64_t base = lo; base < hi; base += blockDim.x) {
        const int64_t i = base + threadIdx.x;
        int32_t v = 0;
        bool g = false, e = false;
        if (i < hi) {
            v = s[i];
            const uint32_t u = tk_key(v);
            g = u > t;
            e = u == t;
        }
        const unsigned bg = __ballot_sync(0xffffffffu, g), be = __ballot_sync(0xffffffffu, e);
        if (lane == 0) { wg[w] = __popc(bg); we[w] = __popc(be); }
        __syncthreads();
        uint32_t pg = run_gt, pe = run_eq;
        for (int x = 0; x < w; x++) { pg += wg[x]; pe += we[x]; }
        if (g) {
            const uint32_t pos = pg + __popc(bg & lt);
            if (pos < (uint32_t)k) { vals[pos] = v; idx[pos] = (int32_t)i; }
        } else if (e) {
            const uint32_t r = pe + __popc(be & lt);
            if (r < need_eq && total_gt + r < (uint32_t)k) { vals[total_gt + r] = v; idx[total_gt + r] = (int32_t)i; }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            uint32_t a = 0, b = 0;
            for (int x = 0; x < TK_THREADS / 32; x++) { a += wg[x]; b += we[x]; }
            run_gt += a; run_eq += b;
        }
        __syncthreads();
    }
}

}  // namespace

size_t topk_workspace_words() { return (size_t)TK_BINS + 16 + 4 * (size_t)TK_MAX_BLOCKS; }

// Select the k best of d_scores[0, n) into d_vals / d_idx (k entries, index order within "greater" then
// "equal"; the caller sorts them).  `ws` holds topk_workspace_words() uint32.  k <= n required.
int topk_device(const int32_t* d_scores, int64_t n, int k, int32_t* d_vals, int32_t* d_idx, uint32_t* ws, int sm_count,
                cudaStream_t st) {
    if (k <= 0 || n <= 0) return 0;
    uint32_t* hist = ws;
    uint32_t* state = ws + TK_BINS;
    uint32_t* cnt = state + 16;
    uint32_t* off = cnt + 2 * TK_MAX_BLOCKS;
    CU(cudaMemsetAsync(ws, 0, (TK_BINS + 16) * sizeof(uint32_t), st));
    const int hgrid = (int)std::max<int64_t>(1, std::min<int64_t>((int64_t)sm_count * 8, (n + TK_THREADS * 8 - 1) / (TK_THREADS * 8)));
    for (int pass = 0; pass < 3; pass++) {
        tk_hist_kernel<<<hgrid, TK_THREADS, 0, st>>>(d_scores, n, state, pass, hist);
        tk_scan_kernel<<<1, 32, 0, st>>>(hist, state, pass, k);
    }
    const int nblocks = (int)std::max<int64_t>(1, std::min<int64_t>(TK_MAX_BLOCKS, (n + TK_THREADS * 16 - 1) / (TK_THREADS * 16)));
    const int64_t chunk = (n + nblocks - 1) / nblocks;
    tk_count_kernel<<<nblocks, TK_THREADS, 0, st>>>(d_scores, n, state, chunk, cnt);
    tk_offsets_kernel<<<1, 256, 0, st>>>(cnt, nblocks, off, state);
    tk_write_kernel<<<nblocks, TK_THREADS, 0, st>>>(d_scores, n, state, chunk, off, k, d_vals, d_idx);
    CU(cudaGetLastError());
    return 0;
}

}  // namespace pyopal
