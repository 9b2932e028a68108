// pyopal_b200/csrc/tgiant.cu -- the longest targets, TRANSPOSED: the target is laid over the lanes.
//
// giant.cu lays the QUERY over a team of warps and streams the target through
// it, so a 35 000-residue target is 35 000 dependent column steps no matter how
// short the query is (2.5 ms for a 100-aa query) -- the floor of every search on
// a small shard.  The DP is symmetric under swapping the two sequences (same
// gap costs in both directions, matrix indexed [query][target] either way), so
// here the roles are swapped: a team of NW warps holds a SLAB of NW*32*K target
// positions (K per lane, H - open and the horizontal gap state in registers),
// the QUERY streams through the team one residue per step as an anti-diagonal
// wavefront, and a target of n residues takes ceil(n / slab) revolutions of
// Lq + NW*32 - 1 steps: 9 x 227 steps instead of 35 213 for the 100-aa query.
//
//   * lane -> lane: SHFL.UP of (H - open, F) of the lane's last target row;
//   * warp -> warp: a ring of 16-byte shared-memory slots {H - open, F, tag}
//     (one 128-bit access, the tag validates the data it travels with, as in
//     giant.cu); the warps free-run, there is no CTA barrier in the loop;
//   * slab -> slab: the last lane's row goes to a shared-memory FIFO of Lq
//     entries that lane 0 of warp 0 reads one revolution later (a progress
//     counter orders it), so the team is a circular pipeline;
//   * every lane builds the profile of its OWN K target positions (int8,
//     score + open, one row per query letter) in shared memory at the start of
//     a slab and reads it back with the same vector LDS as search_kernel: the
//     cell update is kernels.cuh's column_update, 32-bit values, with the
//     one-instruction F chain (the host sends ext > open to giant.cu);
//   * padding rows sit ABOVE target position 0 and are arithmetically
//     transparent exactly like search_kernel's padding above query row 0.
//
// Borders, in the reference's terms (H[i][j], i = query row, j = target column):
// the lanes' initial column is H[-1][j] (0 when the target start is free: HW,
// OV, SW), the row fed to lane 0 is H[i][-1] (0 when the query start is free:
// OV, SW).  Results: SW the running maximum; NW H[Lq-1][n-1]; HW the maximum of
// the last query row; OV that and the last target column.
#include "kernels.cuh"

#include <climits>

namespace pyopal {

namespace {
constexpr int TRING = 8;  // slots between consecutive warps (power of two)

__device__ __forceinline__ int4 t_lds_slot(const int4* p) {
    int4 v;
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void t_sts_slot(int4* p, int g, int f, int tag) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("st.volatile.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(g), "r"(f), "r"(tag), "r"(0) : "memory");
}

struct TLayout {
    size_t prof, ring, bnd, query, mat, ints, total;
};
__host__ __device__ inline TLayout tgiant_layout(int A, int pitch, int warps, int Lq) {
    TLayout L;
    size_t o = 0;
    L.prof = o; o += ((size_t)A * pitch + 15) & ~(size_t)15;
    L.ring = o; o += (size_t)warps * TRING * sizeof(int4);
    L.bnd = o; o += ((size_t)Lq * sizeof(int2) + 15) & ~(size_t)15;
    L.query = o; o += ((size_t)Lq + 15) & ~(size_t)15;
    L.mat = o; o += (size_t)A * 32;
    L.ints = o; o += (size_t)(2 * warps + 4) * sizeof(int);
    L.total = (o + 15) & ~(size_t)15;
    return L;
}
}  // namespace

// P.G = warps of the team (= CTA), P.krows = K, P.kstride / P.pitch = profile layout of the slab,
// P.pad_score = folded score of the padding rows above target position 0.
template <int K, bool LOCAL>
__global__ void __launch_bounds__(256, 1) tgiant_kernel(const __grid_constant__ KParams P) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int NW = P.G;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int Lq = P.Lq, A = P.A;
    const TLayout L = tgiant_layout(A, P.pitch, NW, Lq);
    int8_t* sp = reinterpret_cast<int8_t*>(smem_raw + L.prof);
    int4* ring = reinterpret_cast<int4*>(smem_raw + L.ring);
    volatile int* bnd = reinterpret_cast<volatile int*>(smem_raw + L.bnd);
    uint8_t* sq = smem_raw + L.query;
    int8_t* sm = reinterpret_cast<int8_t*>(smem_raw + L.mat);
    volatile int* cons = reinterpret_cast<volatile int*>(smem_raw + L.ints);
    volatile int* red = cons + NW;
    volatile int* task_s = red + NW;
    volatile int* bprod = task_s + 1;

    const int nwork = *P.nwork;
    if (nwork <= 0) return;
    for (int x = tid; x < Lq; x += blockDim.x) sq[x] = P.query[x];
    for (int x = tid; x < A * 32; x += blockDim.x) {
        const int q = x >> 5, t = x & 31;
        sm[x] = (t < A) ? (int8_t)(P.matrix[q * A + t] + P.go) : (int8_t)0;
    }

    const unsigned FULL = 0xffffffffu;
    const bool top_free = (P.mode != MODE_NW);                        // H[-1][j] = 0: target start free
    const bool left_free = (P.mode == MODE_OV || P.mode == MODE_SW);  // H[i][-1] = 0: query start free
    CellConsts<32> CC;
    CC.init(P);
    const int NEG = CC.NEG;
    const int go = P.go, ge = P.ge;
    const int ttid = w * 32 + lane;
    const int pslot = ttid * P.kstride;
    const int pad8 = P.pad_score & 0xff;

    for (;;) {
        __syncthreads();  // previous task's shared state no longer read; query / matrix staged
        if (tid == 0) { *task_s = atomicAdd(P.counter, 1); *bprod = 0; }
        if (tid < NW) cons[tid] = 0;
        for (int x = tid; x < NW * TRING; x += blockDim.x) t_sts_slot(ring + x, 0, 0, 0);  // no stale tags
        __syncthreads();
        const int task = *task_s;
        if (task >= nwork) break;
        const int entry = P.worklist[task];
        const int pair = entry >> 1, half = entry & 1;
        const int n = P.pair_len[2 * pair + half];
        const uint16_t* tstream = P.stream + P.pair_off[pair];
        const int rows_w = 32 * K;
        const int nw_t = min(NW, (n + rows_w - 1) / rows_w);  // warps this target keeps busy
        const int T = nw_t * rows_w;
        const int nslab = (n + T - 1) / T;
        const int padtop = nslab * T - n;
        const bool last_lane = (ttid == nw_t * 32 - 1);
        int best = 0, res = INT_MIN;
        int cons_seen = 0;

        if (w < nw_t) {
            for (int slab = 0; slab < nslab; ++slab) {
                const int pos0 = slab * T + ttid * K - padtop;  // target position of register 0
                {
                    // int8 profile of this lane's K positions, one row per query letter
                    uint32_t treg[K / 4];
#pragma unroll
                    for (int c = 0; c < K / 4; ++c) {
                        uint32_t tw = 0;
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            const int pos = pos0 + 4 * c + u;
                            unsigned t = 0xffu;
                            if (pos >= 0) t = ((unsigned)tstream[pos] >> (8 * half)) & 0xffu;
                            tw |= t << (8 * u);
                        }
                        treg[c] = tw;
                    }
                    for (int q = 0; q < A; ++q) {
                        const int8_t* row = sm + q * 32;
                        int8_t* dst = sp + (size_t)q * P.pitch + pslot;
#pragma unroll
                        for (int c = 0; c < K / 4; ++c) {
                            uint32_t word = 0;
#pragma unroll
                            for (int u = 0; u < 4; ++u) {
                                const unsigned t = (treg[c] >> (8 * u)) & 0xffu;
                                const unsigned e = (t == 0xffu) ? (unsigned)pad8 : ((unsigned)(int)row[t & 31u] & 0xffu);
                                word |= e << (8 * u);
                            }
                            *reinterpret_cast<uint32_t*>(dst + 4 * c) = word;
                        }
                    }
                }
                int Gl[K], E[K];
#pragma unroll
                for (int r = 0; r < K; ++r) {
                    const int pos = pos0 + r;
                    const int hl = (top_free || pos < 0) ? 0 : -(go + pos * ge);
                    Gl[r] = hl - go;
                    E[r] = NEG;
                }
                int upGprev;
                {
                    const int pos = pos0 - 1;
                    const int hl = (top_free || pos < 0) ? 0 : -(go + pos * ge);
                    upGprev = hl - go;
                }
                int Gout = 0, Fout = NEG;
                const int gbase = slab * Lq;  // rows handed on so far: tags and progress count across slabs
                const int nsteps = Lq + 31;
                for (int s = 0; s < nsteps; ++s) {
                    int upG = __shfl_up_sync(FULL, Gout, 1);
                    int upF = __shfl_up_sync(FULL, Fout, 1);
                    const int i = s - lane;  // query row
                    const bool active = (i >= 0) && (i < Lq);
                    // Hand-offs are polled by the WHOLE warp (every lane reads the same slot: a broadcast), so
                    // the wait loops are warp-uniform and the cell update below is never split by divergence.
                    if (s < Lq) {  // the row lane 0 starts at this step
                        if (w == 0) {
                            if (slab == 0) {
                                if (lane == 0) {
                                    const int hb = left_free ? 0 : -(go + s * ge);
                                    upG = hb - go;
                                    upF = NEG;
                                }
                            } else {
                                const int need = gbase - Lq + s + 1;  // row s of the previous slab
                                while (*bprod < need) {}
                                const int bg = bnd[2 * s], bf = bnd[2 * s + 1];
                                if (lane == 0) { upG = bg; upF = bf; }
                            }
                        } else {
                            const int g0 = gbase + s;
                            const int4* slot = ring + (w - 1) * TRING + (g0 & (TRING - 1));
                            int4 v = t_lds_slot(slot);
                            while (v.z != g0 + 1) v = t_lds_slot(slot);
                            if (lane == 0) { upG = v.x; upF = v.y; cons[w] = g0; }  // rows < g0 are consumed
                        }
                    }
                    if (active) {
                        const unsigned t0 = sq[i];
                        int F = upF, Gup = upG;
                        column_update<K, 32, LOCAL, true, false, true, true>(P, sp, t0, 0u, pslot, CC, E, Gl, upGprev, F, Gup, best);
                        Gout = Gup;
                        Fout = F;
                        upGprev = upG;
                        if (!LOCAL) {
                            if (i == Lq - 1) {  // last query row: every real target position of this lane
                                if (P.mode == MODE_NW) {
                                    if (last_lane && slab == nslab - 1) res = Gl[K - 1] + go;
                                } else {
#pragma unroll
                                    for (int r = 0; r < K; ++r)
                                        if (pos0 + r >= 0) res = max(res, Gl[r] + go);
                                }
                            }
                            // last target column: every query row is a candidate
                            if (P.mode == MODE_OV && last_lane && slab == nslab - 1) res = max(res, Gup + go);
                        }
                    }
                    const int i31 = s - 31;  // the row lane 31 finished at this step
                    if (i31 >= 0 && i31 < Lq) {
                        const int g0 = gbase + i31;
                        if (w < nw_t - 1) {
                            while (g0 - cons_seen >= TRING) cons_seen = cons[w + 1];
                            if (lane == 31) t_sts_slot(ring + w * TRING + (g0 & (TRING - 1)), Gout, Fout, g0 + 1);
                        } else if (slab < nslab - 1) {
                            if (lane == 31) {
                                bnd[2 * i31] = Gout;
                                bnd[2 * i31 + 1] = Fout;
                                __threadfence_block();
                                *bprod = g0 + 1;
                            }
                        }
                    }
                }
            }
        }
        int v = LOCAL ? best : res;
        for (int d = 16; d > 0; d >>= 1) v = max(v, __shfl_xor_sync(FULL, v, d));
        if (lane == 0) red[w] = v;
        __syncthreads();
        if (tid == 0) {
            int m = red[0];
            for (int x = 1; x < nw_t; ++x) m = max(m, red[x]);
            const int ix = P.pair_idx[2 * pair + half];
            if (ix >= P.out_start && ix < P.out_end) P.out[ix - P.out_start] = m;
        }
    }
}

search_kernel_t get_tgiant_kernel(int K, bool local) {
    switch (K) {
        case 16: return local ? tgiant_kernel<16, true> : tgiant_kernel<16, false>;
        case 32: return local ? tgiant_kernel<32, true> : tgiant_kernel<32, false>;
        default: return nullptr;
    }
}

size_t tgiant_smem_bytes(int A, int pitch, int warps, int Lq) { return tgiant_layout(A, pitch, warps, Lq).total; }

}  // namespace pyopal
