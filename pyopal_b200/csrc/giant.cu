// pyopal_b200/csrc/giant.cu -- the longest targets: a team of warps per target.
//
// A target of 35 000 residues is a serial chain of 35 000 column steps for the
// lanes that walk it, so in search_kernel (kernels.cuh) a single warp spends
// tens of milliseconds on it while the rest of the database is long done --
// and that time does not shrink when the database is sharded over more GPUs.
// Two things make the chain short here:
//   * the whole query is laid over a TEAM of 1..32 warps, K = 4 or 8 rows per
//     lane, so every target column is one short step: lanes of a warp hand
//     (H - open, F) of their last row to the next lane with SHFL.UP as in
//     search_kernel, and the last lane of a warp hands them to lane 0 of the
//     next warp through a ring buffer in shared memory with producer /
//     consumer counters (the warps form one long systolic wavefront; no
//     CTA-wide barrier in the column loop);
//   * every lane advances a BLOCK of 4 columns per step, so the per-step
//     overhead (shuffles, predicates, ring hand-off) is paid once per 4
//     columns and the 4 x K cells give the lone resident warp of a scheduler
//     independent instructions to issue back to back (a single warp otherwise
//     pays ~5 clocks per instruction: measured 466 clocks per 4-row step
//     before blocking, 143 clocks per column after);
//   * the CTA asks for all of the SM's shared memory, so no throughput CTA of
//     search_kernel shares the SM: a latency-critical warp that competes for
//     a saturated ALU pipe pays a queueing delay on every dependent
//     instruction.  floor(16 / warps per team) teams share the 512-thread CTA
//     and the profile; surplus warps exit at once.
// W = 32: one target per team, 32-bit values (every algorithm, any score).
// W = 16 (round 2, Smith-Waterman): the two targets of a PAIR per team in one
// u16x2 word, exactly the packed arithmetic of search_kernel -- a giant holds a
// whole SM for milliseconds, so walking two of them for the price of one halves
// the SM time the longest targets take from the bulk kernel (6 % of a 1/8 shard's
// SM time for the 5000-aa query).  A pair whose packed score overflows goes to
// the same rescue list as the bulk kernel's.  Same cell update (column_update)
// and profile as search_kernel.
#include "kernels.cuh"

#include <climits>
#include <type_traits>

namespace pyopal {

namespace {
constexpr int RING = 8;          // blocks of slack between consecutive warps (power of two)
constexpr int GIANT_BLOCK = 4;   // columns per step
}

// One ring slot = {H - open, F, tag, -} of one column, moved with a single 16-byte shared-memory
// access, so the tag (block number + 1) validates the data it travels with: no fence, no flag.
__device__ __forceinline__ int4 lds_slot(const int4* p) {
    int4 v;
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("ld.volatile.shared.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts_slot(int4* p, int g, int f, int tag) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(p);
    asm volatile("st.volatile.shared.v4.s32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(g), "r"(f), "r"(tag), "r"(0) : "memory");
}

__device__ __forceinline__ void team_sync(int team, int nw) {
    if (nw == 1) __syncwarp();
    else asm volatile("bar.sync %0, %1;" ::"r"(team + 1), "r"(nw * 32) : "memory");
}

// P.G = warps per team (<= 16), P.npass = teams per CTA (<= 15 when a team has several warps).
// Each lane advances C columns per step (a block): the per-step overhead (shuffles, predicates,
// loop) is paid once per C columns and the C x K cells of a block give the single resident warp
// of a scheduler independent instructions to issue back to back.
template <int K, int C, bool LOCAL, int W = 32, bool TDADD = true>
__global__ void __launch_bounds__(512, 1) giant_kernel(const __grid_constant__ KParams P) {
    static_assert(C == 2 || C == 4, "block of 2 or 4 columns");
    static_assert(W == 32 || LOCAL, "packed giants: Smith-Waterman only");
    typedef Val<W> V;
    typedef typename V::T val_t;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int nw = P.G, teams = P.npass;
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int team = wid / nw, w = wid - team * nw;
    // layout: profile | per team: ring[nw][RING][C] (int4 slots), cons[nw], red[nw], task
    int8_t* sp = reinterpret_cast<int8_t*>(smem_raw);
    const int prof_bytes = ((P.A + 1) * P.pitch + 15) & ~15;
    const size_t team_bytes = ((size_t)nw * RING * C * sizeof(int4) + (size_t)(2 * nw + 4) * sizeof(int) + 15) & ~(size_t)15;
    unsigned char* tbase = smem_raw + prof_bytes + (size_t)team * team_bytes;

    const int nwork = *P.nwork;
    if (nwork <= 0) return;
    {
        const int total = (P.A + 1) * P.pitch;
        for (int x = tid; x < total; x += blockDim.x) {
            int letter = x / P.pitch, trow = x - letter * P.pitch;
            sp[x] = (int8_t)profile_entry(P, letter, trow);
        }
    }
    __syncthreads();
    if (team >= teams) return;  // surplus warps

    int4* ring = reinterpret_cast<int4*>(tbase);
    volatile int* cons = reinterpret_cast<volatile int*>(tbase + (size_t)nw * RING * C * sizeof(int4));
    volatile int* red = cons + nw;
    volatile int* task_s = red + nw;

    const unsigned FULL = 0xffffffffu;
    const bool top_free = (P.mode != MODE_NW);
    const bool left_free = (P.mode == MODE_OV || P.mode == MODE_SW);
    CellConsts<W> CC;
    CC.init(P);
    const val_t NEG = CC.NEG;
    const int B = P.bias;   // W = 16: values are stored biased; 0 at 32 bit
    const int ttid = w * 32 + lane;
    const int rowbase = ttid * K;
    const bool last_lane = (ttid == nw * 32 - 1);  // holds the last query row
    typedef typename std::conditional<C == 4, uint2, uint32_t>::type codes_t;

    for (;;) {
        team_sync(team, nw);  // previous task's shared state no longer read
        if (ttid == 0) *task_s = atomicAdd(P.counter, 1);
        if (ttid < nw) cons[ttid] = 0;
        for (int x = ttid; x < nw * RING * C; x += nw * 32) sts_slot(ring + x, 0, 0, 0);  // no stale tags
        team_sync(team, nw);
        const int task = *task_s;
        if (task >= nwork) break;
        // W = 32: (pair << 1) | half, one target; W = 16: the pair, both targets (len_a >= len_b columns)
        const int entry = P.worklist[task];
        const int pair = (W == 16) ? entry : (entry >> 1), half = (W == 16) ? 0 : (entry & 1);
        const int ncol = P.pair_len[2 * pair + half];
        // blocks of C columns; the stream is padded with the padding letter to a multiple of 8 columns
        const codes_t* blocks = reinterpret_cast<const codes_t*>(P.stream + P.pair_off[pair]);
        const int nblk = (ncol + C - 1) / C;
        const int nsteps = nblk + 31;

        val_t Gl[K], E[K];
#pragma unroll
        for (int r = 0; r < K; ++r) {
            int qi = rowbase + r - P.padtop;
            int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
            Gl[r] = V::splat(hl - P.go + B);
            E[r] = NEG;
        }
        val_t upGprev;
        {
            int qi = rowbase - 1 - P.padtop;
            int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
            upGprev = V::splat(hl - P.go + B);
        }
        val_t Gout[C], Fout[C];
#pragma unroll
        for (int c = 0; c < C; ++c) { Gout[c] = CC.FLOOR; Fout[c] = NEG; }
        val_t best = CC.FLOOR;   // SW: running maximum (biased at 16 bit; bias 0 at 32 bit)
        int res = INT_MIN;
        int cons_seen = 0;
        val_t nxG[C], nxF[C];     // lane 0 of a warp > 0: the block the next step needs, prefetched from the ring
        bool nx_ready = false;
#pragma unroll
        for (int c = 0; c < C; ++c) { nxG[c] = CC.FLOOR; nxF[c] = NEG; }
        codes_t code_nx = codes_t();
        if (lane == 0 && nblk > 0) code_nx = blocks[0];

        for (int s = 0; s < nsteps; ++s) {
            val_t upG[C], upF[C];
#pragma unroll
            for (int c = 0; c < C; ++c) {
                upG[c] = __shfl_up_sync(FULL, Gout[c], 1);
                upF[c] = __shfl_up_sync(FULL, Fout[c], 1);
            }
            const int b = s - lane;
            const bool active = (b >= 0) && (b < nblk);
            const codes_t code = code_nx;
            if (lane == 0 && active) {
                if (w == 0) {
#pragma unroll
                    for (int c = 0; c < C; ++c) {
                        int ht = top_free ? 0 : -(P.go + (b * C + c) * P.ge);
                        upG[c] = V::splat(ht - P.go + B);
                        upF[c] = NEG;
                    }
                } else {
                    const int4* slots = ring + ((w - 1) * RING + (b & (RING - 1))) * C;
                    while (!nx_ready) {  // normally prefetched during the previous step
                        nx_ready = true;
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            const int4 v = lds_slot(slots + c);
                            nxG[c] = (val_t)v.x; nxF[c] = (val_t)v.y;
                            nx_ready = nx_ready && (v.z == b + 1);
                        }
                    }
#pragma unroll
                    for (int c = 0; c < C; ++c) { upG[c] = nxG[c]; upF[c] = nxF[c]; }
                    cons[w] = b;  // blocks < b are consumed: their slots may be overwritten
                    // prefetch block b + 1; its latency hides behind this step's cells
                    nx_ready = false;
                    if (b + 1 < nblk) {
                        const int4* nslots = ring + ((w - 1) * RING + ((b + 1) & (RING - 1))) * C;
                        bool ok = true;
#pragma unroll
                        for (int c = 0; c < C; ++c) {
                            const int4 v = lds_slot(nslots + c);
                            nxG[c] = (val_t)v.x; nxF[c] = (val_t)v.y;
                            ok = ok && (v.z == b + 2);
                        }
                        nx_ready = ok;
                    }
                }
            }
            {
                const int bn = b + 1;
                if (bn >= 0 && bn < nblk) code_nx = blocks[bn];
            }
            if (active) {
                uint32_t cw[2];
                if (C == 4) { cw[0] = reinterpret_cast<const uint32_t*>(&code)[0]; cw[1] = reinterpret_cast<const uint32_t*>(&code)[1]; }
                else { cw[0] = reinterpret_cast<const uint32_t*>(&code)[0]; cw[1] = 0; }
#pragma unroll
                for (int c = 0; c < C; ++c) {
                    const unsigned t0 = (cw[c >> 1] >> (16 * (c & 1) + 8 * half)) & 0xffu;
                    const unsigned t1 = (W == 16) ? ((cw[c >> 1] >> (16 * (c & 1) + 8)) & 0xffu) : 0u;
                    val_t F = upF[c], Gup = upG[c];
                    column_update<K, W, LOCAL, true, false, TDADD>(P, sp, t0, t1, ttid * P.kstride, CC, E, Gl, upGprev, F, Gup, best);
                    Gout[c] = Gup;
                    Fout[c] = F;
                    upGprev = upG[c];
                    if constexpr (!LOCAL) {
                        const int j = b * C + c;  // columns >= ncol of the last block are padding
                        if (last_lane) {
                            const int h = Gup + P.go;  // last query row
                            if (P.mode == MODE_NW) { if (j == ncol - 1) res = h; }
                            else if (j < ncol) res = max(res, h);
                        }
                        if (P.mode == MODE_OV && j == ncol - 1) {
#pragma unroll
                            for (int r = 0; r < K; ++r)
                                if (rowbase + r >= P.padtop) res = max(res, Gl[r] + P.go);
                        }
                    }
                }
                if (lane == 31 && w < nw - 1) {
                    // slot of block b is free once the consumer is past block b - RING
                    while (b - cons_seen >= RING) cons_seen = cons[w + 1];
                    int4* slots = ring + (w * RING + (b & (RING - 1))) * C;
#pragma unroll
                    for (int c = 0; c < C; ++c) sts_slot(slots + c, (int)Gout[c], (int)Fout[c], b + 1);
                }
            }
        }
        // reduce over the team
        if (W == 16) {
            uint32_t v = (uint32_t)best;
            for (int d = 16; d > 0; d >>= 1) v = __vmaxu2(v, __shfl_xor_sync(FULL, v, d));
            if (lane == 0) red[w] = (int)v;
            team_sync(team, nw);
            if (ttid == 0) {
                uint32_t m = (uint32_t)red[0];
                for (int x = 1; x < nw; ++x) m = __vmaxu2(m, (uint32_t)red[x]);
                const int r0 = (int)(m & 0xffffu) - B, r1 = (int)(m >> 16) - B;
                const int ia = P.pair_idx[2 * pair], ib = P.pair_idx[2 * pair + 1];
                const bool ina = ia >= P.out_start && ia < P.out_end;
                const bool inb = ib >= P.out_start && ib < P.out_end;
                // as in search_kernel: an overflowing LOW half carries into the high half, both are rescored then
                const bool ova = r0 >= P.ovf_thresh;
                const bool ovb = r1 >= P.ovf_thresh || ova;
                if (ina) {
                    if (ova) P.ovf_list[atomicAdd(P.ovf_count, 1)] = pair << 1;
                    else P.out[ia - P.out_start] = r0;
                }
                if (inb) {
                    if (ovb) P.ovf_list[atomicAdd(P.ovf_count, 1)] = (pair << 1) | 1;
                    else P.out[ib - P.out_start] = r1;
                }
            }
        } else {
            int v = LOCAL ? (int)best : res;
            for (int d = 16; d > 0; d >>= 1) v = max(v, __shfl_xor_sync(FULL, v, d));
            if (lane == 0) red[w] = v;
            team_sync(team, nw);
            if (ttid == 0) {
                int m = red[0];
                for (int x = 1; x < nw; ++x) m = max(m, red[x]);
                const int ix = P.pair_idx[2 * pair + half];
                if (ix >= P.out_start && ix < P.out_end) P.out[ix - P.out_start] = m;
            }
        }
    }
}

search_kernel_t get_giant_kernel(int K, bool local) {
    switch (K) {
        case 4: return local ? giant_kernel<4, GIANT_BLOCK, true> : giant_kernel<4, GIANT_BLOCK, false>;
        case 8: return local ? giant_kernel<8, GIANT_BLOCK, true> : giant_kernel<8, GIANT_BLOCK, false>;
        case 16: return local ? giant_kernel<16, GIANT_BLOCK, true> : giant_kernel<16, GIANT_BLOCK, false>;
        default: return nullptr;
    }
}

// Smith-Waterman, the two targets of a pair packed at 16 bit (P.worklist holds pair ids)
search_kernel_t get_giant_kernel_packed(int K, bool tdadd) {
    switch (K) {
        case 4: return tdadd ? giant_kernel<4, GIANT_BLOCK, true, 16, true> : giant_kernel<4, GIANT_BLOCK, true, 16, false>;
        case 8: return tdadd ? giant_kernel<8, GIANT_BLOCK, true, 16, true> : giant_kernel<8, GIANT_BLOCK, true, 16, false>;
        case 16: return tdadd ? giant_kernel<16, GIANT_BLOCK, true, 16, true> : giant_kernel<16, GIANT_BLOCK, true, 16, false>;
        default: return nullptr;
    }
}

size_t giant_smem_bytes(int A, int pitch, int warps_per_team, int teams) {
    const size_t team_bytes = ((size_t)warps_per_team * RING * GIANT_BLOCK * sizeof(int4) + (size_t)(2 * warps_per_team + 4) * sizeof(int) + 15) & ~(size_t)15;
    return (size_t)((((A + 1) * pitch + 15) & ~15)) + (size_t)teams * team_bytes + 16;
}

}  // namespace pyopal
