// pyopal_b200/csrc/dual.cuh -- TWO QUERIES per u16x2 word (Smith-Waterman score, batched searches).
//
// search_kernel packs two TARGETS into the halves of a register; the query profile is int8, so every cell pair
// pays one PRMT to widen (S[q][a], S[q][b]) into a u16x2 word -- one of its 5.5 ALU-pipe instructions.  When
// several queries are searched against the same database (the reference's pattern is a ThreadPool over
// Aligner.align, /root/reference/README.md:116-144; BASELINE.json config 5 is 64 queries per batch) the two
// halves can carry two QUERIES against ONE target instead: both see the same residue, so the profile can hold the
// pair of folded scores pre-packed -- prof[letter][row] = (S_A[row][letter] + open) | (S_B[row][letter] + open) << 16
// -- and the widening disappears: 4.5 ALU-pipe instructions per cell pair, one LDS.128 per four rows.
//
// Everything else is search_kernel's: G lanes x K rows per tile, anti-diagonal wavefront, two sweeps per column,
// passes through a boundary buffer, packed running maximum, overflow list (one per query; an overflowing low half
// carries into the high half, so query B is rescored whenever query A is).  The two queries need not have the
// same length: rows are aligned at the BOTTOM of the tile and the shorter query has more padding rows above row 0
// (score 0 keeps a local alignment at H = 0 there), so the cost is that of the longer one -- the host pairs queries
// of neighbouring length.  Work item = one target: (pair, half) of the pair-interleaved store, taken implicitly
// (task t -> pair first_pair + t / 2, half t % 2: whole-database searches only), so a pair's two targets are walked
// by different groups and the pass boundary is indexed per target (2 * column + half, the 32-bit rung's buffers).
//
// Scope: Smith-Waterman, score mode, short F chain (ext <= open), 32-bit profile words in shared memory (queries up
// to ~1500 rows); the giants, the rescue of overflows and everything else run per query on the other kernels.
#pragma once
#include "kernels.cuh"

#include <climits>

namespace pyopal {

// folded score of tile row `trow` for one of the two queries (see profile_entry)
__device__ __forceinline__ int dual_entry(const KParams& P, const uint8_t* query, int Lq, int padtop, int letter, int trow) {
    const int qi = trow - padtop;
    if (qi < 0 || letter >= P.A) return P.pad_score;
    if (qi >= Lq) return 0;
    return P.matrix[(int)query[qi] * P.A + letter] + P.go;
}

template <int K, bool TDADD>
__global__ void __launch_bounds__(512) dual_kernel(const __grid_constant__ KParams P) {
    static_assert(K % 4 == 0, "K must be a multiple of 4");
    typedef Val<16> V;
    typedef uint32_t val_t;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int nwork = *P.nwork;   // targets: 2 x pairs
    if (nwork <= 0) return;
    // profile: one word per (letter, slot); P.pitch and P.kstride count words, a lane's chunk is K words
    uint32_t* prof = reinterpret_cast<uint32_t*>(smem_raw);
    {
        const int total = (P.A + 1) * P.pitch;
        for (int x = threadIdx.x; x < total; x += blockDim.x) {
            const int letter = x / P.pitch, slot = x - letter * P.pitch;
            const int chunk = slot / P.kstride, within = slot - chunk * P.kstride;
            uint32_t w = 0;
            if (within < P.krows) {
                const int trow = chunk * P.krows + within;
                const int a = dual_entry(P, P.query, P.Lq, P.padtop, letter, trow);
                const int b = dual_entry(P, P.query2, P.Lq2, P.padtop2, letter, trow);
                w = (uint32_t)(uint16_t)(int16_t)a | ((uint32_t)(uint16_t)(int16_t)b << 16);
            }
            prof[x] = w;
        }
        __syncthreads();
    }

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int G = P.G;
    const int p = lane & (G - 1);
    const int grp = lane >> P.logG;
    const int gpw = 32 >> P.logG;
    const int T = G * K;
    const int ntasks = (nwork + gpw - 1) / gpw;
    const int B = P.bias;
    CellConsts<16> C;
    C.init(P);
    const val_t NEG = C.NEG, FLOOR = C.FLOOR;
    const int lastpass = P.npass - 1;

    for (;;) {
        int task = 0;
        if (lane == 0) task = atomicAdd(P.counter, 1);
        task = __shfl_sync(FULL, task, 0);
        if (task >= ntasks) break;

        const int work = task * gpw + grp;
        const bool have = work < nwork;
        const int pair = P.first_pair + (work >> 1);
        const int half = work & 1;
        const int64_t off = have ? P.pair_off[pair] : 0;
        const int ncol = have ? P.pair_len[2 * pair + half] : 0;   // 0 when the pair has no second target
        const int nsteps = __reduce_max_sync(FULL, ncol) + G - 1;
        const uint16_t* const sp = P.stream + off;
        const int sh = 8 * half;
        uint32_t* const bH = P.bndH + (2 * off + half);   // two entries per column: one per target of the pair
        uint32_t* const bF = P.bndF + (2 * off + half);

        val_t best = FLOOR;
        for (int pass = 0; pass < P.npass; ++pass) {
            val_t Gl[K], E[K];
#pragma unroll
            for (int r = 0; r < K; ++r) {
                Gl[r] = V::splat(B - P.go);   // local alignment: free left border
                E[r] = NEG;
            }
            val_t Gout = FLOOR, Fout = NEG, upGprev = V::splat(B - P.go);
            const int pslot = (pass * G + p) * P.kstride;
            unsigned code_nx = 0;
            val_t bG_nx = FLOOR, bF_nx = NEG;
            if (p == 0 && ncol > 0) {
                code_nx = sp[0];
                if (pass != 0) {
                    bG_nx = __ldcg(bH);
                    bF_nx = __ldcg(bF);
                }
            }
            for (int s = 0; s < nsteps; ++s) {
                val_t upG = __shfl_up_sync(FULL, Gout, 1, G);
                val_t upF = __shfl_up_sync(FULL, Fout, 1, G);
                const int j = s - p;
                const bool active = (unsigned)j < (unsigned)ncol;
                const unsigned code = code_nx;
                if (p == 0) {
                    if (pass == 0) {
                        upG = V::splat(B - P.go);   // free top border
                        upF = NEG;
                    } else {
                        upG = bG_nx;
                        upF = bF_nx;
                    }
                }
                {
                    const int jn = j + 1;
                    if ((unsigned)jn < (unsigned)ncol) {
                        code_nx = sp[jn];
                        if (p == 0 && pass != 0) {
                            bG_nx = __ldcg(bH + 2 * jn);
                            bF_nx = __ldcg(bF + 2 * jn);
                        }
                    }
                }
                if (active) {
                    const unsigned t0 = (code >> sh) & 0xffu;
                    val_t F = upF, Gup = upG;
                    column_update<K, 16, true, true, false, TDADD, true, false, false, true>(
                        P, reinterpret_cast<const int8_t*>(prof), t0, 0u, pslot, C, E, Gl, upGprev, F, Gup, best);
                    Gout = Gup;
                    Fout = F;
                    upGprev = upG;
                    if (p == G - 1 && pass != lastpass) {
                        __stcg(bH + 2 * j, Gout);
                        __stcg(bF + 2 * j, Fout);
                    }
                }
            }
            __syncwarp();
        }

        // the group's maxima: low half = query A, high half = query B
        for (int d = 1; d < G; d <<= 1) best = __vmaxu2(best, __shfl_xor_sync(FULL, best, d));
        if (have && p == 0 && ncol > 0) {
            const int ix = P.pair_idx[2 * pair + half];
            if (ix >= P.out_start && ix < P.out_end) {
                const int r0 = V::lo(best) - B, r1 = V::hi(best) - B;
                // an overflowing LOW half carries into the high half through the 32-bit adds
                const bool ova = r0 >= P.ovf_thresh;
                const bool ovb = r1 >= P.ovf_thresh || ova;
                const int e = (pair << 1) | half;
                if (ova) P.ovf_list[atomicAdd(P.ovf_count, 1)] = e;
                else P.out[ix - P.out_start] = r0;
                if (ovb) P.ovf_list2[atomicAdd(P.ovf_count2, 1)] = e;
                else P.out2[ix - P.out_start] = r1;
            }
        }
        (void)T;
    }
}

}  // namespace pyopal
