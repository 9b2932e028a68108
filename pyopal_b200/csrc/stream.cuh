// pyopal_b200/csrc/stream.cuh -- the score kernel as a CONTINUOUS wavefront (16-bit rung).
//
// search_kernel (kernels.cuh) walks one pair per task: a group of G lanes fills its pipeline (G - 1 steps in which
// part of the group idles), streams the pair's columns, drains, and only then claims the next pair -- per pair and
// per pass.  With narrow groups (G = 4) that is 0.8 % of the steps; with the wide groups a small shard needs
// (8 GPUs: G = 16..32) it is 4-8 %, and it is why the 1/8-shard tile ran at 4.6 TCUPS instead of 5.1.
//
// Here a group never drains.  Every lane walks the same sequence of SEGMENTS -- (pair, pass 0), (pair, pass 1), ...,
// (next pair, pass 0), ... -- one step behind the lane above it, and crosses a segment boundary on its own: it
// resets its K rows (a divergent block of moves, off the ALU pipe), switches to the next pass or to the next pair,
// and goes on with column 0 while the lanes below are still finishing the previous segment.  The hand-off between
// lanes (H - open and F of a lane's last row, SHFL.UP) is the same as inside a segment, because the lane above has
// just done the same thing one step earlier.  What else travels down the SHFL.UP chain, one step per lane:
//   * the running maximum of the pair (Smith-Waterman): a lane adds its own at its last column of the pair, the last
//     lane of the group holds the pair's score and writes it out;
//   * the id of the next pair.  Lane 0 of a group claims pairs from the task counter one pair AHEAD, every lane loads
//     the next pair's descriptor itself when it starts the pair before (the load's latency hides behind a pair).
// Pairs are claimed one at a time, longest first, so the end of the kernel is as balanced as in search_kernel.
//
// Several passes: the last lane of the group writes the pass boundary (H - open, F per column) while lane 0 is
// already reading it for the next pass of the same pair, G - 1 + 1 columns behind; a segment is therefore at least
// G + 2 steps long (columns past the pair's end are idle) and every step ends with __syncwarp(), which orders the
// group's global-memory accesses.
//
// Scope: whole-database SCORE searches (pair = first_pair + task: no work list to chase), profile in shared
// memory, short F chain (ext <= open), stream shorter than 2^31 columns; everything else stays on search_kernel.
// Same cell update (column_update), same profile, same overflow list, bit-identical results.
#pragma once
#include "kernels.cuh"

#include <climits>

namespace pyopal {

template <int K, bool LOCAL, bool TDADD>
__global__ void __launch_bounds__(512) stream_kernel(const __grid_constant__ KParams P) {
    static_assert(K % 4 == 0, "K must be a multiple of 4");
    typedef Val<16> V;
    typedef uint32_t val_t;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int nwork = *P.nwork;
    if (nwork <= 0) return;
    int8_t* prof = reinterpret_cast<int8_t*>(smem_raw);
    {
        const int total = (P.A + 1) * P.pitch;
        for (int x = threadIdx.x; x < total; x += blockDim.x) {
            int letter = x / P.pitch, trow = x - letter * P.pitch;
            prof[x] = (int8_t)profile_entry(P, letter, trow);
        }
        __syncthreads();
    }

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int G = P.G;
    const int p = lane & (G - 1);
    const int T = G * K;
    const bool top_free = LOCAL || (P.mode != MODE_NW);
    const bool left_free = LOCAL || (P.mode == MODE_OV);
    const int B = P.bias;
    CellConsts<16> C;
    C.init(P);
    const val_t NEG = C.NEG, FLOOR = C.FLOOR;
    const int npass = P.npass, lastpass = P.npass - 1;
    const int minseg = (npass > 1) ? G + 2 : 1;
    const uint32_t* const off_lo = reinterpret_cast<const uint32_t*>(P.pair_off);  // low words: stream < 2^31 columns

    // ---- claim the first pairs of this group -----------------------------------------------------------------
    // One at a time, each claim waited for before the next is made: the groups that start together then take the
    // (length-sorted) pairs round-robin -- group g gets pairs g, g + NG, g + 2 NG -- instead of three of the longest
    // each, and a CTA that starts late (its SM held a giant) gets what is left, like any later claim.
    unsigned off = 0, nx_off = 0;
    int ncol = 0, len1 = 0, nx_ncol = 0, nx_len1 = 0;
    int t_cur = 0;
    if (p == 0) t_cur = atomicAdd(P.counter, 1);
    t_cur = __shfl_sync(FULL, t_cur, 0, G);
    if (t_cur < nwork) {
        const int pr = P.first_pair + t_cur;
        off = off_lo[2 * pr]; ncol = P.pair_len[2 * pr]; len1 = P.pair_len[2 * pr + 1];
    }
    int t_nx = 0;               // pair after the current one (descriptor in nx_*)
    if (p == 0) t_nx = atomicAdd(P.counter, 1);
    t_nx = __shfl_sync(FULL, t_nx, 0, G);
    if (t_nx < nwork) {
        const int pr = P.first_pair + t_nx;
        nx_off = off_lo[2 * pr]; nx_ncol = P.pair_len[2 * pr]; nx_len1 = P.pair_len[2 * pr + 1];
    }
    int out_t = t_nx;           // what the lane below picks up at its next pair boundary
    bool alive = t_cur < nwork;
    int seglen = alive ? max(ncol, minseg) : INT_MAX;
    int pass = 0;
    int j = -p;                 // column of the current segment this lane works on at this step
    int pslot = p * P.kstride;
    int rowbase = p * K;

    val_t Gl[K], E[K];
    val_t upGprev;
    auto reset_rows = [&]() {
#pragma unroll
        for (int r = 0; r < K; ++r) {
            const int qi = rowbase + r - P.padtop;
            const int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
            Gl[r] = V::splat(hl - P.go + B);
            E[r] = NEG;
        }
        const int qi = rowbase - 1 - P.padtop;
        const int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
        upGprev = V::splat(hl - P.go + B);
    };
    reset_rows();

    val_t best = FLOOR;        // SW: running maximum of this lane's rows over the current pair
    val_t out_best = FLOOR;    // SW: maximum of lanes 0..p over the pair this lane finished last
    int res0 = INT_MIN, res1 = INT_MIN;   // NW / HW / OV (true values), as in search_kernel
    int out_r0 = INT_MIN, out_r1 = INT_MIN;
    val_t Gout = FLOOR, Fout = NEG;
    unsigned code_nx = 0;
    val_t bG_nx = FLOOR, bF_nx = NEG;
    if (p == 0 && alive) code_nx = P.stream[off];

    for (;;) {
        if (!__any_sync(FULL, alive)) break;
        val_t upG = __shfl_up_sync(FULL, Gout, 1, G);
        val_t upF = __shfl_up_sync(FULL, Fout, 1, G);
        const int in_t = __shfl_up_sync(FULL, out_t, 1, G);
        val_t in_best = FLOOR;
        int in_r0 = INT_MIN, in_r1 = INT_MIN;
        if (LOCAL) in_best = __shfl_up_sync(FULL, out_best, 1, G);
        else if (P.mode == MODE_OV) {
            in_r0 = __shfl_up_sync(FULL, out_r0, 1, G);
            in_r1 = __shfl_up_sync(FULL, out_r1, 1, G);
        }
        // a lane without a pair has ncol = 0 and seglen = INT_MAX: never active, never at a boundary
        const bool active = (unsigned)j < (unsigned)ncol;
        const bool last = (j + 1 == seglen);   // this step ends the lane's segment
        const unsigned code = code_nx;
        if (p == 0) {
            if (pass == 0) {
                const int ht = top_free ? 0 : -(P.go + j * P.ge);
                upG = V::splat(ht - P.go + B);
                upF = NEG;
            } else {
                upG = bG_nx;
                upF = bF_nx;
            }
        }
        // software pipeline: the residue pair and the pass boundary of the NEXT step are requested before this
        // step's cells are computed (a whole step of latency to hide them).  Almost always that is the next column
        // of the same segment; at the end of a segment it is column 0 of the next pass or of the next pair (the
        // lane's state itself advances after the cells, below).
        if (!last) {
            const unsigned jn = (unsigned)(j + 1);
            if (jn < (unsigned)ncol) {
                const unsigned x = off + jn;
                code_nx = P.stream[x];
                if (p == 0 && pass != 0) {
                    bG_nx = __ldcg(P.bndH + x);
                    bF_nx = __ldcg(P.bndF + x);
                }
            }
        } else if (pass != lastpass) {
            code_nx = P.stream[off];
            if (p == 0) {
                bG_nx = __ldcg(P.bndH + off);
                bF_nx = __ldcg(P.bndF + off);
            }
        } else if (t_nx < nwork) {
            code_nx = P.stream[nx_off];
        }
        if (active) {
            const unsigned t0 = code & 0xffu, t1 = code >> 8;
            val_t F = upF, Gup = upG;
            column_update<K, 16, LOCAL, true, false, TDADD, true, false, false>(P, prof, t0, t1, pslot, C, E, Gl, upGprev, F, Gup, best);
            Gout = Gup;
            Fout = F;
            upGprev = upG;
            if (p == G - 1) {
                if (pass != lastpass) {
                    const unsigned x = off + (unsigned)j;
                    __stcg(P.bndH + x, (uint32_t)Gout);
                    __stcg(P.bndF + x, (uint32_t)Fout);
                } else if (!LOCAL) {
                    const int h0 = V::lo(Gout) + P.go - B, h1 = V::hi(Gout) + P.go - B;  // last query row: H = G + open
                    if (P.mode == MODE_NW) {
                        if (j == ncol - 1) res0 = h0;
                        if (j == len1 - 1) res1 = h1;
                    } else {
                        res0 = max(res0, h0);
                        if (j < len1) res1 = max(res1, h1);
                    }
                }
            }
            if (!LOCAL && P.mode == MODE_OV && (j == ncol - 1 || j == len1 - 1)) {
                // last target column: every real query row is a candidate
#pragma unroll
                for (int r = 0; r < K; ++r) {
                    if (rowbase + r >= P.padtop) {
                        if (j == ncol - 1) res0 = max(res0, V::lo(Gl[r]) + P.go - B);
                        if (j == len1 - 1) res1 = max(res1, V::hi(Gl[r]) + P.go - B);
                    }
                }
            }
        }
        // ---- advance: next column, next pass or next pair ----------------------------------------------------
        ++j;
        if (last) {
            j = 0;
            if (pass != lastpass) {
                ++pass;
            } else {
                // this lane is done with the pair: fold its part of the result into what came down the chain
                const int pr = P.first_pair + t_cur;
                if (LOCAL) {
                    out_best = (p == 0) ? best : __vmaxu2(best, in_best);
                    best = FLOOR;
                } else {
                    out_r0 = (p == 0 || P.mode != MODE_OV) ? res0 : max(res0, in_r0);
                    out_r1 = (p == 0 || P.mode != MODE_OV) ? res1 : max(res1, in_r1);
                    res0 = res1 = INT_MIN;
                }
                if (p == G - 1) {
                    const int r0 = LOCAL ? V::lo(out_best) - B : out_r0;
                    const int r1 = LOCAL ? V::hi(out_best) - B : out_r1;
                    const int ia = P.pair_idx[2 * pr], ib = P.pair_idx[2 * pr + 1];
                    // an overflowing LOW half carries into the high half: both targets are rescored then
                    const bool ova = LOCAL && r0 >= P.ovf_thresh;
                    const bool ovb = LOCAL && (r1 >= P.ovf_thresh || ova);
                    if (ia >= 0) {
                        if (ova) P.ovf_list[atomicAdd(P.ovf_count, 1)] = pr << 1;
                        else P.out[ia - P.out_start] = r0;
                    }
                    if (ib >= 0) {
                        if (ovb) P.ovf_list[atomicAdd(P.ovf_count, 1)] = (pr << 1) | 1;
                        else P.out[ib - P.out_start] = r1;
                    }
                }
                // next pair: its descriptor was loaded a pair ago; start loading the one after it
                t_cur = t_nx;
                off = nx_off; ncol = nx_ncol; len1 = nx_len1;
                alive = t_cur < nwork;
                if (!alive) ncol = 0;
                // lane 0 claims the pair after the next one here and waits for the answer (a microsecond per pair and
                // group; holding a second claimed pair instead would only lengthen the tail of the kernel); the lanes
                // below pick the id up from the chain
                t_nx = (p == 0) ? atomicAdd(P.counter, 1) : in_t;
                out_t = t_nx;
                if (t_nx < nwork) {
                    const int nr = P.first_pair + t_nx;
                    nx_off = off_lo[2 * nr]; nx_ncol = P.pair_len[2 * nr]; nx_len1 = P.pair_len[2 * nr + 1];
                }
                seglen = alive ? max(ncol, minseg) : INT_MAX;
                pass = 0;
            }
            rowbase = pass * T + p * K;
            pslot = (pass * G + p) * P.kstride;
            reset_rows();
        }
        if (npass > 1) __syncwarp();
    }
}

// kernels_stream_l*.cu; nullptr when the combination is not built
search_kernel_t get_stream_kernel_l1(int K, bool tdadd);
search_kernel_t get_stream_kernel_l0(int K, bool tdadd);

}  // namespace pyopal
