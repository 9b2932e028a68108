// pyopal_b200/csrc/pairdp.cu -- end-location and alignment (full) modes.
//
// Replaces what opalSearchDatabase does for searchType = OPAL_SEARCH_SCORE_END
// and OPAL_SEARCH_ALIGNMENT (/root/reference/src/pyopal/opal.pxd:40-54,
// lib.pyx:73-77, doc lib.pyx:1278-1282; result fields read back at
// lib.pyx:783-1119).  Three device passes, all 32-bit, one warp per target:
//
//   forward : score + end cell (first maximum in the mode's examination
//             order: SW column-major over all cells, HW last row, OV last row
//             then last column, NW the corner);
//   reverse : the same kernel on the reversed prefixes query[0..endQ],
//             target[0..endT], anchored at the end cell (both borders
//             penalised, no floor) -> start cell;
//   fill    : global DP over the [start..end] rectangle writing one direction
//             byte per cell, then one thread per target walks it backwards
//             and emits the operations (M=0 D=1 I=2 X=3, lib.pyx:991-996).
//
// A warp holds 32*K consecutive query rows (K per lane, K = 4 .. 32 picked from
// the query length) in registers and the target streams through it as an
// anti-diagonal wavefront (two shuffles per step for the carried row, one for
// the prefetched residue); longer queries take several passes whose last row
// round-trips through global memory.  Tie-breaks are those of
// oracle/opal_oracle.c.
#include "store.h"

#include <algorithm>
#include <climits>
#include <cstring>

using namespace pyopal;

namespace {

enum : int { END_CORNER = 0, END_LASTROW = 1, END_LASTROW_LASTCOL = 2, END_ALL = 3 };
constexpr int NEGV = -(1 << 29);

struct PairJob {
    int64_t tbase;   // first column of the target in the u16 stream
    int64_t tb_off;  // fill: byte offset of this job's direction matrix
    int32_t half;    // 0: low byte of each u16, 1: high byte
    int32_t q0, qdir, m;  // DP row i    <-> query[q0 + qdir * i]
    int32_t t0, tdir, n;  // DP column c <-> target[t0 + tdir * c]
    int32_t expect;  // reverse / fill: the forward score
    int32_t slot;    // where the results go
    int32_t ops_end; // walk: one past the last byte of this job's ops slot
};

struct PairParams {
    const uint16_t* stream;
    const uint8_t* query;
    const int32_t* matrix;
    int A, go, ge;
    const PairJob* jobs;
    int njobs;
    int32_t* counter;
    int top_free, left_free, floor0, endrule, anchored;
    int32_t* out_score;
    int32_t* out_q;
    int32_t* out_t;
    int32_t* bndH;
    int32_t* bndF;
    uint8_t* tb;
    int32_t* err;
    // bands: a job whose query needs `maxpass` > 1 passes is walked by that many warps at once, pass p one chunk of
    // columns behind pass p - 1 (task = job * maxpass + pass)
    int maxpass;
    int32_t* prog;   // [task]  columns of this band whose last row is in bndH / bndF
    int32_t* done;   // [job]   bands finished
    int32_t* cut;    // [job]   reverse pass: columns still worth walking (INT_MAX until `expect` is met)
    int32_t* slots;  // [task][5] per-band candidates, merged by the band that finishes last
};

__device__ __forceinline__ int ld_volatile(const int32_t* p) { return *reinterpret_cast<const volatile int32_t*>(p); }

__device__ __forceinline__ int border_val(int is_free, int k, int go, int ge) {
    if (k < 0) return 0;
    return is_free ? 0 : -(go + k * ge);
}

// (value desc, column asc, row asc)
__device__ __forceinline__ bool better(int v, int c, int i, int bv, int bc, int bi) {
    return v > bv || (v == bv && (c < bc || (c == bc && i < bi)));
}

// KIND_ALL: locate, every cell is a candidate (SW, forward and reverse); KIND_EDGE: locate on the last row / last
// column / corner (NW, HW, OV, forward and reverse); KIND_FILL: fill the direction matrix.
enum : int { KIND_ALL = 0, KIND_FILL = 1, KIND_EDGE = 2 };

// One warp per job.  Values are kept as G = H - open (the "gap already opened" form of kernels.cuh), the score table
// holds S + open, and the vertical gap is carried as f[r+1] = max(f[r] - min(open, ext), max(diag, e)[r] - open):
// identical to F[r+1] = max(F[r] - ext, H[r] - open) with H[r] = max(diag, e, F[r]), but only ONE dependent
// operation per row stays on the lane's serial chain - a long target is walked by a single warp, so the chain and
// the instruction count per row are what bound it.  SW's "first maximum" is found from a per-column maximum (half
// an instruction per row); the row is looked up only in the rare step that improves the lane's maximum.  Target
// residues and pass-boundary rows are fetched 32 columns at a time, one chunk ahead, and handed to the lane that
// needs them by shuffle, so no global load sits on the chain.
template <int K, int KIND, int C>
__global__ void __launch_bounds__(128) pair_kernel(const __grid_constant__ PairParams P) {
    static_assert(K % 2 == 0 && (KIND != KIND_FILL || K % 4 == 0), "the fill kernel packs four rows per 32-bit store");
    static_assert(C == 1 || C == 2 || C == 4, "columns per step");
    __shared__ int sS[33 * 32];
    __shared__ int2 sbnd[128 * C];  // per warp: the 32 C boundary columns of the current chunk
    const int go = P.go, ge = P.ge;
    const int gmin = min(go, ge);
    for (int x = threadIdx.x; x < P.A * P.A; x += blockDim.x) sS[x] = P.matrix[x] + go;
    for (int x = threadIdx.x; x < P.A; x += blockDim.x) sS[P.A * P.A + x] = NEGV;  // row of the padding rows
    __syncthreads();

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    constexpr int T = 32 * K;
    const int floorv = P.floor0 ? 0 : 2 * NEGV;

    for (;;) {
        int task = 0;
        if (lane == 0) task = atomicAdd(P.counter, 1);
        task = __shfl_sync(FULL, task, 0);
        if (task >= P.njobs * P.maxpass) break;
        const int job = task / P.maxpass, pass = task - job * P.maxpass;
        const PairJob jb = P.jobs[job];
        const int m = jb.m, n = jb.n;
        if (m <= 0 || n <= 0) continue;
        const int npass = (m + T - 1) / T;
        if (pass >= npass) continue;
        const uint8_t* tbytes = reinterpret_cast<const uint8_t*>(P.stream + jb.tbase) + jb.half;
        int32_t* bH = P.bndH + 2 * jb.tbase + jb.half;
        int32_t* bF = P.bndF + 2 * jb.tbase + jb.half;
        const int rr = (m - 1) - (npass - 1) * T;  // tile row of the last query row in the last pass
        const int owner = rr / K, rl = rr % K;
        const bool multi = npass > 1;
        const bool cutting = (KIND == KIND_ALL) && P.anchored;
        const int32_t* prog_prev = P.prog + task - 1;  // read when pass > 0 only
        int32_t* prog_mine = P.prog + task;
        int32_t* cutp = P.cut + job;

        // trackers
        int gv = P.floor0 ? 0 : INT_MIN, gc = -1, gi = -1;              // KIND_ALL
        int lv = INT_MIN, lc = -1;                                         // last row
        int cv = INT_MIN, ci = -1;                                         // last column
        if (P.anchored) lv = border_val(0, m - 1, go, ge);
        int ncols = n;  // may shrink once the reverse pass has met `expect`
        int corner = 0;

        {
            const int i0 = pass * T + lane * K;
            const int nv = min(max(m - i0, 0), K);
            int qa[K], Gl[K], El[K];
#pragma unroll
            for (int r = 0; r < K; ++r) {
                const int i = i0 + r;
                // byte offset of the query letter's row in sS (padding rows: the -inf row)
                qa[r] = 4 * ((r < nv) ? (int)P.query[jb.q0 + jb.qdir * i] * P.A : P.A * P.A);
                Gl[r] = border_val(P.left_free, i, go, ge) - go;
                El[r] = NEGV;
            }
            int Gd0 = border_val(P.left_free, i0 - 1, go, ge) - go;
            int Gout[C], Fout[C];
#pragma unroll
            for (int cc = 0; cc < C; ++cc) { Gout[cc] = 0; Fout[cc] = NEGV; }
            // best G of this lane in this band (KIND_ALL); a local score of 0 is "no alignment", never a candidate
            int pvg = P.floor0 ? -go : 4 * NEGV, pc = -1, pi = -1;
            const bool last = (pass == npass - 1);
            const bool own = last && (lane == owner);
            // a step takes a lane through a group of C consecutive columns; lane l is one group behind lane l - 1
            auto steps_for = [](int cols) { return (cols + C - 1) / C + 31; };
            int nsteps = steps_for(ncols);
            uint32_t tbw[(K + 3) / 4];

            // chunked prefetch: chunk k = groups [32k, 32k + 32), this lane holds group 32k + lane (C residues, packed)
            auto load_res = [&](int grp) -> uint32_t {
                uint32_t w = 0;
#pragma unroll
                for (int cc = 0; cc < C; ++cc) {
                    const int col = C * grp + cc;
                    if (col < n) w |= (uint32_t)tbytes[2 * (int64_t)(jb.t0 + jb.tdir * col)] << (8 * cc);
                }
                return w;
            };
            // boundary chunk of the band above (32 C columns, this lane takes columns lane, lane + 32, ...): wait until
            // it has published them, or the job was cut short (then the columns past the cut are never walked)
            auto load_bnd = [&](int firstgrp, int (&h)[C], int (&f)[C]) {
                int need = min(C * (firstgrp + 32), ncols);
                while (ld_volatile(prog_prev) < need) {
                    if (cutting) {
                        const int cc = ld_volatile(cutp);
                        if (cc < ncols) { ncols = cc; nsteps = min(nsteps, steps_for(ncols)); need = min(need, ncols); }
                    }
                    __nanosleep(40);
                }
                __threadfence();
#pragma unroll
                for (int cc = 0; cc < C; ++cc) {
                    const int col = C * firstgrp + lane + 32 * cc;
                    if (col < need) { h[cc] = __ldcg(bH + 2 * (int64_t)col); f[cc] = __ldcg(bF + 2 * (int64_t)col); }
                }
            };
            // the same without waiting: false when the band above has not got that far yet
            auto try_bnd = [&](int firstgrp, int (&h)[C], int (&f)[C]) -> bool {
                const int need = min(C * (firstgrp + 32), ncols);
                if (ld_volatile(prog_prev) < need) return false;
                __threadfence();
#pragma unroll
                for (int cc = 0; cc < C; ++cc) {
                    const int col = C * firstgrp + lane + 32 * cc;
                    if (col < need) { h[cc] = __ldcg(bH + 2 * (int64_t)col); f[cc] = __ldcg(bF + 2 * (int64_t)col); }
                }
                return true;
            };
            uint32_t rc_prev = 0, rc_cur = load_res(lane), rc_next = load_res(32 + lane);
            int bh_cur[C], bf_cur[C], bh_next[C], bf_next[C];
#pragma unroll
            for (int cc = 0; cc < C; ++cc) { bh_cur[cc] = bf_cur[cc] = bh_next[cc] = bf_next[cc] = 0; }
            // a band follows the one above as closely as the published columns allow: the chunk after the current
            // one is tried for at every half chunk and waited for only when the current one is used up
            bool have_next = true;
            if (pass > 0) {
                load_bnd(0, bh_cur, bf_cur);
                if (C * 32 < ncols) have_next = try_bnd(32, bh_next, bf_next);
            }
            int2* const wbnd = sbnd + (threadIdx.x & ~31) * C;  // this warp's 32 C boundary columns

            int thr = pvg;  // KIND_ALL: best G any lane of the warp holds (candidates below it cannot win)
            for (int s0 = 0; s0 < nsteps; s0 += 32) {
                if (s0 > 0) {
                    // ---- chunk boundary (every 32 steps): everything that touches global memory lives here ----
                    if (cutting && multi) {  // another band may have met `expect`
                        const int cc = ld_volatile(cutp);
                        if (cc < ncols) { ncols = cc; nsteps = min(nsteps, steps_for(ncols)); if (s0 >= nsteps) break; }
                    }
                    rc_prev = rc_cur; rc_cur = rc_next; rc_next = load_res(s0 + 32 + lane);
                    if (pass > 0) {
                        if (!have_next) load_bnd(s0, bh_next, bf_next);
#pragma unroll
                        for (int cc = 0; cc < C; ++cc) { bh_cur[cc] = bh_next[cc]; bf_cur[cc] = bf_next[cc]; }
                        have_next = (C * (s0 + 32) < ncols) ? try_bnd(s0 + 32, bh_next, bf_next) : true;
                    }
                    if (multi && !last && lane == 31 && s0 > 31) {  // lane 31 is done with the groups below s0 - 31
                        __threadfence();
                        *reinterpret_cast<volatile int32_t*>(prog_mine) = min(C * (s0 - 31), ncols);
                    }
                }
                if (pass > 0) {  // lane 0 reads the boundary row of the band above from shared memory
                    __syncwarp();
#pragma unroll
                    for (int cc = 0; cc < C; ++cc) wbnd[lane + 32 * cc] = make_int2(bh_cur[cc], bf_cur[cc]);
                    __syncwarp();
                }
                for (int hb = 0; hb < 2; ++hb) {
                    if (hb == 1 && pass > 0 && !have_next) have_next = try_bnd(s0 + 32, bh_next, bf_next);
                    int lim = min(s0 + 16 * (hb + 1), nsteps);
                    for (int s = s0 + 16 * hb; s < lim; ++s) {
                        const int sl = s - s0;
                        int upG[C], upF[C];
#pragma unroll
                        for (int cc = 0; cc < C; ++cc) {
                            upG[cc] = __shfl_up_sync(FULL, Gout[cc], 1);
                            upF[cc] = __shfl_up_sync(FULL, Fout[cc], 1);
                        }
                        const uint32_t tw = __shfl_sync(FULL, (lane <= sl) ? rc_cur : rc_prev, (sl - lane) & 31);
                        const int grp = s - lane;
                        const int c0 = C * grp;
                        const bool active = (grp >= 0) && (c0 < ncols) && (nv > 0);
                        int cm = 4 * NEGV;
                        int Gs[C][K];  // KIND_ALL: the columns of this step, for the rare path
                        if (active) {
                            // lane 0 takes the row above the band: the border (evaluated by every lane, so the warp
                            // does not diverge in the one-band case) or the last row of the band above
                            if (pass == 0) {
#pragma unroll
                                for (int cc = 0; cc < C; ++cc) {
                                    const int bg = border_val(P.top_free, c0 + cc, go, ge) - go;
                                    upG[cc] = (lane == 0) ? bg : upG[cc];
                                    upF[cc] = (lane == 0) ? bg : upF[cc];  // F of the first row: max(-inf, H above - open)
                                }
                            } else if (lane == 0) {
#pragma unroll
                                for (int cc = 0; cc < C; ++cc) {
                                    const int2 bb = wbnd[C * sl + cc];
                                    upG[cc] = bb.x;
                                    upF[cc] = bb.y;
                                }
                            }
#pragma unroll
                            for (int cc = 0; cc < C; ++cc) {
                                const int c = c0 + cc;
                                const bool colok = (C == 1) || (c < ncols);  // the last group may run past the target
                                const int tl = (C == 1) ? (int)tw : (int)((tw >> (8 * cc)) & 255u);
                                int gd = (cc == 0) ? Gd0 : upG[cc > 0 ? cc - 1 : 0];
                                int gu = upG[cc], f = upF[cc], glast = 0, cmc = 4 * NEGV;
                                const char* const srow = reinterpret_cast<const char*>(sS) + 4 * tl;
#pragma unroll
                                for (int r = 0; r < K; ++r) {
                                    const int sc = *reinterpret_cast<const int*>(srow + qa[r]);
                                    const int gl = Gl[r];
                                    const int e = max(El[r] - ge, gl);
                                    const int d = max(gd + sc, floorv);
                                    const int de = max(d, e);
                                    const int deg = de - go;
                                    const int g = max(f - go, deg);  // = max(d, e, F) - open
                                    if (KIND == KIND_ALL) {
                                        cmc = max(cmc, g);
                                        Gs[cc][r] = g;
                                    } else if (KIND == KIND_EDGE) {
                                        if (r == rl) glast = g;
                                    } else {
                                        const int h = max(de, f);
                                        const uint32_t src = (h == d) ? 0u : (h == e) ? 1u : 2u;
                                        const uint32_t b = src | ((e == gl) ? 4u : 0u) | ((f == gu) ? 8u : 0u);
                                        if ((r & 3) == 0) tbw[r >> 2] = b;
                                        else tbw[r >> 2] |= b << (8 * (r & 3));
                                        if (r == rl) glast = g;
                                    }
                                    f = max(f - gmin, deg);
                                    gd = gl;
                                    Gl[r] = g;
                                    El[r] = e;
                                    gu = g;
                                }
                                const int hlast = glast + go;
                                Gout[cc] = gu;
                                Fout[cc] = f;
                                if (lane == 31 && !last && colok) {
                                    __stcg(bH + 2 * (int64_t)c, gu);
                                    __stcg(bF + 2 * (int64_t)c, f);
                                }
                                if (KIND == KIND_ALL) {
                                    cm = max(cm, colok ? cmc : 4 * NEGV);
                                } else if (KIND == KIND_EDGE) {
                                    if (own && colok) {
                                        if (P.endrule == END_CORNER) {
                                            if (c == n - 1) { lv = hlast; lc = c; }
                                        } else if (hlast > lv) {
                                            lv = hlast; lc = c;
                                        }
                                    }
                                    if (P.endrule == END_LASTROW_LASTCOL && c == n - 1) {
#pragma unroll
                                        for (int r = 0; r < K; ++r)
                                            if (r < nv && Gl[r] + go > cv) { cv = Gl[r] + go; ci = i0 + r; }
                                    }
                                } else {
                                    if (colok) {
                                        uint8_t* colb = P.tb + jb.tb_off + (int64_t)c * ((m + 3) & ~3) + i0;
                                        uint32_t* col = reinterpret_cast<uint32_t*>(colb);
#pragma unroll
                                        for (int w = 0; w < K / 4; ++w)
                                            if (4 * w < nv) col[w] = tbw[w];
                                    }
                                    if (own && c == n - 1) corner = hlast;
                                }
                            }
                            Gd0 = upG[C - 1];
                        }
                        if (KIND == KIND_ALL) {
                            // rare (a warp-uniform branch): some lane's group beats both its own best and the warp's
                            const bool cand = (cm > pvg) && (cm >= thr);
                            if (__any_sync(FULL, cand)) {
                                if (cand) {
                                    int rf = -1, rc = 0;  // the first cell in (column, row) order that holds the maximum
#pragma unroll
                                    for (int cc = C - 1; cc >= 0; --cc)
#pragma unroll
                                        for (int r = K - 1; r >= 0; --r)
                                            if (r < nv && (C == 1 || c0 + cc < ncols) && Gs[cc][r] == cm) { rf = r; rc = cc; }
                                    if (rf >= 0) { pvg = cm; pc = c0 + rc; pi = i0 + rf; }
                                }
                                int w = pvg;
                                for (int d = 16; d > 0; d >>= 1) w = max(w, __shfl_xor_sync(FULL, w, d));
                                thr = w;
                                if (cutting) {
                                    // reverse pass: stop once every lane is past the first column that reached `expect`
                                    const bool hit = (pc >= 0) && (pvg + go == jb.expect);
                                    if (__any_sync(FULL, hit)) {
                                        int cmn = hit ? pc : INT_MAX;
                                        for (int d = 16; d > 0; d >>= 1) cmn = min(cmn, __shfl_xor_sync(FULL, cmn, d));
                                        if (cmn + 1 < ncols) {
                                            ncols = cmn + 1;
                                            if (multi && lane == 0) atomicMin(cutp, ncols);
                                        }
                                        nsteps = min(nsteps, steps_for(ncols));
                                        lim = min(lim, nsteps);
                                    }
                                }
                            }
                        }
                    }
                }
            }
            if (multi && !last) {
                // every column below the (possibly cut) width is in place
                __syncwarp();
                if (lane == 31) {
                    __threadfence();
                    *reinterpret_cast<volatile int32_t*>(prog_mine) = ncols;
                }
            }
            if (KIND == KIND_ALL && pc >= 0 && better(pvg + go, pc, pi, gv, gc, gi)) { gv = pvg + go; gc = pc; gi = pi; }
            __syncwarp();
        }

        if (KIND == KIND_FILL) {
            if (pass == npass - 1) {
                corner = __shfl_sync(FULL, corner, owner);
                if (lane == 0 && corner != jb.expect) atomicExch(P.err, 1 + jb.slot);
            }
            continue;
        }

        // this band's candidates: best cell (KIND_ALL), last-row best (owner lane of the last band), last-column best
        if (KIND == KIND_ALL) {
            for (int d = 16; d > 0; d >>= 1) {
                const int ov = __shfl_xor_sync(FULL, gv, d), oc = __shfl_xor_sync(FULL, gc, d),
                          oi = __shfl_xor_sync(FULL, gi, d);
                if (better(ov, oc, oi, gv, gc, gi)) { gv = ov; gc = oc; gi = oi; }
            }
        } else {
            lv = __shfl_sync(FULL, lv, owner);
            lc = __shfl_sync(FULL, lc, owner);
            for (int d = 16; d > 0; d >>= 1) {
                const int ov = __shfl_xor_sync(FULL, cv, d), oi = __shfl_xor_sync(FULL, ci, d);
                if (ov > cv || (ov == cv && oi < ci)) { cv = ov; ci = oi; }
            }
        }
        if (multi) {
            // hand the candidates to the band that finishes last
            int lastone = 0;
            if (lane == 0) {
                int32_t* sl5 = P.slots + 5 * (int64_t)task;
                if (KIND == KIND_ALL) { sl5[0] = gv; sl5[1] = gc; sl5[2] = gi; }
                else { sl5[0] = cv; sl5[1] = ci; sl5[3] = lv; sl5[4] = lc; }
                __threadfence();
                lastone = (atomicAdd(P.done + job, 1) == npass - 1);
                if (lastone) __threadfence();
            }
            lastone = __shfl_sync(FULL, lastone, 0);
            if (!lastone) continue;
            if (lane == 0) {
                const int32_t* base = P.slots + 5 * (int64_t)job * P.maxpass;
                if (KIND == KIND_ALL) {
                    gv = P.floor0 ? 0 : INT_MIN; gc = -1; gi = -1;
                    for (int p = 0; p < npass; ++p) {
                        const int ov = __ldcg(base + 5 * p), oc = __ldcg(base + 5 * p + 1), oi = __ldcg(base + 5 * p + 2);
                        if (better(ov, oc, oi, gv, gc, gi)) { gv = ov; gc = oc; gi = oi; }
                    }
                } else {
                    cv = INT_MIN; ci = -1;
                    for (int p = 0; p < npass; ++p) {  // bands hold ascending rows: strict > keeps the lowest row
                        const int ov = __ldcg(base + 5 * p), oi = __ldcg(base + 5 * p + 1);
                        if (ov > cv) { cv = ov; ci = oi; }
                    }
                    lv = __ldcg(base + 5 * (npass - 1) + 3);
                    lc = __ldcg(base + 5 * (npass - 1) + 4);
                }
            }
        }
        if (lane == 0) {
            int score, rq, rt;
            if (KIND == KIND_ALL) {
                score = gv; rq = gi; rt = gc;
            } else {
                score = lv; rq = m - 1; rt = lc;
                if (P.endrule == END_LASTROW_LASTCOL) {
                    if (P.anchored) {
                        const int tbv = border_val(0, n - 1, go, ge);
                        if (tbv > score) { score = tbv; rq = -1; rt = n - 1; }
                    }
                    if (cv > score) { score = cv; rq = ci; rt = n - 1; }
                }
            }
            P.out_score[jb.slot] = score;
            P.out_q[jb.slot] = rq;
            P.out_t[jb.slot] = rt;
        }
    }
}

// One thread per job: walk the direction matrix backwards from the corner
// (oracle/opal_oracle.c traceback()): at an H cell the diagonal wins over the
// horizontal gap (E, OPAL_ALIGN_INS) which wins over the vertical gap (F,
// OPAL_ALIGN_DEL); inside a gap, closing it wins over extending it.  The
// operations are written backwards from the end of the job's slot.
__global__ void walk_kernel(const PairParams P, uint8_t* ops, int32_t* ops_len) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= P.njobs) return;
    const PairJob jb = P.jobs[x];
    const int m = jb.m, n = jb.n;
    const int64_t mpad = (m + 3) & ~3;
    const uint8_t* tbytes = reinterpret_cast<const uint8_t*>(P.stream + jb.tbase) + jb.half;
    const uint8_t* tb = P.tb + jb.tb_off;
    uint8_t* w = ops + jb.ops_end;
    int i = m - 1, j = n - 1, state = 0, len = 0;
    while (i >= 0 && j >= 0) {
        const unsigned b = tb[(int64_t)j * mpad + i];
        if (state == 0) {
            const unsigned src = b & 3u;
            if (src == 0) {
                const bool same = P.query[jb.q0 + i] == tbytes[2 * (int64_t)(jb.t0 + j)];
                *--w = same ? OPAL_ALIGN_MATCH : OPAL_ALIGN_MISMATCH;
                ++len; --i; --j;
            } else {
                state = (int)src;
            }
        } else if (state == 1) {
            *--w = OPAL_ALIGN_INS;
            ++len;
            state = (b & 4u) ? 0 : 1;
            --j;
        } else {
            *--w = OPAL_ALIGN_DEL;
            ++len;
            state = (b & 8u) ? 0 : 2;
            --i;
        }
    }
    for (; j >= 0; --j) { *--w = OPAL_ALIGN_INS; ++len; }
    for (; i >= 0; --i) { *--w = OPAL_ALIGN_DEL; ++len; }
    ops_len[jb.slot] = len;
}

// forward jobs straight from the store's tables: job x = sorted position x
// (longest targets first), skipped (m = 0) when outside [start, end)
__global__ void make_forward_jobs(PairJob* jobs, int njobs, const int64_t* pair_off, const int32_t* pair_len,
                                  const int32_t* pair_idx, int Lq, int64_t start, int64_t end) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= njobs) return;
    PairJob jb;
    const int idx = pair_idx[x];
    jb.tbase = pair_off[x >> 1];
    jb.tb_off = 0;
    jb.half = x & 1;
    jb.q0 = 0; jb.qdir = 1; jb.m = Lq;
    jb.t0 = 0; jb.tdir = 1; jb.n = pair_len[x];
    jb.expect = 0;
    jb.slot = 0;
    jb.ops_end = 0;
    if (idx < start || idx >= end) jb.m = 0;
    else jb.slot = (int)(idx - start);
    jobs[x] = jb;
}

// reverse jobs from the forward results (database order, relative to start)
__global__ void make_reverse_jobs(PairJob* jobs, int njobs, const int32_t* score, const int32_t* eq, const int32_t* et) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= njobs) return;
    PairJob jb = jobs[x];
    if (jb.m > 0) {
        const int q = eq[jb.slot], t = et[jb.slot];
        if (q < 0 || t < 0) {
            jb.m = 0;
        } else {
            jb.q0 = q; jb.qdir = -1; jb.m = q + 1;
            jb.t0 = t; jb.tdir = -1; jb.n = t + 1;
            jb.expect = score[jb.slot];
        }
    }
    jobs[x] = jb;
}

__global__ void fill_i32(int32_t* p, int64_t n, int32_t v) {
    for (int64_t x = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; x < n; x += (int64_t)gridDim.x * blockDim.x) p[x] = v;
}

typedef void (*pair_kernel_t)(const PairParams);

// Rows per lane (K) and columns per step (C).  One pass (K = rows / 32, rounded up to 4) does the least work, and is
// right when there are targets enough to fill the machine.  With few targets the longest one is the critical path -
// a single warp walking it column by column - so the query is cut into bands of 32 K rows walked by several warps
// at once, and a lane takes C = 4 columns a step, which spreads the per-step latencies (shuffle round trip,
// branches, shared-memory loads) over four columns.  Both sides of that trade are modelled in clocks, with
// constants measured on B200 (scratch/pairk_probe.py): a lone warp takes about 470 + 13 K clocks per step at C = 1
// and 1090 + 42 K at C = 4, a band follows ~72 steps behind the one above it, a step issues about 8.6 K C + 60
// (+ 40 at C = 4) instructions and a full SM sustains ~2.2 of them per clock; the two terms add because the long
// target's warps share their SMs with everybody else's.
int choose_pair_K(const Tuning& tune, int rows, int64_t njobs, int64_t sum_n, int64_t nmax, int64_t table_jobs, int sm_count, int gran, int* C) {
    *C = 1;
    // `gran`: K comes in steps of 2 for the locate kernels, 4 for the fill kernel (four rows per 32-bit store)
    const int kfit = std::min(32, std::max(4, gran * ((rows + 32 * gran - 1) / (32 * gran))));
    if (tune.pair_k >= 4) {
        const int k = std::min(32, tune.pair_k / gran * gran);
        if (tune.pair_c == 4 && k <= 8) *C = 4;
        return k;
    }
    double best = 1e300;
    int bk = kfit;
    for (int c = 1; c <= 4; c += 3) {
        for (int k = 4; k <= (c == 1 ? kfit : std::min(kfit, 8)); k += (c == 1 ? gran : 4)) {
            const int64_t bands = (rows + 32 * k - 1) / (32 * k);
            if (k < kfit && table_jobs * bands > ((int64_t)1 << 25)) continue;  // band bookkeeping: 24 B per task
            if (c == 4 && bands > 24) continue;
            const double step = (c == 1) ? 470.0 + 13.0 * k : 1090.0 + 42.0 * k;
            const double instr = 8.6 * k * c + (c == 1 ? 60.0 : 100.0);
            const double crit = ((double)nmax / c + 72.0 * (double)(bands - 1)) * step;
            const double thr = (double)bands * ((double)sum_n / c + 31.0 * (double)njobs) * instr / (2.2 * sm_count);
            const double cost = crit + thr;
            if (cost < best * 0.97) { best = cost; bk = k; *C = c; }
        }
    }
    return bk;
}

template <int KIND>
pair_kernel_t pair_kernel_for(int k, int c) {
    if (c == 4 && k == 4) return pair_kernel<4, KIND, 4>;
    if (c == 4 && k == 8) return pair_kernel<8, KIND, 4>;
    if constexpr (KIND != KIND_FILL) {
        switch (k) {
            case 6: return pair_kernel<6, KIND, 1>;
            case 10: return pair_kernel<10, KIND, 1>;
            case 14: return pair_kernel<14, KIND, 1>;
            case 18: return pair_kernel<18, KIND, 1>;
            case 22: return pair_kernel<22, KIND, 1>;
            case 26: return pair_kernel<26, KIND, 1>;
            case 30: return pair_kernel<30, KIND, 1>;
            default: break;
        }
    }
    switch (k) {
        case 4: return pair_kernel<4, KIND, 1>;
        case 8: return pair_kernel<8, KIND, 1>;
        case 12: return pair_kernel<12, KIND, 1>;
        case 16: return pair_kernel<16, KIND, 1>;
        case 20: return pair_kernel<20, KIND, 1>;
        case 24: return pair_kernel<24, KIND, 1>;
        case 28: return pair_kernel<28, KIND, 1>;
        default: return pair_kernel<32, KIND, 1>;
    }
}

// Reset the band bookkeeping for `P.njobs` jobs of at most `rows` query rows and launch.
int launch_pair(pyopal_cuda_db* db, Workspace* ws, pair_kernel_t fn, PairParams P, int rows, int K, cudaStream_t st) {
    if (P.njobs <= 0) return 0;
    P.maxpass = std::max(1, (rows + 32 * K - 1) / (32 * K));
    const int64_t ntasks = (int64_t)P.njobs * P.maxpass;
    if (ntasks >= ((int64_t)1 << 31) - 1024) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "too many target x band tasks");
    P.prog = P.done = P.cut = P.slots = nullptr;
    if (P.maxpass > 1) {
        if (ws->pd_team.reserve((size_t)(6 * ntasks + 2 * (int64_t)P.njobs + 16))) return PYOPAL_CUDA_ERR_NOMEM;
        P.prog = ws->pd_team.p;
        P.done = P.prog + ntasks;
        P.cut = P.done + P.njobs;
        P.slots = P.cut + P.njobs;
        CU(cudaMemsetAsync(P.prog, 0, (size_t)(ntasks + P.njobs) * sizeof(int32_t), st));
        fill_i32<<<256, 256, 0, st>>>(P.cut, P.njobs, INT_MAX);
    }
    CU(cudaMemsetAsync(P.counter, 0, sizeof(int32_t), st));
    int per_sm = 0;
    CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, 0));
    if (per_sm < 1) return fail(PYOPAL_CUDA_ERR_INTERNAL, "pair kernel does not fit on an SM");
    // every block is resident (a band may wait for the band above it, which was claimed earlier)
    int64_t grid = std::min<int64_t>((int64_t)db->sm_count * per_sm, (ntasks + 3) / 4);
    fn<<<(unsigned)std::max<int64_t>(grid, 1), 128, 0, st>>>(P);
    CU(cudaGetLastError());
    ws->stats[0]++;
    return 0;
}

struct ModeCfg {
    int top_free, left_free, floor0, endrule;
};
ModeCfg cfg_for_mode(int mode) {
    switch (mode) {
        case MODE_NW: return {0, 0, 0, END_CORNER};
        case MODE_HW: return {1, 0, 0, END_LASTROW};
        case MODE_OV: return {1, 1, 0, END_LASTROW_LASTCOL};
        default: return {1, 1, 1, END_ALL};
    }
}

// closed forms for an empty query / empty target (oracle/opal_oracle.c dp_end, m == 0 || n == 0)
void degenerate_end(int mode, int m, int n, int go, int ge, int* score, int* eq, int* et) {
    *eq = -1; *et = -1; *score = 0;
    if (mode == MODE_NW) {
        if (!(m == 0 && n == 0)) *score = -(go + (std::max(m, n) - 1) * ge);
        *eq = m - 1; *et = n - 1;
    } else if (mode == MODE_HW) {
        *score = (m == 0) ? 0 : -(go + (m - 1) * ge);
        *eq = m - 1; *et = -1;
    }
}

}  // namespace

// host-built forward jobs for the sorted positions in `pos` (targets the packed kernel leaves to this file)
static void host_forward_jobs(const pyopal_cuda_db* db, const int32_t* pos, size_t count, bool packed_entries, int Lq,
                              int64_t start, std::vector<PairJob>& out) {
    out.clear();
    out.reserve(count);
    for (size_t i = 0; i < count; i++) {
        const int32_t x = pos[i];  // 2 * pair + half either way
        (void)packed_entries;
        PairJob jb;
        jb.tbase = db->h_pair_off[(size_t)(x >> 1)];
        jb.tb_off = 0;
        jb.half = x & 1;
        jb.q0 = 0; jb.qdir = 1; jb.m = Lq;
        jb.t0 = 0; jb.tdir = 1; jb.n = db->h_pair_len[(size_t)x];
        jb.expect = 0;
        jb.slot = (int32_t)(db->h_pair_idx[(size_t)x] - start);
        jb.ops_end = 0;
        out.push_back(jb);
    }
}

int pyopal_end_full_search(pyopal_cuda_db* db, Workspace* ws, const unsigned char* query, int Lq, int go, int ge,
                           const int* M, int A, int searchType, int mode, int64_t start, int64_t end, int32_t* scores,
                           int32_t* endQ, int32_t* endT, int32_t* startQ, int32_t* startT, unsigned char* ops,
                           const int64_t* ops_offsets, int32_t* ops_len) {
    const int64_t nout = end - start;
    if (nout <= 0) return 0;
    const bool full = (searchType == OPAL_SEARCH_ALIGNMENT);
    if (!endQ || !endT) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null end-location output");
    if (full && (!startQ || !startT || !ops || !ops_offsets || !ops_len))
        return fail(PYOPAL_CUDA_ERR_ARGUMENT, "null alignment output");
    for (int i = 0; i < 12; i++) ws->stats[i] = 0;
    cudaStream_t st = ws->stream;
    if (ws->used) CU(cudaStreamWaitEvent(st, ws->ev_free, 0));
    ws->used = true;
    struct Mark {
        Workspace* w; cudaStream_t s;
        ~Mark() { cudaEventRecord(w->ev_free, s); }
    } mark{ws, st};
    CU(cudaEventRecord(ws->ev0, st));

    // 32-bit overflow cannot happen below this bound (values stay exact otherwise)
    {
        int smax = INT_MIN, smin = INT_MAX;
        for (int i = 0; i < A * A; i++) { smax = std::max(smax, M[i]); smin = std::min(smin, M[i]); }
        const int64_t mag = std::max<int64_t>({(int64_t)go, (int64_t)ge, 1, std::llabs((long long)smax), std::llabs((long long)smin)});
        const int64_t longest = db->npairs ? db->h_pair_len[0] : 0;
        if (mag * ((int64_t)Lq + longest + 40) >= (1LL << 29))
            return fail(OPAL_ERR_OVERFLOW, "scores may exceed the 32-bit range");
    }

    std::vector<int32_t> h_sc((size_t)nout), h_eq((size_t)nout, -1), h_et((size_t)nout, -1);
    std::vector<int32_t> h_sq((size_t)nout, -1), h_st((size_t)nout, -1), h_rs((size_t)nout, 0);
    const int njobs = (int)(2 * db->npairs);
    const bool device_work = (Lq > 0 && njobs > 0);
    int32_t* d_res = nullptr;  // [score | eq | et | rscore | rq | rt] x nout, then counter, err
    PairJob* d_jobs = nullptr;
    PairParams P;
    memset(&P, 0, sizeof(P));

    if (device_work) {
        if (ws->pd_i32.reserve((size_t)nout * 6 + 64)) return PYOPAL_CUDA_ERR_NOMEM;
        if (ws->pd_query.reserve((size_t)Lq + 32 * 32 * 4 + 16)) return PYOPAL_CUDA_ERR_NOMEM;
        d_res = ws->pd_i32.p;
        int32_t* d_matrix = reinterpret_cast<int32_t*>(ws->pd_query.p);
        uint8_t* d_query = ws->pd_query.p + 32 * 32 * 4;
        CU(cudaMemcpyAsync(d_query, query, (size_t)Lq, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(d_matrix, M, (size_t)A * A * sizeof(int32_t), cudaMemcpyHostToDevice, st));
        int32_t* d_ctl = d_res + (size_t)nout * 6;
        CU(cudaMemsetAsync(d_ctl, 0, 32 * sizeof(int32_t), st));
        fill_i32<<<256, 256, 0, st>>>(d_res, nout * 6, -1);
        CU(cudaGetLastError());
        P.stream = db->d_stream; P.query = d_query; P.matrix = d_matrix;
        P.A = A; P.go = go; P.ge = ge;
        P.err = d_ctl + 1;
        const ModeCfg c = cfg_for_mode(mode);
        P.top_free = c.top_free; P.left_free = c.left_free; P.floor0 = c.floor0; P.endrule = c.endrule; P.anchored = 0;
        P.out_score = d_res; P.out_q = d_res + nout; P.out_t = d_res + 2 * nout;
        const int tpb = 256;

        // locate launch over a host-built job list (the 32-bit kernels of this file)
        std::vector<PairJob> hjobs0;
        auto locate = [&](const std::vector<PairJob>& jobs, size_t job_off_bytes, int32_t* counter, cudaStream_t s2) -> int {
            if (jobs.empty()) return 0;
            int64_t sum_n = 0, nmax = 0;
            for (const PairJob& jb : jobs) { sum_n += jb.n; nmax = std::max<int64_t>(nmax, jb.n); }
            int Cc = 1;
            const int K = choose_pair_K(db->tune, Lq, (int64_t)jobs.size(), sum_n, nmax, (int64_t)jobs.size(), db->sm_count, 2, &Cc);
            pair_kernel_t kloc = (c.endrule == END_ALL) ? pair_kernel_for<KIND_ALL>(K, Cc) : pair_kernel_for<KIND_EDGE>(K, Cc);
            ws->stats[5] = Cc; ws->stats[6] = K; ws->stats[7] = (Lq + 32 * K - 1) / (32 * K);
            if (Lq > 32 * K)
                if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                    return PYOPAL_CUDA_ERR_NOMEM;
            PairJob* dj = reinterpret_cast<PairJob*>(ws->pd_jobs.p + job_off_bytes);
            CU(cudaMemcpyAsync(dj, jobs.data(), jobs.size() * sizeof(PairJob), cudaMemcpyHostToDevice, s2));
            PairParams Q = P;
            Q.jobs = dj; Q.njobs = (int)jobs.size(); Q.counter = counter;
            Q.bndH = reinterpret_cast<int32_t*>(ws->bnd32H.p); Q.bndF = reinterpret_cast<int32_t*>(ws->bnd32F.p);
            return launch_pair(db, ws, kloc, Q, Lq, K, s2);
        };

        bool packed_done = false;
        if (!db->tune.no_end16) {
            // ---- forward pass in the packed 16-bit kernels (kernels.cuh, END): score + end cell at score-mode
            // cost; what they cannot hold (the longest targets, the static 32-bit rung, 16-bit overflows) is
            // located by this file's 32-bit kernel, concurrently on a high-priority stream
            EndPlan ep;
            ep.d_endq = d_res + nout; ep.d_endt = d_res + 2 * nout;
            // reserve the job area first: [all-target jobs (full mode) | spill jobs | overflow jobs]
            if (ws->pd_jobs.reserve((size_t)njobs * sizeof(PairJob) * 2 + 4096)) return PYOPAL_CUDA_ERR_NOMEM;
            CU(cudaEventRecord(ws->ev_aux0, st));
            bool spilled = false;
            ep.launch_spill = [&]() -> int {
                host_forward_jobs(db, ep.spill.data(), ep.spill.size(), false, Lq, start, hjobs0);
                CU(cudaStreamWaitEvent(ws->side2, ws->ev_aux0, 0));
                int r = locate(hjobs0, (size_t)njobs * sizeof(PairJob), d_ctl, ws->side2);
                if (r) return r;
                CU(cudaEventRecord(ws->ev_aux1, ws->side2));
                spilled = true;
                return 0;
            };
            const int64_t launches_before = ws->stats[0];
            int rc = enqueue_score_search_ex(db, ws, query, Lq, go, ge, M, A, mode, start, end, d_res, st, &ep);
            if (rc) return rc;
            (void)launches_before;
            packed_done = true;
            if (spilled) CU(cudaStreamWaitEvent(st, ws->ev_aux1, 0));
            if (ep.packed && mode == MODE_SW) {
                // 16-bit overflows: the list was produced on the device; fetch it and locate those at 32 bit
                int32_t novf = 0;
                CU(cudaMemcpyAsync(&novf, ep.d_ovf_count, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
                CU(cudaStreamSynchronize(st));
                ws->stats[3] = novf;
                if (novf > 0) {
                    std::vector<int32_t> ovf((size_t)novf);
                    CU(cudaMemcpy(ovf.data(), ep.d_ovf, (size_t)novf * sizeof(int32_t), cudaMemcpyDeviceToHost));
                    std::vector<PairJob> hj;
                    host_forward_jobs(db, ovf.data(), ovf.size(), true, Lq, start, hj);
                    rc = locate(hj, (size_t)njobs * sizeof(PairJob) + hjobs0.size() * sizeof(PairJob), d_ctl + 8, st);
                    if (rc) return rc;
                }
            }
            d_jobs = reinterpret_cast<PairJob*>(ws->pd_jobs.p);
            if (full && mode != MODE_NW) {
                make_forward_jobs<<<(njobs + tpb - 1) / tpb, tpb, 0, st>>>(d_jobs, njobs, db->d_pair_off, db->d_pair_len,
                                                                          db->d_pair_idx, Lq, start, end);
                CU(cudaGetLastError());
            }
        }
        if (!packed_done) {
            if (ws->pd_jobs.reserve((size_t)njobs * sizeof(PairJob))) return PYOPAL_CUDA_ERR_NOMEM;
            d_jobs = reinterpret_cast<PairJob*>(ws->pd_jobs.p);
            int64_t sum_n = 0, nmax = 0, nlive = 0;
            for (int64_t k = 0; k < nout; k++) {
                const int64_t len = db->h_lengths[(size_t)(start + k)];
                sum_n += len; nmax = std::max(nmax, len); nlive += (len > 0);
            }
            int Cc = 1;
            const int K = choose_pair_K(db->tune, Lq, nlive, sum_n, nmax, njobs, db->sm_count, 2, &Cc);
            pair_kernel_t kloc = (c.endrule == END_ALL) ? pair_kernel_for<KIND_ALL>(K, Cc) : pair_kernel_for<KIND_EDGE>(K, Cc);
            ws->stats[5] = Cc; ws->stats[6] = K; ws->stats[7] = (Lq + 32 * K - 1) / (32 * K);
            if (Lq > 32 * K)
                if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                    return PYOPAL_CUDA_ERR_NOMEM;
            make_forward_jobs<<<(njobs + tpb - 1) / tpb, tpb, 0, st>>>(d_jobs, njobs, db->d_pair_off, db->d_pair_len,
                                                                      db->d_pair_idx, Lq, start, end);
            CU(cudaGetLastError());
            PairParams Q = P;
            Q.jobs = d_jobs; Q.njobs = njobs; Q.counter = d_ctl;
            Q.bndH = reinterpret_cast<int32_t*>(ws->bnd32H.p); Q.bndF = reinterpret_cast<int32_t*>(ws->bnd32F.p);
            int rc = launch_pair(db, ws, kloc, Q, Lq, K, st);
            if (rc) return rc;
        }
        if (full && mode != MODE_NW) {
            // reverse pass over every target, anchored at its end cell (the band-team kernel of this file)
            int64_t sum_n = 0, nmax = 0, nlive = 0;
            for (int64_t k = 0; k < nout; k++) {
                const int64_t len = db->h_lengths[(size_t)(start + k)];
                sum_n += len; nmax = std::max(nmax, len); nlive += (len > 0);
            }
            int Cc = 1;
            const int K = choose_pair_K(db->tune, Lq, nlive, sum_n, nmax, njobs, db->sm_count, 2, &Cc);
            pair_kernel_t kloc = (c.endrule == END_ALL) ? pair_kernel_for<KIND_ALL>(K, Cc) : pair_kernel_for<KIND_EDGE>(K, Cc);
            ws->stats[5] = Cc; ws->stats[6] = K; ws->stats[7] = (Lq + 32 * K - 1) / (32 * K);
            if (Lq > 32 * K)
                if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                    return PYOPAL_CUDA_ERR_NOMEM;
            make_reverse_jobs<<<(njobs + tpb - 1) / tpb, tpb, 0, st>>>(d_jobs, njobs, d_res, d_res + nout, d_res + 2 * nout);
            CU(cudaGetLastError());
            PairParams R = P;
            R.jobs = d_jobs; R.njobs = njobs; R.counter = d_ctl + 16;
            R.bndH = reinterpret_cast<int32_t*>(ws->bnd32H.p); R.bndF = reinterpret_cast<int32_t*>(ws->bnd32F.p);
            R.top_free = 0; R.left_free = 0; R.floor0 = 0; R.anchored = 1;
            R.out_score = d_res + 3 * nout; R.out_q = d_res + 4 * nout; R.out_t = d_res + 5 * nout;
            int rc = launch_pair(db, ws, kloc, R, Lq, K, st);
            if (rc) return rc;
        }
        CU(cudaMemcpyAsync(h_sc.data(), d_res, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h_eq.data(), d_res + nout, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(h_et.data(), d_res + 2 * nout, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
        if (full && mode != MODE_NW) {
            CU(cudaMemcpyAsync(h_rs.data(), d_res + 3 * nout, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h_sq.data(), d_res + 4 * nout, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
            CU(cudaMemcpyAsync(h_st.data(), d_res + 5 * nout, (size_t)nout * 4, cudaMemcpyDeviceToHost, st));
        }
        CU(cudaStreamSynchronize(st));
    }
    // degenerate targets (empty query or empty target): closed forms
    for (int64_t k = 0; k < nout; k++) {
        const int n = db->h_lengths[(size_t)(start + k)];
        if (Lq == 0 || n == 0) {
            int s, q, t;
            degenerate_end(mode, Lq, n, go, ge, &s, &q, &t);
            h_sc[(size_t)k] = s; h_eq[(size_t)k] = q; h_et[(size_t)k] = t;
        }
    }
    for (int64_t k = 0; k < nout; k++) {
        scores[k] = h_sc[(size_t)k];
        endQ[k] = h_eq[(size_t)k];
        endT[k] = h_et[(size_t)k];
    }
    if (!full) {
        CU(cudaEventRecord(ws->ev1, st));
        CU(cudaEventSynchronize(ws->ev1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, ws->ev0, ws->ev1));
        ws->stats[4] = (int64_t)(ms * 1000.0f);
        return 0;
    }

    // ---- start cells -----------------------------------------------------------
    struct Todo { int64_t k; int sq, stt, m, n; };
    std::vector<Todo> todo;
    for (int64_t k = 0; k < nout; k++) {
        const int eq = h_eq[(size_t)k], et = h_et[(size_t)k];
        ops_len[k] = 0;
        startQ[k] = -1; startT[k] = -1;
        if (eq < 0 || et < 0) continue;  // SW score 0 / degenerate: no alignment
        int sq = 0, stt = 0;
        if (mode != MODE_NW) {
            if (h_rs[(size_t)k] != h_sc[(size_t)k])
                return fail(PYOPAL_CUDA_ERR_INTERNAL, "reverse pass does not reproduce the forward score");
            sq = eq - h_sq[(size_t)k];
            stt = et - h_st[(size_t)k];
        }
        startQ[k] = sq; startT[k] = stt;
        if (sq > eq || stt > et) {
            // the alignment is a single gap (oracle/opal_oracle.c search_one)
            int len = (sq > eq) ? (et - stt + 1) : (eq - sq + 1);
            if (sq > eq && stt > et) len = 0;
            memset(ops + ops_offsets[k], (sq > eq) ? OPAL_ALIGN_INS : OPAL_ALIGN_DEL, (size_t)std::max(len, 0));
            ops_len[k] = std::max(len, 0);
            continue;
        }
        todo.push_back({k, sq, stt, eq - sq + 1, et - stt + 1});
    }

    // ---- direction matrices + walk, in batches bounded by device memory ----------
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    const int64_t budget = std::max<int64_t>((int64_t)1 << 28, std::min<int64_t>((int64_t)(free_b / 2), (int64_t)8 << 30));
    const std::vector<int32_t>& where = db->h_where;  // database index -> sorted position
    // big rectangles first so that the dynamic scheduler starts them early
    std::stable_sort(todo.begin(), todo.end(), [](const Todo& a, const Todo& b) {
        return (int64_t)a.m * a.n > (int64_t)b.m * b.n;
    });
    size_t pos = 0;
    std::vector<PairJob> hjobs;
    std::vector<int32_t> hlen;
    std::vector<unsigned char> hops;
    while (pos < todo.size()) {
        hjobs.clear();
        int64_t tb_bytes = 0, ops_bytes = 0, sum_tn = 0, max_tn = 0;
        int maxm = 0;
        size_t p2 = pos;
        while (p2 < todo.size()) {
            const Todo& t = todo[p2];
            const int64_t need = (int64_t)((t.m + 3) & ~3) * t.n;
            if (!hjobs.empty() && (tb_bytes + need > budget || hjobs.size() >= (size_t)1 << 22)) break;
            if (need > (int64_t)free_b - ((int64_t)1 << 28))
                return fail(PYOPAL_CUDA_ERR_NOMEM, "alignment matrix of one target does not fit in device memory");
            const int32_t x = where[(size_t)(start + t.k)];
            PairJob jb;
            jb.tbase = db->h_pair_off[(size_t)(x >> 1)];
            jb.tb_off = tb_bytes;
            jb.half = x & 1;
            jb.q0 = t.sq; jb.qdir = 1; jb.m = t.m;
            jb.t0 = t.stt; jb.tdir = 1; jb.n = t.n;
            jb.expect = h_sc[(size_t)t.k];
            jb.slot = (int32_t)hjobs.size();
            ops_bytes += t.m + t.n;
            if (ops_bytes > INT32_MAX - 16) { if (hjobs.empty()) return fail(PYOPAL_CUDA_ERR_ARGUMENT, "alignment too long"); break; }
            jb.ops_end = (int32_t)ops_bytes;
            hjobs.push_back(jb);
            tb_bytes += need;
            maxm = std::max(maxm, t.m);
            sum_tn += t.n;
            max_tn = std::max<int64_t>(max_tn, t.n);
            p2++;
        }
        const int nb = (int)hjobs.size();
        if (ws->pd_jobs.reserve((size_t)nb * sizeof(PairJob)) || ws->pd_tb.reserve((size_t)tb_bytes + 16) ||
            ws->pd_ops.reserve((size_t)ops_bytes + 16) || ws->pd_i32.reserve((size_t)nb + 16))
            return PYOPAL_CUDA_ERR_NOMEM;
        d_jobs = reinterpret_cast<PairJob*>(ws->pd_jobs.p);
        int32_t* d_len = ws->pd_i32.p;
        int32_t* d_ctl = d_len + nb;
        int Cc = 1;
        const int K = choose_pair_K(db->tune, maxm, nb, sum_tn, max_tn, nb, db->sm_count, 4, &Cc);
        pair_kernel_t kfill = pair_kernel_for<KIND_FILL>(K, Cc);
        if (maxm > 32 * K)
            if (ws->bnd32H.reserve((size_t)db->ncols * 2 + 16) || ws->bnd32F.reserve((size_t)db->ncols * 2 + 16))
                return PYOPAL_CUDA_ERR_NOMEM;
        CU(cudaMemcpyAsync(d_jobs, hjobs.data(), (size_t)nb * sizeof(PairJob), cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(d_ctl, 0, 16 * sizeof(int32_t), st));
        PairParams F = P;
        F.stream = db->d_stream; F.query = P.query; F.matrix = P.matrix;
        F.A = A; F.go = go; F.ge = ge;
        F.jobs = d_jobs; F.njobs = nb; F.counter = d_ctl; F.err = d_ctl + 1;
        F.bndH = reinterpret_cast<int32_t*>(ws->bnd32H.p); F.bndF = reinterpret_cast<int32_t*>(ws->bnd32F.p);
        F.top_free = 0; F.left_free = 0; F.floor0 = 0; F.endrule = END_CORNER; F.anchored = 0;
        F.tb = ws->pd_tb.p;
        int rc = launch_pair(db, ws, kfill, F, maxm, K, st);
        if (rc) return rc;
        walk_kernel<<<(nb + 63) / 64, 64, 0, st>>>(F, ws->pd_ops.p, d_len);
        CU(cudaGetLastError());
        ws->stats[0]++;
        hlen.resize((size_t)nb + 2);
        hops.resize((size_t)ops_bytes + 1);
        CU(cudaMemcpyAsync(hlen.data(), d_len, (size_t)nb * 4, cudaMemcpyDeviceToHost, st));
        int32_t herr[2] = {0, 0};
        CU(cudaMemcpyAsync(herr, d_ctl, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(hops.data(), ws->pd_ops.p, (size_t)ops_bytes, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (herr[1] != 0) return fail(PYOPAL_CUDA_ERR_INTERNAL, "traceback does not reproduce the forward score");
        for (int b = 0; b < nb; b++) {
            const Todo& t = todo[pos + (size_t)b];
            const int len = hlen[(size_t)b];
            memcpy(ops + ops_offsets[t.k], hops.data() + hjobs[(size_t)b].ops_end - len, (size_t)len);
            ops_len[t.k] = len;
        }
        pos = p2;
    }
    CU(cudaEventRecord(ws->ev1, st));
    CU(cudaEventSynchronize(ws->ev1));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, ws->ev0, ws->ev1));
    ws->stats[4] = (int64_t)(ms * 1000.0f);
    return 0;
}
