// pyopal_b200/csrc/kernels.cuh -- sm_100a database-search kernels.
//
// Replaces the SIMD inner loop of opalSearchDatabase (declared at
// /root/reference/src/pyopal/opal.pxd:40-54; the source lives in the absent
// vendor/opal submodule).  Not a port: the CPU library runs one target per
// SIMD lane down whole query columns; here
//
//   * the QUERY is register-tiled: a group of G threads (G = 4..32 lanes of
//     one warp) holds G*K consecutive query rows, K rows per thread, H-open
//     and E of the previous column in registers;
//   * the TARGET streams through the group one column per step, thread p
//     working on column (step - p): an anti-diagonal wavefront whose only
//     communication is two warp shuffles per step (H and F of the row above);
//   * in the 16-bit rung every 32-bit lane carries TWO targets (u16x2, values
//     stored with a bias B so that they are non-negative): maxima are DPX
//     instructions on the ALU pipe (VIADDMNMX.U16x2, VIMNMX3.U16x2,
//     VIMNMX.U16x2) and -- because no half can go negative or overflow --
//     the two additions of a cell (Hdiag + S, H - open) are plain 32-bit
//     IMADs on the otherwise idle FMA pipe; the database is stored
//     pair-interleaved so one u16 load yields the residues of both targets;
//   * a column step is two sweeps over the thread's K rows: bottom-up
//     (E and max(Hdiag + S, E), independent rows, results overwrite the
//     previous column in place) then top-down (the F chain), so no register
//     rotation is needed between steps;
//   * substitution scores come from a query profile in shared memory
//     (int8, score+gap_open folded), fetched K rows at a time with vector LDS
//     and widened with one PRMT per cell pair;
//   * queries longer than G*K rows take several passes; the last row of a
//     pass (H, F per column) round-trips through an L2-resident buffer;
//   * the 32-bit rung is the same code with one target per lane; 16-bit
//     overflow is detected from the running maximum (scores stay exact, so the
//     ladder is result-invisible, lib.pyx:1283-1289).
//
// Padding rows sit ABOVE query row 0 and are arithmetically transparent (see
// build_profile), so the last query row is always the last register of the
// last thread of the last pass: NW/HW/OV read their result from it without
// any per-cell bookkeeping.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace pyopal {

enum : int { MODE_NW = 0, MODE_HW = 1, MODE_OV = 2, MODE_SW = 3 };

struct KParams {
    // pair-interleaved store
    const uint16_t* stream;   // column c of pair k at stream[pair_off[k] + c]: lo byte = target a, hi = target b
    const int64_t* pair_off;  // [npairs]
    const int32_t* pair_len;  // [2*npairs] len_a >= len_b (len_b = 0 if absent)
    const int32_t* pair_idx;  // [2*npairs] database index of a, b (-1 if absent)
    // work
    const int32_t* worklist;  // W=16: pair ids; W=32: (pair << 1) | half
    const int32_t* nwork;     // device-resident count (lets a kernel produce the next kernel's list)
    int32_t* counter;         // task counter, zeroed before launch
    // query
    const uint8_t* query;
    const int32_t* matrix;    // A x A, row = query letter, column = target letter (pyx.in:86)
    const int8_t* gprofile;   // global-memory profile when it does not fit shared memory
    const int32_t* wprofile;  // int32 global profile for matrices / gap costs outside the int8 range
    int Lq, A, go, ge;
    int G, logG, npass, pitch, padtop;
    int krows, kstride;       // profile layout: a lane's K rows occupy kstride >= K bytes (odd number of words)
    int mode;
    int neg;                  // "-infinity" for this rung (true value; stored as neg + bias)
    int bias;                 // W=16: values are stored as value + bias in each u16 half; W=32: 0
    unsigned one;             // 1, as a run-time value: forces IMAD (FMA pipe) for the 32-bit adds
    int pad_score;            // folded score of padding rows / the padding letter
    int ovf_thresh;           // 16-bit rung, SW: running max >= this => rescore at 32 bit
    // results
    int32_t* out;             // scores in database order, entry (index - out_start)
    int32_t* out_q;           // END kernels: end cell (query row, target column), same indexing
    int32_t* out_t;
    int64_t out_start, out_end;
    int32_t* ovf_list;        // (pair << 1) | half of overflowed targets
    int32_t* ovf_count;
    // pass boundary (row above the tile): W=16 indexed like `stream`, W=32 at 2*column + half
    uint32_t* bndH;
    uint32_t* bndF;
    int first_pair;           // stream_kernel (stream.cuh): task t is pair first_pair + t
    // dual_kernel (dual.cuh): a second query in the high halves, its padding, results and overflow list
    const uint8_t* query2;
    int Lq2, padtop2;
    int32_t* out2;
    int32_t* ovf_list2;
    int32_t* ovf_count2;
};

// ---------------------------------------------------------------------------
// value traits: one s16x2 word (two targets) or one s32 (one target) per lane
template <int W> struct Val;

// d = a + b on the FMA pipe (a * one + b); exact for biased u16x2 words as long
// as neither half leaves [0, 65535] and the low half of the addend is >= 0, or
// the addend is a borrow-adjusted constant (see delta())
__device__ __forceinline__ uint32_t imad_add(uint32_t a, uint32_t b, unsigned one) {
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
}

template <> struct Val<16> {
    typedef uint32_t T;
    // value v (already biased, in [0, 65535]) in both halves
    static __device__ __forceinline__ T splat(int v) { return (uint32_t)(uint16_t)v * 0x00010001u; }
    // a signed delta for the packed adder of VIADDMNMX (each half wraps mod 2^16)
    static __device__ __forceinline__ T pdelta(int v) { return (uint32_t)(uint16_t)(int16_t)v * 0x00010001u; }
    // the same delta for a 32-bit add: v + v * 65536 in two's complement (borrow-adjusted)
    static __device__ __forceinline__ T delta(int v) { return (uint32_t)(v * 0x00010001); }
    static __device__ __forceinline__ T addmax(T a, T b, T c) { return __viaddmax_u16x2(a, b, c); }
    static __device__ __forceinline__ T add(T a, T b, unsigned one) { return imad_add(a, b, one); }
    static __device__ __forceinline__ T max2(T a, T b) { return __vmaxu2(a, b); }
    static __device__ __forceinline__ T max3(T a, T b, T c) { return __vimax3_u16x2(a, b, c); }
    static __device__ __forceinline__ int lo(T v) { return (int)(v & 0xffffu); }
    static __device__ __forceinline__ int hi(T v) { return (int)(v >> 16); }
};

template <> struct Val<32> {
    typedef int T;
    static __device__ __forceinline__ T splat(int v) { return v; }
    static __device__ __forceinline__ T pdelta(int v) { return v; }
    static __device__ __forceinline__ T delta(int v) { return v; }
    static __device__ __forceinline__ T addmax(T a, T b, T c) { return __viaddmax_s32(a, b, c); }
    static __device__ __forceinline__ T add(T a, T b, unsigned one) { return (int)imad_add((uint32_t)a, (uint32_t)b, one); }
    static __device__ __forceinline__ T max2(T a, T b) { return max(a, b); }
    static __device__ __forceinline__ T max3(T a, T b, T c) { return __vimax3_s32(a, b, c); }
    static __device__ __forceinline__ int lo(T v) { return v; }
    static __device__ __forceinline__ int hi(T v) { return v; }
};

// Profile: prof[letter * pitch + slot], letter in [0, A] (A = padding letter).
// The tile rows are stored lane chunk by lane chunk: chunk q = tile rows
// [q * krows, (q + 1) * krows) starts at slot q * kstride, where kstride is
// krows or krows + 4 so that a chunk is an ODD number of 32-bit words -- the
// lanes of a group then read their chunks of one letter from distinct banks
// (with kstride = K = 32 the 32 lanes of a group hit only 4 banks: measured
// 20 % slower than K = 28 or 36).  Entry = S[q][letter] + gap_open so the
// kernel can keep H-open instead of H (one subtract per cell feeds both E and
// F).  Tile rows < padtop are padding ABOVE the query: with a free top border
// (HW/OV/SW) score 0 keeps them at H = 0; with NW's penalised border a score
// of -max(open, ext) makes each of them reproduce the border row exactly.
__device__ __forceinline__ int profile_entry(const KParams& P, int letter, int slot) {
    const int chunk = slot / P.kstride, within = slot - chunk * P.kstride;
    if (within >= P.krows) return 0;  // bank padding, never read
    const int trow = chunk * P.krows + within;
    int qi = trow - P.padtop;
    if (qi < 0 || letter >= P.A) return P.pad_score;
    if (qi >= P.Lq) return 0;  // slots past the last chunk (pitch rounding): never read, and the query ends here
    return P.matrix[(int)P.query[qi] * P.A + letter] + P.go;
}

// PRMT selectors (generic mode: nibble bit 3 replicates the selected byte's
// sign).  __byte_perm masks that bit away, hence the inline PTX.
template <unsigned SEL>
__device__ __forceinline__ uint32_t prmt(uint32_t x, uint32_t y) {
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(x), "r"(y), "n"(SEL));
    return d;
}
// (sext16(byte u of x), sext16(byte u of y))
__device__ __forceinline__ uint32_t widen_pair(uint32_t x, uint32_t y, int u) {
    switch (u) {
        case 0: return prmt<0xC480>(x, y);
        case 1: return prmt<0xD591>(x, y);
        case 2: return prmt<0xE6A2>(x, y);
        default: return prmt<0xF7B3>(x, y);
    }
}
// sext32(byte u of x)
__device__ __forceinline__ int widen_one(uint32_t x, int u) {
    switch (u) {
        case 0: return (int)prmt<0x8880>(x, 0u);
        case 1: return (int)prmt<0x9991>(x, 0u);
        case 2: return (int)prmt<0xAAA2>(x, 0u);
        default: return (int)prmt<0xBBB3>(x, 0u);
    }
}

// Per-launch constants of the cell update, in the rung's value type.
template <int W> struct CellConsts {
    typename Val<W>::T NEGEXT;    // -ext, packed adder of VIADDMNMX
    typename Val<W>::T NEGOPEN;   // -open, 32-bit adder (IMAD)
    typename Val<W>::T PNEGOPEN;  // -open, packed adder
    typename Val<W>::T NEG;       // "-infinity"
    typename Val<W>::T FLOOR;     // local alignment: H >= 0 (the bias)
    unsigned one;
    __device__ __forceinline__ void init(const KParams& P) {
        typedef Val<W> V;
        NEGEXT = V::pdelta(-P.ge);
        NEGOPEN = V::delta(-P.go);
        PNEGOPEN = V::pdelta(-P.go);
        NEG = V::splat(P.neg + P.bias);
        FLOOR = V::splat(P.bias);
        one = P.one;
    }
};

// END mode (score + end cell) in the 16-bit rung, local alignment.  The end cell of Smith-Waterman is the FIRST
// maximum in (target column, query row) order (oracle/opal_oracle.c dp_end: `h > best`, columns outer).  A thread
// sees the columns of its K-row strip in ascending order, so it keeps one candidate per target (value, column,
// row) and replaces it when its strip's column maximum `cm` beats it -- or, in a later pass, ties it at an earlier
// column.  The row is looked up among the K registers only in that case: a divergent, rare path.  To keep it rare
// `thr` carries the best value any lane of the group has seen for the target (one butterfly shuffle per step):
// a strip maximum below it cannot be the target's maximum and is not even looked at.  The group's candidates are
// merged (value desc, column asc, row asc) when the target is done.
struct EndTrack {
    uint32_t thr;            // packed: best value known for the two targets (biased), >= bias + 1
    int bv0, bc0, bi0;       // candidate of the low-half target: biased value, column, tile row
    int bv1, bc1, bi1;       // high half
};

template <int K>
__device__ __forceinline__ void end_event(const uint32_t (&Gl)[K], uint32_t cm, EndTrack& T, int j, int rowbase, int go,
                                          bool c0, bool c1) {
    // between the sweeps Gl[r] holds M(r) - open in each half
    if (c0) {
        const int v = (int)(cm & 0xffffu);
        if (v > T.bv0 || (v == T.bv0 && j < T.bc0)) {
            const unsigned want = (unsigned)(v - go);
            int row = 0;
#pragma unroll
            for (int r = K - 1; r >= 0; --r)
                if ((Gl[r] & 0xffffu) == want) row = r;
            T.bv0 = v; T.bc0 = j; T.bi0 = rowbase + row;
        }
    }
    if (c1) {
        const int v = (int)(cm >> 16);
        if (v > T.bv1 || (v == T.bv1 && j < T.bc1)) {
            const unsigned want = (unsigned)(v - go);
            int row = 0;
#pragma unroll
            for (int r = K - 1; r >= 0; --r)
                if ((Gl[r] >> 16) == want) row = r;
            T.bv1 = v; T.bc1 = j; T.bi1 = rowbase + row;
        }
    }
    T.thr = __vmaxu2(T.thr, cm);
}

// One column of K rows held by one thread.  In: E, Gl = (E, H - open) of the
// previous column, upGprev = H - open of the row above in the previous column,
// F / Gup = (F, H - open) of the row above in this column.  Out: E, Gl of this
// column, F / Gup of the thread's last row (for the thread below), best.
//
// BFIX (long F chain only): with gap_extend > gap_open the padding rows above query row 0 cannot reproduce NW's
// penalised top border (re-opening a gap is then cheaper than extending it, so their E path beats the border
// value); the thread that holds the LAST padding row (register `fixrow`) overwrites its output with the border
// H[-1][j] - open (`fixG`) and an empty F, which is exactly what the first query row must see.
// DUAL (dual.cuh): `prof` holds one 32-bit word per (letter, row) with the folded scores of TWO queries already in the
// two u16 halves; t0 is the residue of the ONE target both are aligned to.  No widening instruction per cell pair.
template <int K, int W, bool LOCAL, bool PSMEM, bool WIDE, bool TDADD, bool SHORTF_ = (W == 16), bool BFIX = false,
          bool END = false, bool DUAL = false>
__device__ __forceinline__ void column_update(const KParams& P, const int8_t* prof, unsigned t0, unsigned t1, int pslot,
                                              const CellConsts<W>& C, typename Val<W>::T (&E)[K],
                                              typename Val<W>::T (&Gl)[K], typename Val<W>::T upGprev,
                                              typename Val<W>::T& F, typename Val<W>::T& Gup,
                                              typename Val<W>::T& best, int fixrow = -1,
                                              typename Val<W>::T fixG = 0, EndTrack* track = nullptr, int col = 0,
                                              int rowbase = 0) {
    static_assert(!BFIX || !SHORTF_, "the border fix lives in the long F chain");
    static_assert(!END || (W == 16 && SHORTF_ && LOCAL), "end-cell tracking: 16-bit rung, local alignment");
    typename Val<W>::T cm = C.FLOOR;  // END: maximum of this column over the thread's rows
    typedef Val<W> V;
    typedef typename V::T val_t;
    constexpr bool ADDPATH = TDADD || (W == 32);
    constexpr bool SHORTF = SHORTF_;          // one-instruction F chain: needs ext <= open (the host checks)
    const val_t NEGEXT = C.NEGEXT, NEGOPEN = C.NEGOPEN, PNEGOPEN = C.PNEGOPEN, FLOOR = C.FLOOR;
    const unsigned one = C.one;
    const int8_t* pa = prof + t0 * P.pitch + pslot;   // pslot: first profile slot of this lane's chunk
    const int8_t* pb = prof + t1 * P.pitch + pslot;
    // ---- sweep 1, bottom-up: E and M = max(Hdiag + S, E [, 0]); rows are
    // independent and M overwrites the previous column's G in place
#pragma unroll
    for (int c = K / 4 - 1; c >= 0; --c) {
        uint32_t wa = 0, wb = 0;
        int4 sw = make_int4(0, 0, 0, 0);
        if (WIDE) {
            sw = __ldg(reinterpret_cast<const int4*>(P.wprofile + (size_t)t0 * P.pitch + pslot) + c);
        } else if (DUAL) {
            // pitch and pslot count 32-bit words here; lane chunks start on 16-byte boundaries
            sw = *(reinterpret_cast<const int4*>(reinterpret_cast<const uint32_t*>(prof) + t0 * P.pitch + pslot) + c);
        } else if (PSMEM) {
            wa = *reinterpret_cast<const uint32_t*>(pa + 4 * c);
            if (W == 16) wb = *reinterpret_cast<const uint32_t*>(pb + 4 * c);
        } else {
            wa = __ldg(reinterpret_cast<const uint32_t*>(pa + 4 * c));
            if (W == 16) wb = __ldg(reinterpret_cast<const uint32_t*>(pb + 4 * c));
        }
#pragma unroll
        for (int u = 3; u >= 0; --u) {
            const int r = 4 * c + u;
            val_t S;
            if (WIDE || DUAL) S = (val_t)(u == 0 ? sw.x : u == 1 ? sw.y : u == 2 ? sw.z : sw.w);
            else if (W == 16) S = (val_t)widen_pair(wa, wb, u);
            else S = (val_t)widen_one(wa, u);
            const val_t Gd = (r == 0) ? upGprev : Gl[r - 1];
            E[r] = V::addmax(E[r], NEGEXT, Gl[r]);   // max(E - ext, Hleft - open)
            val_t M;
            if (ADDPATH) {
                const val_t Td = V::add(Gd, S, one);     // Hdiag + S  (S carries the +open)
                M = LOCAL ? V::max3(Td, E[r], FLOOR) : V::max2(Td, E[r]);
            } else {
                M = V::addmax(Gd, S, E[r]);
                if (LOCAL && SHORTF) M = V::max2(M, FLOOR);
            }
            if (SHORTF) {
                // F never sets a new maximum, so the running best only looks at M
                if (END) cm = V::max2(cm, M);
                else if (LOCAL) best = V::max2(best, M);
                Gl[r] = V::add(M, NEGOPEN, one);         // M - open
            } else {
                Gl[r] = M;
            }
        }
    }
    if constexpr (END) {
        // rare: this strip's column maximum reaches the best known for one of the two targets
        const uint32_t x = __vmaxu2((uint32_t)cm, track->thr) ^ (uint32_t)cm;  // half == 0 <=> cm >= thr there
        const bool c0 = (x & 0xffffu) == 0u, c1 = (x >> 16) == 0u;
        if (c0 || c1) end_event<K>(Gl, (uint32_t)cm, *track, col, rowbase, P.go, c0, c1);
    }
    // ---- sweep 2, top-down: the F chain, H, G = H - open
    if (SHORTF) {
        // ext <= open: F(r+1) = max(F(r) - ext, M(r) - open) does not need H(r), so the
        // serial chain is one instruction per row; G(r) = max(M(r) - open, F(r) - open)
        F = V::addmax(F, NEGEXT, Gup);               // F(0) = max(Fup - ext, Hup - open)
#pragma unroll
        for (int r = 0; r < K; ++r) {
            const val_t MO = Gl[r];
            Gup = V::addmax(F, PNEGOPEN, MO);        // G(r) = max(F(r), M(r)) - open
            Gl[r] = Gup;
            if (r < K - 1) F = V::addmax(F, NEGEXT, MO);   // F(r+1)
        }
    } else {
#pragma unroll
        for (int r = 0; r < K; ++r) {
            F = V::addmax(F, NEGEXT, Gup);               // max(F - ext, Hup - open)
            const val_t Hn = (LOCAL && !ADDPATH) ? V::max3(Gl[r], F, FLOOR) : V::max2(Gl[r], F);
            if (LOCAL) best = V::max2(best, Hn);
            Gup = V::add(Hn, NEGOPEN, one);
            if (BFIX && r == fixrow) { Gup = fixG; F = C.NEG; }
            Gl[r] = Gup;
        }
    }
}

// TDADD: every profile entry (S + open, padding included) is >= 0, so the
// 16-bit rung may form Hdiag + S with a 32-bit IMAD; otherwise that add stays
// in the packed adder of VIADDMNMX.  (Ignored by the 32-bit rung.)
template <int K, int W, bool LOCAL, bool PSMEM, bool WIDE = false, bool TDADD = true, bool BFIX = false, bool END = false>
__global__ void __launch_bounds__(512) search_kernel(const __grid_constant__ KParams P) {
    static_assert(!BFIX || (W == 32 && !LOCAL), "border fix: 32-bit rung, NW");
    static_assert(!END || W == 16, "end-cell tracking lives in the 16-bit rung");
    static_assert(K % 4 == 0, "K must be a multiple of 4");
    static_assert(!WIDE || (W == 32 && !PSMEM), "wide profile: 32-bit rung, global memory only");
    typedef Val<W> V;
    typedef typename V::T val_t;
    extern __shared__ __align__(16) unsigned char smem_raw[];

    const int nwork = *P.nwork;
    if (nwork <= 0) return;  // e.g. the rescue launch when nothing overflowed

    const int8_t* prof;
    if (PSMEM) {
        int8_t* sp = reinterpret_cast<int8_t*>(smem_raw);
        const int total = (P.A + 1) * P.pitch;
        for (int x = threadIdx.x; x < total; x += blockDim.x) {
            int letter = x / P.pitch, trow = x - letter * P.pitch;
            sp[x] = (int8_t)profile_entry(P, letter, trow);
        }
        __syncthreads();
        prof = sp;
    } else {
        prof = P.gprofile;
    }

    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int G = P.G;
    const int p = lane & (G - 1);
    const int grp = lane >> P.logG;
    const int gpw = 32 >> P.logG;
    const int T = G * K;
    const int ntasks = (nwork + gpw - 1) / gpw;
    // local alignment (SW) has both borders free: constants there, so the border arithmetic compiles out
    const bool top_free = LOCAL || (P.mode != MODE_NW);
    const bool left_free = LOCAL || (P.mode == MODE_OV);
    const int B = P.bias;
    CellConsts<W> C;
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
        const int entry = have ? P.worklist[work] : 0;
        const int pair = (W == 16) ? entry : (entry >> 1);
        const int half = entry & 1;
        const int64_t off = have ? P.pair_off[pair] : 0;
        const int lenA = have ? P.pair_len[2 * pair] : 0;
        const int lenB = have ? P.pair_len[2 * pair + 1] : 0;
        // columns this group walks; len0/len1 = lengths of the target(s) in the lo/hi half
        const int len0 = (W == 16) ? lenA : (half ? lenB : lenA);
        const int len1 = (W == 16) ? lenB : 0;
        const int ncol = len0;
        const int nsteps = __reduce_max_sync(FULL, ncol) + G - 1;
        // per-task base pointers: the step loop then indexes them with 32-bit columns (one IMAD.WIDE on the FMA
        // pipe per access instead of 64-bit adds on the ALU pipe)
        const uint16_t* const sp = P.stream + off;
        constexpr int BS = (W == 16) ? 1 : 2;   // boundary entries per column
        uint32_t* const bH = P.bndH + ((W == 16) ? off : (2 * off + half));
        uint32_t* const bF = P.bndF + ((W == 16) ? off : (2 * off + half));

        val_t best = FLOOR;                   // SW: running max of H
        int res0 = INT_MIN, res1 = INT_MIN;   // NW/HW/OV (true values)
        // END mode: SW candidates (EndTrack); HW/OV: column of the last-row maximum, best of the last column
        EndTrack track;
        track.thr = (uint32_t)V::splat(B + 1);
        track.bv0 = track.bv1 = B; track.bc0 = track.bc1 = -1; track.bi0 = track.bi1 = 0;
        int lc0 = -1, lc1 = -1, cv0 = INT_MIN, cv1 = INT_MIN, ci0 = -1, ci1 = -1;
        int sh_d = 1;                          // rotating butterfly distance for sharing track.thr

        for (int pass = 0; pass < P.npass; ++pass) {
            const int rowbase = pass * T + p * K;  // tile row of register 0
            val_t Gl[K], E[K];
#pragma unroll
            for (int r = 0; r < K; ++r) {
                int qi = rowbase + r - P.padtop;
                int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
                Gl[r] = V::splat(hl - P.go + B);
                E[r] = NEG;
            }
            val_t Gout = FLOOR, Fout = NEG, upGprev;
            {
                int qi = rowbase - 1 - P.padtop;
                int hl = (left_free || qi < 0) ? 0 : -(P.go + qi * P.ge);
                upGprev = V::splat(hl - P.go + B);
            }
            // software pipeline: the residue pair and the pass boundary of step s+1 are
            // fetched while step s computes
            unsigned code_nx = 0;
            val_t bG_nx = FLOOR, bF_nx = NEG;
            if (p == 0 && ncol > 0) {
                code_nx = sp[0];
                if (pass != 0) {
                    bG_nx = (val_t)__ldcg(bH);
                    bF_nx = (val_t)__ldcg(bF);
                }
            }
            for (int s = 0; s < nsteps; ++s) {
                val_t upG = __shfl_up_sync(FULL, Gout, 1, G);
                val_t upF = __shfl_up_sync(FULL, Fout, 1, G);
                if (END && LOCAL) {
                    track.thr = __vmaxu2(track.thr, __shfl_xor_sync(FULL, track.thr, sh_d));
                    sh_d <<= 1;
                    if (sh_d >= G) sh_d = 1;
                }
                const int j = s - p;
                const bool active = (unsigned)j < (unsigned)ncol;
                const unsigned code = code_nx;
                if (p == 0) {
                    if (pass == 0) {
                        int ht = top_free ? 0 : -(P.go + j * P.ge);
                        upG = V::splat(ht - P.go + B);
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
                            bG_nx = (val_t)__ldcg(bH + BS * jn);
                            bF_nx = (val_t)__ldcg(bF + BS * jn);
                        }
                    }
                }
                if (active) {
                    const unsigned t0 = (W == 16) ? (code & 0xffu) : (half ? (code >> 8) : (code & 0xffu));
                    const unsigned t1 = code >> 8;
                    val_t F = upF, Gup = upG;
                    if (BFIX) {
                        const int fr = P.padtop - 1 - rowbase;  // register that holds the last padding row, if any
                        const bool mine = (pass == 0) && (fr >= 0) && (fr < K);
                        column_update<K, W, LOCAL, PSMEM, WIDE, TDADD, false, BFIX>(
                            P, prof, t0, t1, (pass * G + p) * P.kstride, C, E, Gl, upGprev, F, Gup, best, mine ? fr : -1,
                            V::splat(-(P.go + j * P.ge) - P.go + B));
                    } else {
                        column_update<K, W, LOCAL, PSMEM, WIDE, TDADD, (W == 16), false, (END && LOCAL)>(
                            P, prof, t0, t1, (pass * G + p) * P.kstride, C, E, Gl, upGprev, F, Gup, best, -1, 0, &track, j, rowbase);
                    }
                    Gout = Gup;
                    Fout = F;
                    upGprev = upG;
                    if (p == G - 1) {
                        if (pass != lastpass) {
                            __stcg(bH + BS * j, (uint32_t)Gout);
                            __stcg(bF + BS * j, (uint32_t)Fout);
                        } else if (!LOCAL) {
                            // last query row: H = G + open
                            const int h0 = V::lo(Gout) + P.go - B, h1 = V::hi(Gout) + P.go - B;
                            if (P.mode == MODE_NW) {
                                if (j == len0 - 1) res0 = h0;
                                if (j == len1 - 1) res1 = h1;
                            } else if (END) {
                                // first maximum of the last row (oracle: `H[m - 1] > best`, columns ascending)
                                if (j < len0 && h0 > res0) { res0 = h0; lc0 = j; }
                                if (j < len1 && h1 > res1) { res1 = h1; lc1 = j; }
                            } else {
                                if (j < len0) res0 = max(res0, h0);
                                if (j < len1) res1 = max(res1, h1);
                            }
                        }
                    }
                    if (!LOCAL && P.mode == MODE_OV && (j == len0 - 1 || j == len1 - 1)) {
                        // last target column: every real query row is a candidate
#pragma unroll
                        for (int r = 0; r < K; ++r) {
                            if (rowbase + r >= P.padtop) {
                                if (END) {
                                    // lowest row of the last-column maximum (rows ascend within a thread)
                                    const int v0 = V::lo(Gl[r]) + P.go - B, v1 = V::hi(Gl[r]) + P.go - B;
                                    if (j == len0 - 1 && v0 > cv0) { cv0 = v0; ci0 = rowbase + r - P.padtop; }
                                    if (W == 16 && j == len1 - 1 && v1 > cv1) { cv1 = v1; ci1 = rowbase + r - P.padtop; }
                                } else {
                                    if (j == len0 - 1) res0 = max(res0, V::lo(Gl[r]) + P.go - B);
                                    if (W == 16 && j == len1 - 1) res1 = max(res1, V::hi(Gl[r]) + P.go - B);
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
        }

        // reduce over the group, lane p == 0 publishes
        int r0, r1;
        int eq0 = -1, et0 = -1, eq1 = -1, et1 = -1;   // END: end cell (query row, target column)
        if (END && LOCAL) {
            for (int d = 1; d < G; d <<= 1) {
                const int ov0 = __shfl_xor_sync(FULL, track.bv0, d), oc0 = __shfl_xor_sync(FULL, track.bc0, d),
                          oi0 = __shfl_xor_sync(FULL, track.bi0, d);
                if (ov0 > track.bv0 || (ov0 == track.bv0 && (oc0 < track.bc0 || (oc0 == track.bc0 && oi0 < track.bi0)))) {
                    track.bv0 = ov0; track.bc0 = oc0; track.bi0 = oi0;
                }
                const int ov1 = __shfl_xor_sync(FULL, track.bv1, d), oc1 = __shfl_xor_sync(FULL, track.bc1, d),
                          oi1 = __shfl_xor_sync(FULL, track.bi1, d);
                if (ov1 > track.bv1 || (ov1 == track.bv1 && (oc1 < track.bc1 || (oc1 == track.bc1 && oi1 < track.bi1)))) {
                    track.bv1 = ov1; track.bc1 = oc1; track.bi1 = oi1;
                }
            }
            r0 = track.bv0 - B; r1 = track.bv1 - B;
            if (track.bc0 >= 0) { eq0 = track.bi0 - P.padtop; et0 = track.bc0; }   // score 0: no alignment, (-1, -1)
            if (track.bc1 >= 0) { eq1 = track.bi1 - P.padtop; et1 = track.bc1; }
        } else if (END) {
            // NW: the corner.  HW: first last-row maximum (held by the group's last lane).  OV: the last row,
            // then the last column where it is strictly better (oracle/opal_oracle.c dp_end).
            const int lv0 = __shfl_sync(FULL, res0, G - 1, G), lv1 = __shfl_sync(FULL, res1, G - 1, G);
            const int lcol0 = __shfl_sync(FULL, lc0, G - 1, G), lcol1 = __shfl_sync(FULL, lc1, G - 1, G);
            r0 = lv0; r1 = lv1;
            eq0 = eq1 = P.Lq - 1;
            if (P.mode == MODE_NW) { et0 = len0 - 1; et1 = len1 - 1; }
            else { et0 = lcol0; et1 = lcol1; }
            if (P.mode == MODE_OV) {
                for (int d = 1; d < G; d <<= 1) {
                    const int ov0 = __shfl_xor_sync(FULL, cv0, d), oi0 = __shfl_xor_sync(FULL, ci0, d);
                    if (ov0 > cv0 || (ov0 == cv0 && oi0 < ci0)) { cv0 = ov0; ci0 = oi0; }
                    const int ov1 = __shfl_xor_sync(FULL, cv1, d), oi1 = __shfl_xor_sync(FULL, ci1, d);
                    if (ov1 > cv1 || (ov1 == cv1 && oi1 < ci1)) { cv1 = ov1; ci1 = oi1; }
                }
                if (cv0 > r0) { r0 = cv0; eq0 = ci0; et0 = len0 - 1; }
                if (cv1 > r1) { r1 = cv1; eq1 = ci1; et1 = len1 - 1; }
            }
        } else {
            if (LOCAL) { r0 = V::lo(best) - B; r1 = V::hi(best) - B; }
            else { r0 = res0; r1 = res1; }
            for (int d = 1; d < G; d <<= 1) {
                r0 = max(r0, __shfl_xor_sync(FULL, r0, d));
                r1 = max(r1, __shfl_xor_sync(FULL, r1, d));
            }
        }
        if (have && p == 0) {
            if (W == 16) {
                const int ia = P.pair_idx[2 * pair], ib = P.pair_idx[2 * pair + 1];
                const bool ina = ia >= P.out_start && ia < P.out_end;
                const bool inb = ib >= P.out_start && ib < P.out_end;
                // An overflowing LOW half carries into the high half through the 32-bit adds:
                // when target a overflows, target b is rescored as well.
                const bool ova = LOCAL && r0 >= P.ovf_thresh;
                const bool ovb = LOCAL && (r1 >= P.ovf_thresh || ova);
                if (ina) {
                    if (ova) P.ovf_list[atomicAdd(P.ovf_count, 1)] = pair << 1;
                    else {
                        P.out[ia - P.out_start] = r0;
                        if (END) { P.out_q[ia - P.out_start] = eq0; P.out_t[ia - P.out_start] = et0; }
                    }
                }
                if (inb) {
                    if (ovb) P.ovf_list[atomicAdd(P.ovf_count, 1)] = (pair << 1) | 1;
                    else {
                        P.out[ib - P.out_start] = r1;
                        if (END) { P.out_q[ib - P.out_start] = eq1; P.out_t[ib - P.out_start] = et1; }
                    }
                }
            } else {
                const int ix = P.pair_idx[2 * pair + half];
                if (ix >= P.out_start && ix < P.out_end) P.out[ix - P.out_start] = r0;
            }
        }
    }
}

// Materialise the profile in global memory (queries too long for shared
// memory, or scores outside int8).
__global__ void build_profile_kernel(const KParams P, int8_t* narrow, int32_t* wide);

typedef void (*search_kernel_t)(const KParams);

// defined in kernels_w*_l*.cu; nullptr when the combination is not built
search_kernel_t get_search_kernel_w16_l1(int K, bool psmem, bool tdadd);
search_kernel_t get_search_kernel_w16_l0(int K, bool psmem, bool tdadd);
// the same with end-cell tracking (END mode): kernels_w16_l*_end.cu
search_kernel_t get_search_kernel_w16_l1_end(int K, bool psmem, bool tdadd);
search_kernel_t get_search_kernel_w16_l0_end(int K, bool psmem, bool tdadd);
search_kernel_t get_search_kernel_w32_l1(int K, bool psmem);
search_kernel_t get_search_kernel_w32_l0(int K, bool psmem);
search_kernel_t get_search_kernel_wide(bool local);  // K = 32, W = 32
// NW with gap_extend > gap_open: K = 32, W = 32, profile in global memory (int8, or int32 when `wide`)
search_kernel_t get_search_kernel_bfix(bool wide);
// stream.cuh: the 16-bit score kernel as a continuous wavefront (whole-database searches)
search_kernel_t get_stream_kernel_l1(int K, bool tdadd);
search_kernel_t get_stream_kernel_l0(int K, bool tdadd);
// dual.cuh: two queries against one target per u16x2 word (Smith-Waterman, batched searches)
search_kernel_t get_dual_kernel(int K, bool tdadd);
// giant.cu: a team of warps per target (KParams.G = warps per team, KParams.npass = teams per CTA), K = 4 or 8
search_kernel_t get_giant_kernel(int K, bool local);
search_kernel_t get_giant_kernel_packed(int K, bool tdadd);   // SW, both targets of a pair at 16 bit (KParams.worklist: pair ids)
size_t giant_smem_bytes(int A, int pitch, int warps_per_team, int teams);
// tgiant.cu: the longest targets with the TARGET laid over the lanes of a team (KParams.G = warps), K = 16 or 32
search_kernel_t get_tgiant_kernel(int K, bool local);
size_t tgiant_smem_bytes(int A, int pitch, int warps, int Lq);

}  // namespace pyopal
