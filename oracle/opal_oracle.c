/*
 * oracle/opal_oracle.c -- TEST INFRASTRUCTURE ONLY (see opal_oracle.h).
 *
 * Plain scalar affine-gap DP, 64-bit accumulators, no SIMD, no ladder.  It
 * restates what opalSearchDatabase (opal.pxd:40-54) is documented to compute:
 *   algorithms  lib.pyx:1290-1295 / _align.py:71-76  (nw, hw, ov, sw)
 *   search type lib.pyx:1278-1282                    (score, end, full)
 *   result use  lib.pyx:783-1119                     (fields read back)
 * Gap of length n costs open + (n-1)*ext (confirmed by
 * the golden scores 44 / 47 / 41,31,23).
 *
 * PARITY: pinned by the reference's own vectors for
 * NW/SW scores, end/start cells of those pairs and one traceback tie; every
 * other tie-break below is a documented choice ("parity unpinned").
 */
#include "opal_oracle.h"

#include <limits.h>
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef long long i64;
#define NEG_INF (LLONG_MIN / 4)

/* ---- documented, switchable tie-break choices ---------------------------
 * END cell: first cell in column-major order (target index, then query
 * index) that strictly improves the running best.
 * OV: last-row candidates of every column first, then the last column's
 * other rows top to bottom.
 * TRACEBACK at an H cell, walking backwards: diagonal first (pinned by
 * lib.pyx:1006-1010 -> 'IMMMXMIM'), then E (OPAL_ALIGN_INS, target-only
 * step), then F (OPAL_ALIGN_DEL, query-only step).
 * Inside a gap: prefer closing the gap (return to H) over extending it. */
#define TB_E_BEFORE_F 1
#define TB_CLOSE_GAP_FIRST 1

enum { END_CORNER, END_LASTROW, END_LASTROW_LASTCOL, END_ALL };
typedef struct {
    int top_free;  /* H[-1][j] = 0 : target prefix free            */
    int left_free; /* H[i][-1] = 0 : query prefix free             */
    int floor0;    /* local alignment: H >= 0                      */
    int endrule;
    int anchored;  /* reverse pass: border cells are end candidates too */
} Cfg;

static Cfg cfg_for_mode(int mode) {
    Cfg c;
    switch (mode) {
        case ORACLE_MODE_NW: c = (Cfg){0, 0, 0, END_CORNER, 0}; break;
        case ORACLE_MODE_HW: c = (Cfg){1, 0, 0, END_LASTROW, 0}; break;
        case ORACLE_MODE_OV: c = (Cfg){1, 1, 0, END_LASTROW_LASTCOL, 0}; break;
        default: c = (Cfg){1, 1, 1, END_ALL, 0}; break;
    }
    return c;
}

/* Reverse pass: the alignment is anchored at the reversed origin (= the end
 * cell), so both borders are penalised and there is no floor; where it may
 * stop follows the mode's end rule.  A forward HW/OV optimum may consume no
 * target (or no query) residue at all -- e.g. a long query against a 2-residue
 * target: every query residue deleted.  Seen from the anchor that alignment
 * stops ON the border, so in the anchored pass the border cell of the last
 * row (column -1) and of the last column (row -1) are candidates as well, and
 * are examined first.  (Degenerate case; what Opal reports there is unknown.) */
static Cfg cfg_for_reverse(int mode) {
    Cfg c = cfg_for_mode(mode);
    c.top_free = 0;
    c.left_free = 0;
    c.floor0 = 0;
    c.anchored = 1;
    return c;
}

static inline i64 border(int is_free, int k, int go, int ge) {
    if (k < 0) return 0;
    return is_free ? 0 : -((i64)go + (i64)k * ge);
}
static inline i64 max2(i64 a, i64 b) { return a > b ? a : b; }

/* Score + end cell.  Rows = query, columns = target (test_aligner.py:43-45). */
static void dp_end(const unsigned char* q, int m, const unsigned char* t, int n, int go, int ge,
                   const int* S, int A, Cfg c, i64* score, int* eq, int* et) {
    i64 best = c.floor0 ? 0 : NEG_INF;
    int bq = -1, bt = -1;
    if (m == 0 || n == 0) {
        /* Degenerate inputs (not exercised by the reference): the DP borders. */
        switch (c.endrule) {
            case END_CORNER:
                best = (m == 0) ? border(c.top_free, n - 1, go, ge) : border(c.left_free, m - 1, go, ge);
                if (m == 0 && n == 0) best = 0;
                bq = m - 1; bt = n - 1;
                break;
            case END_LASTROW:
                best = (m == 0) ? 0 : border(c.left_free, m - 1, go, ge);
                bq = m - 1; bt = -1;
                break;
            default: best = 0; break;
        }
        *score = best; *eq = bq; *et = bt;
        return;
    }
    i64* H = (i64*)malloc(sizeof(i64) * (size_t)m);
    i64* E = (i64*)malloc(sizeof(i64) * (size_t)m);
    if (c.anchored && (c.endrule == END_LASTROW || c.endrule == END_LASTROW_LASTCOL)) {
        best = border(c.left_free, m - 1, go, ge); bq = m - 1; bt = -1;
    }
    for (int j = 0; j < n; j++) {
        i64 hdiag = border(c.top_free, j - 1, go, ge);
        i64 hup = border(c.top_free, j, go, ge);
        i64 f = NEG_INF;
        const int tj = t[j];
        for (int i = 0; i < m; i++) {
            i64 hleft = (j == 0) ? border(c.left_free, i, go, ge) : H[i];
            i64 eleft = (j == 0) ? NEG_INF : E[i];
            i64 e = max2(eleft - ge, hleft - go);
            f = max2(f - ge, hup - go);
            i64 h = hdiag + S[q[i] * A + tj];
            h = max2(h, max2(e, f));
            if (c.floor0 && h < 0) h = 0;
            hdiag = hleft;
            H[i] = h; E[i] = e; hup = h;
            if (c.endrule == END_ALL && h > best) { best = h; bq = i; bt = j; }
        }
        if ((c.endrule == END_LASTROW || c.endrule == END_LASTROW_LASTCOL) && H[m - 1] > best) {
            best = H[m - 1]; bq = m - 1; bt = j;
        }
    }
    if (c.endrule == END_LASTROW_LASTCOL && c.anchored && border(c.top_free, n - 1, go, ge) > best) {
        best = border(c.top_free, n - 1, go, ge); bq = -1; bt = n - 1;
    }
    if (c.endrule == END_LASTROW_LASTCOL)
        for (int i = 0; i < m - 1; i++)
            if (H[i] > best) { best = H[i]; bq = i; bt = n - 1; }
    if (c.endrule == END_CORNER) { best = H[m - 1]; bq = m - 1; bt = n - 1; }
    free(H); free(E);
    *score = best; *eq = bq; *et = bt;
}

static unsigned char* reversed_copy(const unsigned char* s, int n) {
    unsigned char* r = (unsigned char*)malloc((size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) r[i] = s[n - 1 - i];
    return r;
}

/* Global alignment of q[0..m) x t[0..n) with traceback; returns ops in
 * forward order.  Both m, n >= 1. */
static int traceback(const unsigned char* q, int m, const unsigned char* t, int n, int go, int ge,
                     const int* S, int A, i64 expect, unsigned char** ops_out, int* len_out) {
    size_t cells = (size_t)m * (size_t)n;
    i64* H = (i64*)malloc(sizeof(i64) * cells);
    i64* E = (i64*)malloc(sizeof(i64) * cells);
    i64* F = (i64*)malloc(sizeof(i64) * cells);
    if (!H || !E || !F) { free(H); free(E); free(F); return -1; }
#define AT(M, i, j) M[(size_t)(j) * (size_t)m + (size_t)(i)]
#define HB(i, j) ((i) < 0 ? border(0, (j), go, ge) : ((j) < 0 ? border(0, (i), go, ge) : AT(H, i, j)))
    for (int j = 0; j < n; j++)
        for (int i = 0; i < m; i++) {
            i64 e = max2((j > 0 ? AT(E, i, j - 1) : NEG_INF) - ge, HB(i, j - 1) - go);
            i64 f = max2((i > 0 ? AT(F, i - 1, j) : NEG_INF) - ge, HB(i - 1, j) - go);
            i64 d = ((i == 0 && j == 0) ? 0 : HB(i - 1, j - 1)) + S[q[i] * A + t[j]];
            AT(E, i, j) = e; AT(F, i, j) = f; AT(H, i, j) = max2(d, max2(e, f));
        }
    int rc = (AT(H, m - 1, n - 1) == expect) ? 0 : -2;
    unsigned char* ops = (unsigned char*)malloc((size_t)m + (size_t)n);
    int len = 0, i = m - 1, j = n - 1, state = 0; /* 0=H 1=E 2=F */
    while (i >= 0 && j >= 0) {
        if (state == 0) {
            i64 d = ((i == 0 && j == 0) ? 0 : HB(i - 1, j - 1)) + S[q[i] * A + t[j]];
            i64 h = AT(H, i, j);
            if (h == d) {
                ops[len++] = (q[i] == t[j]) ? ORACLE_ALIGN_MATCH : ORACLE_ALIGN_MISMATCH;
                i--; j--;
            } else if (TB_E_BEFORE_F ? (h == AT(E, i, j)) : (h != AT(F, i, j))) {
                state = 1;
            } else {
                state = 2;
            }
        } else if (state == 1) {
            i64 e = AT(E, i, j);
            int can_close = (e == HB(i, j - 1) - go);
            int can_ext = (j > 0) && (e == AT(E, i, j - 1) - ge);
            ops[len++] = ORACLE_ALIGN_INS;
            state = (TB_CLOSE_GAP_FIRST ? can_close : !can_ext) ? 0 : 1;
            j--;
        } else {
            i64 f = AT(F, i, j);
            int can_close = (f == HB(i - 1, j) - go);
            int can_ext = (i > 0) && (f == AT(F, i - 1, j) - ge);
            ops[len++] = ORACLE_ALIGN_DEL;
            state = (TB_CLOSE_GAP_FIRST ? can_close : !can_ext) ? 0 : 2;
            i--;
        }
    }
    while (j >= 0) { ops[len++] = ORACLE_ALIGN_INS; j--; }
    while (i >= 0) { ops[len++] = ORACLE_ALIGN_DEL; i--; }
    for (int a = 0, b = len - 1; a < b; a++, b--) { unsigned char x = ops[a]; ops[a] = ops[b]; ops[b] = x; }
#undef AT
#undef HB
    free(H); free(E); free(F);
    *ops_out = ops; *len_out = len;
    return rc;
}

static int fits32(i64 v) { return v <= INT_MAX && v >= INT_MIN; }

static int search_one(const unsigned char* q, int m, const unsigned char* t, int n, int go, int ge,
                      const int* S, int A, int searchType, int mode, OracleSearchResult* r) {
    i64 score; int eq, et;
    dp_end(q, m, t, n, go, ge, S, A, cfg_for_mode(mode), &score, &eq, &et);
    if (!fits32(score)) return ORACLE_ERR_OVERFLOW;
    r->score = (int)score; r->scoreSet = 1;
    if (searchType == ORACLE_SEARCH_SCORE) return 0;
    r->endLocationQuery = eq; r->endLocationTarget = et;
    if (searchType == ORACLE_SEARCH_SCORE_END) return 0;
    /* full: start cell by a reverse pass over the prefixes ending at the end
     * cell, then a global traceback over [start..end]. */
    r->alignment = NULL; r->alignmentLength = 0;
    if (eq < 0 || et < 0) { /* SW score 0 / degenerate: no alignment */
        r->startLocationQuery = -1; r->startLocationTarget = -1;
        return 0;
    }
    int sq = 0, st = 0;
    if (mode != ORACLE_MODE_NW) {
        unsigned char* rq = reversed_copy(q, eq + 1);
        unsigned char* rt = reversed_copy(t, et + 1);
        i64 rscore; int req, ret;
        dp_end(rq, eq + 1, rt, et + 1, go, ge, S, A, cfg_for_reverse(mode), &rscore, &req, &ret);
        free(rq); free(rt);
        if (rscore != score) return 98;
        sq = eq - req; st = et - ret;
    }
    r->startLocationQuery = sq; r->startLocationTarget = st;
    if (sq > eq || st > et) {
        /* degenerate: the alignment is one gap (see cfg_for_reverse) */
        int len = (sq > eq) ? (et - st + 1) : (eq - sq + 1);
        if (sq > eq && st > et) len = 0;
        if (len > 0 && score != -((i64)go + (i64)(len - 1) * ge)) return 97;
        r->alignment = (unsigned char*)malloc((size_t)(len > 0 ? len : 1));
        memset(r->alignment, (sq > eq) ? ORACLE_ALIGN_INS : ORACLE_ALIGN_DEL, (size_t)len);
        r->alignmentLength = len;
        return 0;
    }
    int rc = traceback(q + sq, eq - sq + 1, t + st, et - st + 1, go, ge, S, A, score, &r->alignment,
                       &r->alignmentLength);
    return rc == 0 ? 0 : 97;
}

int oracle_opalSearchDatabase(unsigned char query[], int queryLength, unsigned char* db[], int dbLength,
                              int dbSeqLengths[], int gapOpen, int gapExt, int* scoreMatrix,
                              int alphabetLength, OracleSearchResult* results[], const int searchType,
                              int mode, int overflowMethod) {
    (void)overflowMethod; /* the ladder is result-invisible (lib.pyx:1283-1289) */
    if (mode < ORACLE_MODE_NW || mode > ORACLE_MODE_SW) return ORACLE_ERR_INVALID_MODE;
    if (searchType < ORACLE_SEARCH_SCORE || searchType > ORACLE_SEARCH_ALIGNMENT) return ORACLE_ERR_INVALID_MODE;
    for (int k = 0; k < dbLength; k++) {
        int rc = search_one(query, queryLength, db[k], dbSeqLengths[k], gapOpen, gapExt, scoreMatrix,
                            alphabetLength, searchType, mode, results[k]);
        if (rc) return rc;
    }
    return 0;
}

/* ---- flat-array front end (ctypes, CPU baseline) ------------------------ */

typedef struct {
    const unsigned char* query; int m;
    const unsigned char* residues; const i64* offsets; const int* lengths;
    int lo, hi, go, ge; const int* S; int A, searchType, mode;
    int *scores, *endQ, *endT, *startQ, *startT;
    unsigned char* ops; const i64* ops_offsets; int* ops_len;
    int rc;
} Job;

static void* run_job(void* arg) {
    Job* jb = (Job*)arg;
    for (int k = jb->lo; k < jb->hi; k++) {
        OracleSearchResult r;
        memset(&r, 0, sizeof r);
        r.endLocationQuery = r.endLocationTarget = r.startLocationQuery = r.startLocationTarget = -1;
        int rc = search_one(jb->query, jb->m, jb->residues + jb->offsets[k], jb->lengths[k], jb->go, jb->ge,
                            jb->S, jb->A, jb->searchType, jb->mode, &r);
        if (rc) { jb->rc = rc; free(r.alignment); return NULL; }
        jb->scores[k] = r.score;
        if (jb->endQ) jb->endQ[k] = r.endLocationQuery;
        if (jb->endT) jb->endT[k] = r.endLocationTarget;
        if (jb->startQ) jb->startQ[k] = r.startLocationQuery;
        if (jb->startT) jb->startT[k] = r.startLocationTarget;
        if (jb->ops_len) jb->ops_len[k] = r.alignmentLength;
        if (jb->ops && r.alignment) memcpy(jb->ops + jb->ops_offsets[k], r.alignment, (size_t)r.alignmentLength);
        free(r.alignment);
    }
    return NULL;
}

int oracle_search_flat(const unsigned char* query, int queryLength, const unsigned char* residues,
                       const long long* offsets, const int* lengths, int n, int gapOpen, int gapExt,
                       const int* scoreMatrix, int alphabetLength, int searchType, int mode,
                       int* scores, int* endQ, int* endT, int* startQ, int* startT,
                       unsigned char* ops, const long long* ops_offsets, int* ops_len, int nthreads) {
    if (mode < ORACLE_MODE_NW || mode > ORACLE_MODE_SW) return ORACLE_ERR_INVALID_MODE;
    if (searchType < ORACLE_SEARCH_SCORE || searchType > ORACLE_SEARCH_ALIGNMENT) return ORACLE_ERR_INVALID_MODE;
    if (n <= 0) return 0;
    if (nthreads < 1) nthreads = 1;
    if (nthreads > n) nthreads = n;
    /* contiguous chunks by sequence count, as _align.py:151-170 */
    int chunk = n / nthreads;
    int nchunks = (n + chunk - 1) / chunk;
    Job* jobs = (Job*)calloc((size_t)nchunks, sizeof(Job));
    pthread_t* th = (pthread_t*)calloc((size_t)nchunks, sizeof(pthread_t));
    for (int c = 0; c < nchunks; c++) {
        Job* jb = &jobs[c];
        jb->query = query; jb->m = queryLength; jb->residues = residues; jb->offsets = offsets;
        jb->lengths = lengths; jb->lo = c * chunk; jb->hi = (c + 1) * chunk > n ? n : (c + 1) * chunk;
        jb->go = gapOpen; jb->ge = gapExt; jb->S = scoreMatrix; jb->A = alphabetLength;
        jb->searchType = searchType; jb->mode = mode; jb->scores = scores; jb->endQ = endQ; jb->endT = endT;
        jb->startQ = startQ; jb->startT = startT; jb->ops = ops; jb->ops_offsets = ops_offsets;
        jb->ops_len = ops_len; jb->rc = 0;
        if (nchunks == 1) run_job(jb);
        else pthread_create(&th[c], NULL, run_job, jb);
    }
    int rc = 0;
    for (int c = 0; c < nchunks; c++) {
        if (nchunks > 1) pthread_join(th[c], NULL);
        if (jobs[c].rc && !rc) rc = jobs[c].rc;
    }
    free(jobs); free(th);
    return rc;
}

void oracle_free(void* p) { free(p); }
