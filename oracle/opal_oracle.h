/*
 * oracle/opal_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * Scalar CPU restatement of the Opal database-search semantics that PyOpal
 * binds at /root/reference/src/pyopal/opal.pxd:40-54 (opalSearchDatabase) and
 * calls from /root/reference/src/pyopal/platform/pyx.in:77-91.
 *
 * The arithmetic itself lives in the un-vendored submodule
 * althonos/opal @ bc2f807cf99436a8467d26eb6b812936d7e003bb, which is absent
 * from /root/reference (vendor/opal is empty), so this file restates the
 * published algorithm (affine-gap SW / NW / HW / OV, SWIPE-style database
 * search) and is pinned by every golden value the reference's own tests and
 * doctests hold for the path (tests/test_oracle_golden.py).  What those
 * vectors do not pin (HW/OV values, tie-breaks for end/start cells and most
 * traceback ties) is "parity unpinned"; each such decision is one named
 * constant below.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library.  The product path
 * (pyopal_b200) never does.
 */
#ifndef OPAL_ORACLE_H
#define OPAL_ORACLE_H

#ifdef __cplusplus
extern "C" {
#endif

/* Same constant names as opal.pxd:5-24.  Values: upstream opal.h is absent;
 * the alignment ops are pinned by lib.pyx:991-996 (M=0 D=1 I=2 X=3), the rest
 * only have to be self-consistent. */
#define ORACLE_ERR_OVERFLOW 1
#define ORACLE_ERR_NO_SIMD_SUPPORT 2
#define ORACLE_ERR_INVALID_MODE 3

#define ORACLE_MODE_NW 0
#define ORACLE_MODE_HW 1
#define ORACLE_MODE_OV 2
#define ORACLE_MODE_SW 3

#define ORACLE_OVERFLOW_SIMPLE 0
#define ORACLE_OVERFLOW_BUCKETS 1

#define ORACLE_SEARCH_SCORE 0
#define ORACLE_SEARCH_SCORE_END 1
#define ORACLE_SEARCH_ALIGNMENT 2

#define ORACLE_ALIGN_MATCH 0
#define ORACLE_ALIGN_DEL 1
#define ORACLE_ALIGN_INS 2
#define ORACLE_ALIGN_MISMATCH 3

/* Field-for-field the struct at opal.pxd:26-34. */
typedef struct {
    int scoreSet;
    int score;
    int endLocationTarget;
    int endLocationQuery;
    int startLocationTarget;
    int startLocationQuery;
    unsigned char* alignment; /* malloc()ed by the oracle; free with oracle_free */
    int alignmentLength;
} OracleSearchResult;

/* Same argument list as opalSearchDatabase (opal.pxd:40-54). */
int oracle_opalSearchDatabase(unsigned char query[], int queryLength, unsigned char* db[], int dbLength,
                              int dbSeqLengths[], int gapOpen, int gapExt, int* scoreMatrix,
                              int alphabetLength, OracleSearchResult* results[], const int searchType,
                              int mode, int overflowMethod);

/* Flat-array convenience for ctypes / the CPU baseline: the database is one
 * packed residue buffer + int64 offsets; outputs are plain arrays.  `ops`
 * (may be NULL) receives the alignment of target k at ops + ops_offsets[k],
 * at most ops_cap[k] bytes.  nthreads > 1 splits targets in contiguous chunks
 * by count exactly as _align.py:151-170 does. */
int oracle_search_flat(const unsigned char* query, int queryLength, const unsigned char* residues,
                       const long long* offsets, const int* lengths, int n, int gapOpen, int gapExt,
                       const int* scoreMatrix, int alphabetLength, int searchType, int mode,
                       int* scores, int* endQ, int* endT, int* startQ, int* startT,
                       unsigned char* ops, const long long* ops_offsets, int* ops_len, int nthreads);

void oracle_free(void* p);

/* opal_simd.c: inter-sequence SIMD Smith-Waterman scores (the CPU baseline of bench.py). */
int oracle_sw_scores_simd(const unsigned char* query, int m, const unsigned char* residues, const long long* offsets,
                          const int* lengths, int n, int go, int ge, const int* S, int A, int* scores, int threads);

#ifdef __cplusplus
}
#endif
#endif
