/*
 * include/pyopal_cuda.h -- C-ABI of the B200 (sm_100a) database-search backend.
 *
 * This is the drop-in boundary for PyOpal's hot path.  Everything here is
 * extern "C", plain pointers and sizes; no C++/torch types, no exceptions, and
 * every entry point may be called without the GIL from several host threads.
 *
 * Reference interfaces replaced (paths relative to /root/reference/):
 *   - opalSearchDatabase + OpalSearchResult + OPAL_* constants
 *         src/pyopal/opal.pxd:3-54         (the C declarations Cython binds)
 *         src/pyopal/lib.pxd:40-54         (searchfn_t: the plugin fn pointer)
 *         src/pyopal/platform/pyx.in:77-91 (the single call site)
 *   - opalInitSearchResult / opalSearchResultIsEmpty / opalSearchResultSetScore
 *         src/pyopal/opal.pxd:36-38        (declared, never called by pyopal)
 *   - the per-sequence heap store the call reads through
 *         src/pyopal/lib.pxd:95-98, src/pyopal/lib.pyx:501-508
 *     becomes the device-resident, length-sorted, pair-interleaved store
 *     behind pyopal_cuda_db_* below.
 *
 * Numeric values of the OPAL_* constants: upstream opal.h is not in the
 * reference tree (vendor/opal is an empty submodule); only the alignment ops
 * are pinned (M=0 D=1 I=2 X=3, lib.pyx:991-996).  The rest follow upstream
 * Opal as recalled in SURVEY.md App. A.2 and only need to be self-consistent,
 * because PyOpal never exposes them.
 */
#ifndef PYOPAL_CUDA_H
#define PYOPAL_CUDA_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- opal.pxd:5-24 ------------------------------------------------------ */
#define OPAL_ERR_OVERFLOW 1
#define OPAL_ERR_NO_SIMD_SUPPORT 2 /* here: "no usable CUDA device" -> RuntimeError, pyx.in:102-103 */
#define OPAL_ERR_INVALID_MODE 3
/* additions of this backend; pyx.in:106-107 maps any other non-zero code to RuntimeError */
#define PYOPAL_CUDA_ERR_CUDA 16     /* a CUDA runtime call failed; see pyopal_cuda_last_error() */
#define PYOPAL_CUDA_ERR_NOMEM 17    /* host or device allocation failed -> MemoryError */
#define PYOPAL_CUDA_ERR_ARGUMENT 18 /* NULL / out-of-range argument */
#define PYOPAL_CUDA_ERR_INTERNAL 19 /* self-check failed (e.g. traceback score mismatch) */
#define PYOPAL_CUDA_ERR_NCCL 20     /* NCCL missing or a collective failed; see pyopal_cuda_last_error() */

#define OPAL_MODE_NW 0
#define OPAL_MODE_HW 1
#define OPAL_MODE_OV 2
#define OPAL_MODE_SW 3

#define OPAL_OVERFLOW_SIMPLE 0
#define OPAL_OVERFLOW_BUCKETS 1

#define OPAL_SEARCH_SCORE 0
#define OPAL_SEARCH_SCORE_END 1
#define OPAL_SEARCH_ALIGNMENT 2

#define OPAL_ALIGN_MATCH 0
#define OPAL_ALIGN_DEL 1
#define OPAL_ALIGN_INS 2
#define OPAL_ALIGN_MISMATCH 3

/* ---- opal.pxd:26-34, field for field ------------------------------------ */
typedef struct {
    int scoreSet;
    int score;
    int endLocationTarget;
    int endLocationQuery;
    int startLocationTarget;
    int startLocationQuery;
    unsigned char* alignment; /* malloc()ed by the callee in ALIGNMENT mode, as upstream Opal does */
    int alignmentLength;
} OpalSearchResult;

/* opal.pxd:36-38 */
void opalInitSearchResult(OpalSearchResult* result);
int opalSearchResultIsEmpty(const OpalSearchResult result);
void opalSearchResultSetScore(OpalSearchResult* result, int score);

/* opal.pxd:40-54 / lib.pxd:40-54.  Generic path: host pointers in, results
 * out; the database is staged to the current device for the duration of the
 * call, so it works for any BaseDatabase subclass (lib.pyx:337-346). */
int opalSearchDatabase(unsigned char query[], int queryLength, unsigned char* db[], int dbLength,
                       int dbSeqLengths[], int gapOpen, int gapExt, int* scoreMatrix, int alphabetLength,
                       OpalSearchResult* results[], const int searchType, int mode, int overflowMethod);

/* opal.pxd:56-67: Smith-Waterman SCORE search over `char` sequences (declared by the interface PyOpal binds,
 * never called by it); same staging as opalSearchDatabase. */
int opalSearchDatabaseCharSW(char query[], int queryLength, char** db, int dbLength, int dbSeqLengths[],
                             int gapOpen, int gapExt, int* scoreMatrix, int alphabetLength,
                             OpalSearchResult* results[]);

/* ---- device-resident store (replaces lib.pxd:95-98 for the search) ------ */
typedef struct pyopal_cuda_db pyopal_cuda_db;

int pyopal_cuda_device_count(void);
/* name (<= cap bytes), SM count, total memory of `device`. */
int pyopal_cuda_device_info(int device, char* name, int cap, int* sm_count, int64_t* total_mem);
/* PCI bus id ("0000:1b:00.0") of `device`: lets a caller find the same GPU in NVML whatever CUDA_VISIBLE_DEVICES says */
int pyopal_cuda_device_pci_bus_id(int device, char* out, int cap);
/* thread-local message of the last failing call on this thread */
const char* pyopal_cuda_last_error(void);

/* Build the store on `device` from a packed host residue buffer:
 * sequence k is residues[offsets[k] .. offsets[k]+lengths[k]).  Residues are
 * alphabet indices < alphabetLength <= 32 (lib.pxd:28-32). */
int pyopal_cuda_db_create(int device, const unsigned char* residues, const int64_t* offsets,
                          const int32_t* lengths, int64_t n, int alphabetLength, pyopal_cuda_db** out);
/* Same, from the reference's own layout: an array of per-sequence pointers
 * (BaseDatabase.get_sequences / get_lengths, lib.pxd:90-92). */
int pyopal_cuda_db_create_from_pointers(int device, const unsigned char* const* sequences,
                                        const int32_t* lengths, int64_t n, int alphabetLength,
                                        pyopal_cuda_db** out);
void pyopal_cuda_db_destroy(pyopal_cuda_db* db);
int64_t pyopal_cuda_db_size(const pyopal_cuda_db* db);
int64_t pyopal_cuda_db_total_length(const pyopal_cuda_db* db);
int64_t pyopal_cuda_db_device_bytes(const pyopal_cuda_db* db);

/* One query against targets [start, end) of the store (lib.pyx:1364-1383).
 * Output arrays are HOST buffers of (end - start) entries, in database order.
 * endQ/endT may be NULL in SCORE mode; startQ/startT/ops* may be NULL unless
 * ALIGNMENT.  In ALIGNMENT mode `ops` receives target k's operations at
 * ops[ops_offsets[k] ..) (caller sizes slot k as queryLength + length_k, the
 * upper bound) and ops_len[k] their count. */
int pyopal_cuda_search(pyopal_cuda_db* db, const unsigned char* query, int queryLength, int gapOpen,
                       int gapExt, const int* scoreMatrix, int alphabetLength, int searchType, int mode,
                       int overflowMethod, int64_t start, int64_t end, int32_t* scores, int32_t* endQ,
                       int32_t* endT, int32_t* startQ, int32_t* startT, unsigned char* ops,
                       const int64_t* ops_offsets, int32_t* ops_len);

/* SCORE-mode search whose result stays on the device: `d_scores` is a DEVICE
 * pointer to (end - start) int32.  `stream` is a cudaStream_t (0 = the legacy
 * default stream); the call returns once the work is enqueued on it.  Every
 * call takes a workspace from the handle's pool; when a workspace is reused
 * its next search is ordered (by event) after this one, on whatever stream,
 * so concurrent calls from several threads or streams are safe. */
int pyopal_cuda_search_scores_device(pyopal_cuda_db* db, const unsigned char* query, int queryLength,
                                     int gapOpen, int gapExt, const int* scoreMatrix, int alphabetLength,
                                     int mode, int64_t start, int64_t end, int32_t* d_scores, void* stream);

/* Batched SCORE-mode search: nq queries (queries[offsets[i] .. +lengths[i]))
 * against the whole store; scores is a HOST buffer of nq * size int32,
 * row-major by query. */
int pyopal_cuda_search_batch(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                             const int32_t* q_lengths, int nq, int gapOpen, int gapExt,
                             const int* scoreMatrix, int alphabetLength, int mode, int32_t* scores);

/* The same with a device-side top-k per query (topk.cu): scores / indices are HOST buffers of nq * k entries,
 * row-major by query, each row sorted by score descending, ties by ascending database index; k <= size. */
int pyopal_cuda_search_batch_topk(pyopal_cuda_db* db, const unsigned char* queries, const int64_t* q_offsets,
                                  const int32_t* q_lengths, int nq, int gapOpen, int gapExt, const int* scoreMatrix,
                                  int alphabetLength, int mode, int k, int32_t* scores, int32_t* indices);

/* The k best hits of a SCORE-mode search, selected on the device (topk.cu): `scores` / `indices` are HOST
 * buffers of k entries, filled with min(k, end - start) hits sorted by score descending, ties by ascending
 * database index (absolute, like ScoreResult.target_index, pyx.in:69); *n_hits receives their number.
 * Additive: the reference returns every target (pyx.in:65-72) and leaves the sort to the caller
 * (docs/examples/abc.ipynb cells 23-24). */
int pyopal_cuda_search_topk(pyopal_cuda_db* db, const unsigned char* query, int queryLength, int gapOpen,
                            int gapExt, const int* scoreMatrix, int alphabetLength, int mode, int64_t start,
                            int64_t end, int k, int32_t* scores, int32_t* indices, int32_t* n_hits);

/* The same selection over a DEVICE array of n int32 scores (e.g. the gathered scores of a sharded search):
 * d_scores_out / d_indices_out are DEVICE buffers of k entries, written UNSORTED (every score above the k-th
 * largest, then the lowest-index ties); k <= n.  Enqueued on `stream`. */
int pyopal_cuda_topk_device(pyopal_cuda_db* db, const int32_t* d_scores, int64_t n, int k, int32_t* d_scores_out,
                            int32_t* d_indices_out, void* stream);

/* Statistics of the last search on this handle (for bench.py / profiles):
 * stats[0]=kernel launches, [1]=targets scored at 16 bit, [2]=targets scored
 * at 32 bit (static), [3]=targets escalated 16->32 after overflow,
 * [4]=device time of the last search in microseconds (CUDA events),
 * [5]=threads per target group G, [6]=query rows per thread K, [7]=passes,
 * [8]/[9]/[10]=device microseconds of the side stream (32-bit static + 16-bit
 * tail kernels) / the 16-bit bulk kernel / the 32-bit rescore kernel of the last search (0 until its stream
 * has completed), [11]=1 if the query profile was held in shared memory. */
int pyopal_cuda_last_stats(const pyopal_cuda_db* db, int64_t* stats, int n);

/* Host helper for packed databases: offsets[i] = sum of lengths[0..i), *total = sum of all (either may be NULL). */
/* Order in which a batch of queries is searched: the caller's, or with pair_queries != 0 (Smith-Waterman batches)
 * sorted by length (descending) with neighbours within 0.7 of each other paired: a pair goes down as one pass of the
 * two-queries-per-register kernel, which costs the longer query twice over.
 * units[2 * u], units[2 * u + 1] = the query indices of unit u (second = -1: searched alone); returns the number of
 * units (<= nq).  Pure host code; every rank of a multi-GPU handle derives the same units. */
int pyopal_host_batch_units(const int32_t* q_lengths, int nq, int pair_queries, int32_t* units);

void pyopal_host_offsets_from_lengths(const int32_t* lengths, int64_t n, int64_t* offsets, int64_t* total);

/* ---- one database over several GPUs (multi.cu) ---------------------------------------------------------
 * Replaces the ThreadPool fan-out of src/pyopal/_align.py:145-172: the database is dealt to the GPUs by
 * cumulative cells (length-sorted, snake order), every GPU searches its shard, the per-shard score arrays are
 * exchanged with ONE ncclAllGather over NVLink and un-permuted to database order on the device.  NCCL is bound
 * at run time (dlopen libnccl.so.2; PYOPAL_NCCL_LIB or pyopal_cuda_nccl_load() name another copy). */
typedef struct pyopal_cuda_mdb pyopal_cuda_mdb;

int pyopal_cuda_nccl_load(const char* path);          /* optional: bind this libnccl instead of the default one; a copy the
                                                          process has already mapped (e.g. PyTorch's) always wins */
int pyopal_cuda_nccl_loaded(void);                    /* 1 once NCCL is bound */
int pyopal_cuda_nccl_version(void);                   /* e.g. 22809, 0 if NCCL cannot be loaded */
int pyopal_cuda_nccl_unique_id(unsigned char id[128]); /* ncclGetUniqueId: rank 0 creates it, the caller distributes it */

/* owner[i] = rank of target i under the cell-balanced snake deal (pure host code). */
int pyopal_cuda_deal(const int32_t* lengths, int64_t n, int world, int32_t* owner);

/* ONE process driving `ndev` devices (ncclCommInitAll, one host thread per device): what Aligner.align uses. */
int pyopal_cuda_mdb_create(const int* devices, int ndev, const unsigned char* residues, const int64_t* offsets,
                           const int32_t* lengths, int64_t n, int alphabetLength, pyopal_cuda_mdb** out);
/* One process PER GPU (ncclCommInitRank): this process is `rank` of `world` and brings its own shard, whose
 * targets have the (ascending) database indices `global_index` (NULL with world 1: the identity).  Collective. */
int pyopal_cuda_mdb_create_rank(int device, int rank, int world, const unsigned char nccl_id[128],
                                const unsigned char* residues, const int64_t* offsets, const int32_t* lengths,
                                int64_t n_local, const int32_t* global_index, int64_t n_all, int alphabetLength,
                                pyopal_cuda_mdb** out);
/* The same when every rank holds the whole packed database on the host: the deal is computed inside. */
int pyopal_cuda_mdb_create_rank_dealt(int device, int rank, int world, const unsigned char nccl_id[128],
                                      const unsigned char* residues, const int64_t* offsets, const int32_t* lengths,
                                      int64_t n, int alphabetLength, pyopal_cuda_mdb** out);
void pyopal_cuda_mdb_destroy(pyopal_cuda_mdb* m);
int64_t pyopal_cuda_mdb_size(const pyopal_cuda_mdb* m);
int pyopal_cuda_mdb_world(const pyopal_cuda_mdb* m);
int pyopal_cuda_mdb_local_ranks(const pyopal_cuda_mdb* m);
int64_t pyopal_cuda_mdb_local_size(const pyopal_cuda_mdb* m, int local_rank);
int64_t pyopal_cuda_mdb_local_residues(const pyopal_cuda_mdb* m, int local_rank);
pyopal_cuda_db* pyopal_cuda_mdb_shard(pyopal_cuda_mdb* m, int local_rank); /* borrowed: statistics only */

/* pyopal_cuda_search over all GPUs: host outputs in database order (on the first local rank's process).
 * SCORE goes through the NCCL gather; END / ALIGNMENT need a single-process handle (per-target results, no
 * exchange).  In the one-process-per-GPU layout every rank must make the same call. */
int pyopal_cuda_mdb_search(pyopal_cuda_mdb* m, const unsigned char* query, int queryLength, int gapOpen, int gapExt,
                           const int* scoreMatrix, int alphabetLength, int searchType, int mode, int overflowMethod,
                           int64_t start, int64_t end, int32_t* scores, int32_t* endQ, int32_t* endT, int32_t* startQ,
                           int32_t* startT, unsigned char* ops, const int64_t* ops_offsets, int32_t* ops_len);
/* Batched SCORE search, stream-ordered on every GPU (Smith-Waterman queries of similar length two per pass):
 * scores = HOST nq * (end - start) int32 (row-major by query) or NULL to leave the results on the devices (timing). */
int pyopal_cuda_mdb_search_batch(pyopal_cuda_mdb* m, const unsigned char* queries, const int64_t* q_offsets,
                                 const int32_t* q_lengths, int nq, int gapOpen, int gapExt, const int* scoreMatrix,
                                 int alphabetLength, int mode, int64_t start, int64_t end, int32_t* scores);
/* The same with a device-side top-k per query over the gathered scores: HOST nq * k (score, index) rows. */
int pyopal_cuda_mdb_search_batch_topk(pyopal_cuda_mdb* m, const unsigned char* queries, const int64_t* q_offsets,
                                      const int32_t* q_lengths, int nq, int gapOpen, int gapExt, const int* scoreMatrix,
                                      int alphabetLength, int mode, int k, int32_t* scores, int32_t* indices);
/* Device-side clock (CUDA events on the handle's own streams) around a sequence of calls; end() returns the
 * longest elapsed time over this process's GPUs, in milliseconds. */
int pyopal_cuda_mdb_time_begin(pyopal_cuda_mdb* m);
int pyopal_cuda_mdb_time_end(pyopal_cuda_mdb* m, double* ms);
int64_t pyopal_cuda_mdb_last_launches(const pyopal_cuda_mdb* m); /* kernels + collectives of the last call, this process */

/* Integer-ALU issue microbenchmark used for the cell-update roofline
 * (SURVEY.md 8(d)): runs the six-instruction 16-bit cell mix dependency-free
 * on `device` and returns thread-instructions per second. */
int pyopal_cuda_int_issue_rate(int device, double* thread_instr_per_s, double* sm_clock_mhz);

#ifdef __cplusplus
}
#endif
#endif
