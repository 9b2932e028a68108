# cython: language_level=3, boundscheck=False, wraparound=False, initializedcheck=False
"""Cython binding of the C-ABI search entry point (include/pyopal_cuda.h).

The reference's host layer is Cython: every ``search${SIMD}`` plugin is one
``with nogil:`` block around ``opalSearchDatabase`` (platform/pyx.in:76-91).
This is that block for the device-resident store: typed memoryviews over the
caller's arrays, the GIL released around ``pyopal_cuda_search``.  It is what
INTEGRATION.md section 2 ("fast variant") describes; ``platform/cuda.py`` uses
it when the extension is built and the equivalent ``ctypes`` call otherwise.
"""

from libc.stdint cimport int32_t, int64_t, uintptr_t

cdef extern from "pyopal_cuda.h" nogil:
    ctypedef struct pyopal_cuda_db:
        pass
    const char* pyopal_cuda_last_error()
    int pyopal_cuda_device_count()
    int pyopal_cuda_search(pyopal_cuda_db* db, const unsigned char* query, int queryLength,
                           int gapOpen, int gapExt, const int* scoreMatrix, int alphabetLength,
                           int searchType, int mode, int overflowMethod, int64_t start, int64_t end,
                           int32_t* scores, int32_t* endQ, int32_t* endT, int32_t* startQ, int32_t* startT,
                           unsigned char* ops, const int64_t* ops_offsets, int32_t* ops_len)
    int pyopal_cuda_search_scores_device(pyopal_cuda_db* db, const unsigned char* query, int queryLength,
                                         int gapOpen, int gapExt, const int* scoreMatrix, int alphabetLength,
                                         int mode, int64_t start, int64_t end, int32_t* d_scores, void* stream)


def device_count():
    return pyopal_cuda_device_count()


def last_error():
    cdef const char* msg = pyopal_cuda_last_error()
    return msg.decode("utf-8", "replace") if msg != NULL else ""


def search(uintptr_t handle, const unsigned char[::1] query, int gap_open, int gap_extend,
           const int[::1] matrix not None, int alphabet_length, int search_type, int algorithm, int overflow,
           int64_t start, int64_t end, int[::1] scores, int[::1] end_q=None, int[::1] end_t=None,
           int[::1] start_q=None, int[::1] start_t=None, unsigned char[::1] ops=None,
           const long long[::1] ops_offsets=None, int[::1] ops_len=None):
    """One query against targets [start, end) of the store behind ``handle``; returns the C return code.

    Mirrors pyx.in:76-91: plain pointers into the caller's buffers, no GIL while the device works.
    """
    cdef pyopal_cuda_db* db = <pyopal_cuda_db*> handle
    cdef int qlen = <int> query.shape[0] if query is not None else 0
    cdef const unsigned char* q = &query[0] if qlen > 0 else NULL
    cdef int32_t* p_scores = <int32_t*> &scores[0] if scores is not None and scores.shape[0] > 0 else NULL
    cdef int32_t* p_eq = <int32_t*> &end_q[0] if end_q is not None and end_q.shape[0] > 0 else NULL
    cdef int32_t* p_et = <int32_t*> &end_t[0] if end_t is not None and end_t.shape[0] > 0 else NULL
    cdef int32_t* p_sq = <int32_t*> &start_q[0] if start_q is not None and start_q.shape[0] > 0 else NULL
    cdef int32_t* p_st = <int32_t*> &start_t[0] if start_t is not None and start_t.shape[0] > 0 else NULL
    cdef unsigned char* p_ops = &ops[0] if ops is not None and ops.shape[0] > 0 else NULL
    cdef const int64_t* p_oo = <const int64_t*> &ops_offsets[0] if ops_offsets is not None and ops_offsets.shape[0] > 0 else NULL
    cdef int32_t* p_ol = <int32_t*> &ops_len[0] if ops_len is not None and ops_len.shape[0] > 0 else NULL
    cdef int rc
    with nogil:
        rc = pyopal_cuda_search(db, q, qlen, gap_open, gap_extend, &matrix[0], alphabet_length, search_type,
                                algorithm, overflow, start, end, p_scores, p_eq, p_et, p_sq, p_st, p_ops, p_oo, p_ol)
    return rc


def search_scores_device(uintptr_t handle, const unsigned char[::1] query, int gap_open, int gap_extend,
                         const int[::1] matrix not None, int alphabet_length, int algorithm, int64_t start,
                         int64_t end, uintptr_t d_scores, uintptr_t stream):
    """Enqueue a score search whose output stays on the device (``d_scores``: device pointer)."""
    cdef pyopal_cuda_db* db = <pyopal_cuda_db*> handle
    cdef int qlen = <int> query.shape[0] if query is not None else 0
    cdef const unsigned char* q = &query[0] if qlen > 0 else NULL
    cdef int rc
    with nogil:
        rc = pyopal_cuda_search_scores_device(db, q, qlen, gap_open, gap_extend, &matrix[0], alphabet_length,
                                              algorithm, start, end, <int32_t*> d_scores, <void*> stream)
    return rc
