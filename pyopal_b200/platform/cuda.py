"""The CUDA (sm_100a) platform module: ``searchCUDA``.

This is the fifth sibling of the reference's per-ISA plugin modules
``pyopal.platform.{sse2,sse4,avx2,neon}``, all generated from
/root/reference/src/pyopal/platform/pyx.in.  ``searchCUDA`` has the signature
of ``search${SIMD}`` (pyx.in:16-27), builds the result list the same way
(pyx.in:62-72), maps return codes to the same exceptions (pyx.in:102-107) and
calls the C-ABI in ``libpyopal_cuda.so`` (include/pyopal_cuda.h) where the
template calls ``opalSearchDatabase`` (pyx.in:77-91).

There is no CPU fallback: if the shared library is missing, or no CUDA device
is visible, the call raises ``RuntimeError`` -- the class the reference raises
for "no supported SIMD backend available" (lib.pyx:1226-1227).
"""

import array
import ctypes
import os
import threading

UINT32_MAX = 0xFFFFFFFF

# include/pyopal_cuda.h
OPAL_ERR_OVERFLOW = 1
OPAL_ERR_NO_SIMD_SUPPORT = 2
OPAL_ERR_INVALID_MODE = 3
PYOPAL_CUDA_ERR_CUDA = 16
PYOPAL_CUDA_ERR_NOMEM = 17
PYOPAL_CUDA_ERR_ARGUMENT = 18
PYOPAL_CUDA_ERR_INTERNAL = 19
PYOPAL_CUDA_ERR_NCCL = 20

# below this many residues a database stays on one GPU unless PYOPAL_CUDA_DEVICES names several
MULTI_GPU_MIN_RESIDUES = 8_000_000

OPAL_MODE_NW, OPAL_MODE_HW, OPAL_MODE_OV, OPAL_MODE_SW = 0, 1, 2, 3
OPAL_OVERFLOW_SIMPLE, OPAL_OVERFLOW_BUCKETS = 0, 1
OPAL_SEARCH_SCORE, OPAL_SEARCH_SCORE_END, OPAL_SEARCH_ALIGNMENT = 0, 1, 2
OPAL_ALIGN_MATCH, OPAL_ALIGN_DEL, OPAL_ALIGN_INS, OPAL_ALIGN_MISMATCH = 0, 1, 2, 3

_LIB_NAME = "libpyopal_cuda.so"
_lib = None
_lib_lock = threading.Lock()

_c_i32p = ctypes.POINTER(ctypes.c_int32)
_c_i64p = ctypes.POINTER(ctypes.c_int64)
_c_u8p = ctypes.POINTER(ctypes.c_ubyte)


class OpalSearchResult(ctypes.Structure):
    """/root/reference/src/pyopal/opal.pxd:26-34."""

    _fields_ = [
        ("scoreSet", ctypes.c_int),
        ("score", ctypes.c_int),
        ("endLocationTarget", ctypes.c_int),
        ("endLocationQuery", ctypes.c_int),
        ("startLocationTarget", ctypes.c_int),
        ("startLocationQuery", ctypes.c_int),
        ("alignment", ctypes.POINTER(ctypes.c_ubyte)),
        ("alignmentLength", ctypes.c_int),
    ]


try:  # Cython binding of the hot call, built in-tree by csrc/build_host.sh; ctypes otherwise
    from . import _cuda_cy as _cy
except ImportError:  # pragma: no cover - extension not built
    _cy = None


def library_path():
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), _LIB_NAME)


def load_library():
    """Load ``libpyopal_cuda.so`` and declare every prototype of the header."""
    global _lib
    if _lib is not None:
        return _lib
    with _lib_lock:
        if _lib is not None:
            return _lib
        path = library_path()
        if not os.path.exists(path):
            raise RuntimeError(
                f"CUDA backend not built: {path} is missing (run `python -c 'import __graft_entry__ as g; g.build()'`); "
                "there is no CPU fallback"
            )
        lib = ctypes.CDLL(path)
        vp, ci, i64 = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64
        lib.pyopal_cuda_device_count.restype = ci
        lib.pyopal_cuda_device_count.argtypes = []
        lib.pyopal_cuda_device_info.restype = ci
        lib.pyopal_cuda_device_info.argtypes = [ci, ctypes.c_char_p, ci, ctypes.POINTER(ci), _c_i64p]
        lib.pyopal_cuda_device_pci_bus_id.restype = ci
        lib.pyopal_cuda_device_pci_bus_id.argtypes = [ci, ctypes.c_char_p, ci]
        lib.pyopal_cuda_last_error.restype = ctypes.c_char_p
        lib.pyopal_cuda_last_error.argtypes = []
        lib.pyopal_cuda_db_create.restype = ci
        lib.pyopal_cuda_db_create.argtypes = [ci, vp, vp, vp, i64, ci, ctypes.POINTER(vp)]
        lib.pyopal_cuda_db_create_from_pointers.restype = ci
        lib.pyopal_cuda_db_create_from_pointers.argtypes = [ci, vp, vp, i64, ci, ctypes.POINTER(vp)]
        lib.pyopal_cuda_db_destroy.restype = None
        lib.pyopal_cuda_db_destroy.argtypes = [vp]
        for name in ("pyopal_cuda_db_size", "pyopal_cuda_db_total_length", "pyopal_cuda_db_device_bytes"):
            getattr(lib, name).restype = i64
            getattr(lib, name).argtypes = [vp]
        lib.pyopal_cuda_search.restype = ci
        lib.pyopal_cuda_search.argtypes = [vp, vp, ci, ci, ci, vp, ci, ci, ci, ci, i64, i64] + [vp] * 8
        lib.pyopal_cuda_search_scores_device.restype = ci
        lib.pyopal_cuda_search_scores_device.argtypes = [vp, vp, ci, ci, ci, vp, ci, ci, i64, i64, vp, vp]
        lib.pyopal_cuda_search_batch.restype = ci
        lib.pyopal_cuda_search_batch.argtypes = [vp, vp, vp, vp, ci, ci, ci, vp, ci, ci, vp]
        lib.pyopal_cuda_search_topk.restype = ci
        lib.pyopal_cuda_search_topk.argtypes = [vp, vp, ci, ci, ci, vp, ci, ci, i64, i64, ci, vp, vp, ctypes.POINTER(ci)]
        lib.pyopal_cuda_topk_device.restype = ci
        lib.pyopal_cuda_topk_device.argtypes = [vp, vp, i64, ci, vp, vp, vp]
        lib.pyopal_cuda_last_stats.restype = ci
        lib.pyopal_cuda_last_stats.argtypes = [vp, _c_i64p, ci]
        lib.pyopal_cuda_int_issue_rate.restype = ci
        lib.pyopal_cuda_int_issue_rate.argtypes = [ci, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]
        lib.opalSearchDatabase.restype = ci
        lib.opalSearchDatabase.argtypes = [vp, ci, vp, ci, vp, ci, ci, vp, ci, vp, ci, ci, ci]
        lib.opalSearchDatabaseCharSW.restype = ci
        lib.opalSearchDatabaseCharSW.argtypes = [vp, ci, vp, ci, vp, ci, ci, vp, ci, vp]
        lib.pyopal_cuda_search_batch_topk.restype = ci
        lib.pyopal_cuda_search_batch_topk.argtypes = [vp, vp, vp, vp, ci, ci, ci, vp, ci, ci, ci, vp, vp]
        lib.pyopal_host_offsets_from_lengths.restype = None
        lib.pyopal_host_offsets_from_lengths.argtypes = [vp, i64, vp, _c_i64p]
        # several GPUs (csrc/multi.cu)
        lib.pyopal_cuda_nccl_load.restype = ci
        lib.pyopal_cuda_nccl_load.argtypes = [ctypes.c_char_p]
        lib.pyopal_cuda_nccl_version.restype = ci
        lib.pyopal_cuda_nccl_version.argtypes = []
        lib.pyopal_cuda_nccl_unique_id.restype = ci
        lib.pyopal_cuda_nccl_unique_id.argtypes = [vp]
        lib.pyopal_cuda_deal.restype = ci
        lib.pyopal_cuda_deal.argtypes = [vp, i64, ci, vp]
        lib.pyopal_cuda_mdb_create.restype = ci
        lib.pyopal_cuda_mdb_create.argtypes = [vp, ci, vp, vp, vp, i64, ci, ctypes.POINTER(vp)]
        lib.pyopal_cuda_mdb_create_rank.restype = ci
        lib.pyopal_cuda_mdb_create_rank.argtypes = [ci, ci, ci, vp, vp, vp, vp, i64, vp, i64, ci, ctypes.POINTER(vp)]
        lib.pyopal_cuda_mdb_create_rank_dealt.restype = ci
        lib.pyopal_cuda_mdb_create_rank_dealt.argtypes = [ci, ci, ci, vp, vp, vp, vp, i64, ci, ctypes.POINTER(vp)]
        lib.pyopal_cuda_mdb_destroy.restype = None
        lib.pyopal_cuda_mdb_destroy.argtypes = [vp]
        lib.pyopal_cuda_mdb_size.restype = i64
        lib.pyopal_cuda_mdb_size.argtypes = [vp]
        for name in ("pyopal_cuda_mdb_world", "pyopal_cuda_mdb_local_ranks"):
            getattr(lib, name).restype = ci
            getattr(lib, name).argtypes = [vp]
        for name in ("pyopal_cuda_mdb_local_size", "pyopal_cuda_mdb_local_residues"):
            getattr(lib, name).restype = i64
            getattr(lib, name).argtypes = [vp, ci]
        lib.pyopal_cuda_mdb_shard.restype = vp
        lib.pyopal_cuda_mdb_shard.argtypes = [vp, ci]
        lib.pyopal_cuda_mdb_search.restype = ci
        lib.pyopal_cuda_mdb_search.argtypes = [vp, vp, ci, ci, ci, vp, ci, ci, ci, ci, i64, i64] + [vp] * 8
        lib.pyopal_cuda_mdb_search_batch.restype = ci
        lib.pyopal_cuda_mdb_search_batch.argtypes = [vp, vp, vp, vp, ci, ci, ci, vp, ci, ci, i64, i64, vp]
        lib.pyopal_cuda_mdb_search_batch_topk.restype = ci
        lib.pyopal_cuda_mdb_search_batch_topk.argtypes = [vp, vp, vp, vp, ci, ci, ci, vp, ci, ci, ci, vp, vp]
        lib.pyopal_cuda_mdb_time_begin.restype = ci
        lib.pyopal_cuda_mdb_time_begin.argtypes = [vp]
        lib.pyopal_cuda_mdb_time_end.restype = ci
        lib.pyopal_cuda_mdb_time_end.argtypes = [vp, ctypes.POINTER(ctypes.c_double)]
        lib.pyopal_cuda_mdb_last_launches.restype = i64
        lib.pyopal_cuda_mdb_last_launches.argtypes = [vp]
        lib.opalInitSearchResult.restype = None
        lib.opalInitSearchResult.argtypes = [ctypes.POINTER(OpalSearchResult)]
        lib.opalSearchResultIsEmpty.restype = ci
        lib.opalSearchResultIsEmpty.argtypes = [OpalSearchResult]
        lib.opalSearchResultSetScore.restype = None
        lib.opalSearchResultSetScore.argtypes = [ctypes.POINTER(OpalSearchResult), ci]
        _lib = lib
    return _lib


def last_error():
    msg = load_library().pyopal_cuda_last_error()
    return msg.decode("utf-8", "replace") if msg else ""


def check(retcode):
    """Return-code -> exception mapping of pyx.in:102-107 (plus this backend's codes)."""
    if retcode == 0:
        return
    detail = last_error()
    if retcode == OPAL_ERR_NO_SIMD_SUPPORT:
        raise RuntimeError("no supported CUDA device available" + (f" ({detail})" if detail else ""))
    if retcode == OPAL_ERR_OVERFLOW:
        raise OverflowError("overflow detected while computing alignment scores")
    if retcode == PYOPAL_CUDA_ERR_NOMEM:
        raise MemoryError(detail or "allocation failed in the CUDA backend")
    if retcode == PYOPAL_CUDA_ERR_NCCL:
        raise RuntimeError("NCCL failure in the multi-GPU path" + (f": {detail}" if detail else ""))
    raise RuntimeError(f"failed to align to Opal database (code={retcode})" + (f": {detail}" if detail else ""))


def device_count():
    return load_library().pyopal_cuda_device_count()


def device_info(device=0):
    lib = load_library()
    name = ctypes.create_string_buffer(256)
    sms = ctypes.c_int(0)
    mem = ctypes.c_int64(0)
    check(lib.pyopal_cuda_device_info(device, name, 256, ctypes.byref(sms), ctypes.byref(mem)))
    return {"name": name.value.decode(), "sm_count": sms.value, "total_memory": mem.value}


def device_pci_bus_id(device=0):
    buf = ctypes.create_string_buffer(64)
    check(load_library().pyopal_cuda_device_pci_bus_id(device, buf, 64))
    return buf.value.decode()


def default_device():
    """Device of this process: ``PYOPAL_CUDA_DEVICE`` or ``LOCAL_RANK`` (one process per GPU), else 0."""
    for var in ("PYOPAL_CUDA_DEVICE", "LOCAL_RANK"):
        if os.environ.get(var, "").isdigit():
            n = device_count()
            return int(os.environ[var]) % n if n > 0 else 0
    return 0


def devices_forced():
    return bool(os.environ.get("PYOPAL_CUDA_DEVICES", "").strip())


def default_devices():
    """The GPUs a `Database` spreads over: ``PYOPAL_CUDA_DEVICES`` ("0,1,2" or "all"), else the one device of a
    one-process-per-GPU launch (``PYOPAL_CUDA_DEVICE`` / ``LOCAL_RANK``), else every visible device."""
    n = device_count()
    if n <= 0:
        return [0]
    env = os.environ.get("PYOPAL_CUDA_DEVICES", "").strip()
    if env and env != "all":
        picked = []
        for tok in env.split(","):
            tok = tok.strip()
            if not tok.isdigit() or int(tok) >= n:
                raise ValueError(f"PYOPAL_CUDA_DEVICES names device {tok!r}; {n} visible")
            if int(tok) not in picked:
                picked.append(int(tok))
        return picked or [0]
    if not env:
        for var in ("PYOPAL_CUDA_DEVICE", "LOCAL_RANK"):
            if os.environ.get(var, "").isdigit():
                return [int(os.environ[var]) % n]
    return list(range(n))


class _PyBuffer(ctypes.Structure):
    _fields_ = [("buf", ctypes.c_void_p), ("obj", ctypes.py_object), ("len", ctypes.c_ssize_t),
                ("itemsize", ctypes.c_ssize_t), ("readonly", ctypes.c_int), ("ndim", ctypes.c_int),
                ("format", ctypes.c_char_p), ("shape", ctypes.c_void_p), ("strides", ctypes.c_void_p),
                ("suboffsets", ctypes.c_void_p), ("internal", ctypes.c_void_p)]


def _buffer_address(obj):
    """Address of any C-contiguous buffer, read-only ones included (mmap, memoryview of bytes)."""
    view = _PyBuffer()
    api = ctypes.pythonapi
    api.PyObject_GetBuffer.argtypes = [ctypes.py_object, ctypes.POINTER(_PyBuffer), ctypes.c_int]
    api.PyObject_GetBuffer.restype = ctypes.c_int
    api.PyBuffer_Release.argtypes = [ctypes.POINTER(_PyBuffer)]
    api.PyBuffer_Release.restype = None
    if api.PyObject_GetBuffer(obj, ctypes.byref(view), 0) != 0:  # PyBUF_SIMPLE
        raise TypeError(f"expected a contiguous buffer, found {type(obj).__name__}")
    addr = view.buf
    api.PyBuffer_Release(ctypes.byref(view))
    return addr


def _addr(buf):
    """Address of a bytes / bytearray / array.array / memoryview / numpy / ctypes buffer without copying."""
    if buf is None:
        return None
    if isinstance(buf, bytes):
        return ctypes.cast(ctypes.c_char_p(buf), ctypes.c_void_p).value
    if isinstance(buf, array.array):
        return buf.buffer_info()[0]
    if isinstance(buf, ctypes.Array):
        return ctypes.addressof(buf)
    if hasattr(buf, "ctypes"):  # numpy array (must be C-contiguous)
        return buf.ctypes.data
    return _buffer_address(buf)


def as_int32_array(values):
    """``values`` as a C ``int32`` buffer: buffers of 4-byte integers are used as they are, anything else is copied."""
    if isinstance(values, array.array) and values.typecode == "i":
        return values
    try:
        view = memoryview(values)
        if view.itemsize == 4 and view.format in ("i", "l", "<i", "=i") and view.c_contiguous:
            return view.cast("B").cast("i")
    except TypeError:
        pass
    import operator

    return array.array("i", [operator.index(v) for v in values])


def offsets_from_lengths(lengths):
    """``(offsets array('q'), total)`` of a packed database, computed in C."""
    lens = as_int32_array(lengths)
    n = len(lens)
    offsets = array.array("q", bytes(8 * n))
    total = ctypes.c_int64(0)
    load_library().pyopal_host_offsets_from_lengths(_addr(lens) if n else None, n, _addr(offsets) if n else None,
                                                    ctypes.byref(total))
    return offsets, total.value


def has_byte_ge(data, limit):
    """True if any byte of the buffer is >= ``limit``; C speed, chunked so a multi-GB database is never copied whole."""
    view = memoryview(data).cast("B")
    allowed = bytes(range(min(max(limit, 0), 256)))
    step = 1 << 26
    for lo in range(0, view.nbytes, step):
        if bytes(view[lo : lo + step]).translate(None, allowed):
            return True
    return False


def make_store(residues, offsets, lengths, alphabet_length, devices=None):
    """Device store of a packed database over ``devices`` (one device: `DeviceStore`, several: `MultiDeviceStore`)."""
    devices = default_devices() if devices is None else list(devices)
    if len(devices) > 1:
        return MultiDeviceStore(residues, offsets, lengths, alphabet_length, devices)
    return DeviceStore(residues, offsets, lengths, alphabet_length, device=devices[0])


def make_store_from_sequences(sequences, alphabet_length, devices=None):
    devices = default_devices() if devices is None else list(devices)
    if len(devices) > 1:
        lengths = array.array("i", map(len, sequences))
        offsets, _ = offsets_from_lengths(lengths)
        return MultiDeviceStore(b"".join(sequences), offsets, lengths, alphabet_length, devices)
    return DeviceStore.from_sequences(sequences, alphabet_length, devices[0])


class DeviceStore:
    """Device-resident, length-sorted, pair-interleaved copy of a database.

    Replaces, for the search, the per-sequence heap store of
    /root/reference/src/pyopal/lib.pxd:95-98 (see include/pyopal_cuda.h).
    """

    def __init__(self, residues, offsets, lengths, alphabet_length, device=None):
        lib = load_library()
        self._lib = lib
        self._handle = ctypes.c_void_p(None)
        self.device = default_device() if device is None else device
        n = len(lengths)
        check(
            lib.pyopal_cuda_db_create(
                self.device, _addr(residues), _addr(offsets), _addr(lengths), n, alphabet_length,
                ctypes.byref(self._handle),
            )
        )
        self.size = n
        self.alphabet_length = alphabet_length

    @classmethod
    def from_sequences(cls, sequences, alphabet_length, device=None):
        """Build from a list of encoded ``bytes`` (one per target)."""
        lengths = array.array("i", map(len, sequences))
        offsets, _ = offsets_from_lengths(lengths)
        return cls(b"".join(sequences), offsets, lengths, alphabet_length, device)

    def close(self):
        handle = getattr(self, "_handle", None)
        if handle:
            self._handle = None
            self._lib.pyopal_cuda_db_destroy(handle)

    __del__ = close

    @property
    def handle(self):
        return self._handle

    def device_bytes(self):
        return self._lib.pyopal_cuda_db_device_bytes(self._handle)

    def stats(self):
        buf = (ctypes.c_int64 * 12)()
        check(self._lib.pyopal_cuda_last_stats(self._handle, buf, 12))
        return dict(zip(_STAT_KEYS, buf))

    def search(self, query, gap_open, gap_extend, int_matrix, search_type, algorithm, overflow=OPAL_OVERFLOW_BUCKETS,
               start=0, end=None, lengths=None):
        """Run one query; returns arrays (scores, endQ, endT, startQ, startT, ops, ops_offsets, ops_len)."""
        end = self.size if end is None or end > self.size else end
        n = max(end - start, 0)
        scores, eq, et, sq, st, ops, oo, ol = _result_arrays(n, search_type, len(query), lengths, start)
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        if _cy is not None and isinstance(query, (bytes, bytearray)):
            # the Cython binding (platform/_cuda_cy.pyx): typed memoryviews, `with nogil` around the C call
            rc = _cy.search(self._handle.value or 0, query, gap_open, gap_extend, matrix, alen, search_type, algorithm,
                            overflow, start, end, scores, eq, et, sq, st, ops, oo, ol)
        else:
            rc = self._lib.pyopal_cuda_search(
                self._handle, _addr(query), len(query), gap_open, gap_extend, _addr(matrix), alen, search_type,
                algorithm, overflow, start, end, _addr(scores), _addr(eq), _addr(et), _addr(sq), _addr(st),
                _addr(ops), _addr(oo), _addr(ol),
            )
        check(rc)
        return scores, eq, et, sq, st, ops, oo, ol

    def search_scores_device(self, query, gap_open, gap_extend, int_matrix, algorithm, d_scores, stream=0,
                             start=0, end=None):
        """Enqueue a score search whose output stays on the device (``d_scores``: device pointer)."""
        end = self.size if end is None or end > self.size else end
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        check(
            self._lib.pyopal_cuda_search_scores_device(
                self._handle, _addr(query), len(query), gap_open, gap_extend, _addr(matrix), alen, algorithm,
                start, end, d_scores, stream,
            )
        )

    def search_topk(self, query, gap_open, gap_extend, int_matrix, algorithm, k, start=0, end=None):
        """The ``k`` best hits of a score search, selected on the device: ``(scores, indices)`` as
        ``array('i')``, score descending, ties by ascending (absolute) database index."""
        end = self.size if end is None or end > self.size else end
        k = max(0, min(int(k), max(end - start, 0)))
        scores = array.array("i", bytes(4 * k))
        indices = array.array("i", bytes(4 * k))
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        n_hits = ctypes.c_int(0)
        check(
            self._lib.pyopal_cuda_search_topk(
                self._handle, _addr(query), len(query), gap_open, gap_extend, _addr(matrix), alen, algorithm,
                start, end, k, _addr(scores), _addr(indices), ctypes.byref(n_hits),
            )
        )
        return scores[: n_hits.value], indices[: n_hits.value]

    def topk_device(self, d_scores, n, k, d_scores_out, d_indices_out, stream=0):
        """Device-side selection over ``n`` int32 scores at device pointer ``d_scores`` (unsorted output)."""
        check(self._lib.pyopal_cuda_topk_device(self._handle, d_scores, n, k, d_scores_out, d_indices_out, stream))

    def kernel_times(self):
        """Resolve the per-kernel CUDA events of the last search (call after synchronising its stream)."""
        return self.stats()

    def search_batch(self, queries, gap_open, gap_extend, int_matrix, algorithm):
        """Score-mode search of several encoded queries; returns one ``array('i')`` of nq*size scores."""
        nq = len(queries)
        blob, qo, ql = _pack_queries(queries)
        scores = array.array("i", bytes(4 * nq * self.size))
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        check(
            self._lib.pyopal_cuda_search_batch(
                self._handle, _addr(blob), _addr(qo), _addr(ql), nq, gap_open, gap_extend, _addr(matrix), alen,
                algorithm, _addr(scores),
            )
        )
        return scores

    def search_batch_topk(self, queries, gap_open, gap_extend, int_matrix, algorithm, k):
        """The ``k`` best (score, database index) of every query, selected on the device: two ``array('i')`` of
        nq * k entries (rows sorted by score descending, ties by ascending index)."""
        k = max(1, min(int(k), self.size))
        blob, qo, ql = _pack_queries(queries)
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        vals = array.array("i", bytes(4 * len(queries) * k))
        idx = array.array("i", bytes(4 * len(queries) * k))
        check(
            self._lib.pyopal_cuda_search_batch_topk(
                self._handle, _addr(blob), _addr(qo), _addr(ql), len(queries), gap_open, gap_extend, _addr(matrix),
                alen, algorithm, k, _addr(vals), _addr(idx),
            )
        )
        return vals, idx


class MultiDeviceStore:
    """One database over several GPUs of this process (csrc/multi.cu): cell-balanced deal, one host thread per
    device, NCCL all-gather of the per-shard scores, database order restored on the device.

    Same search interface as `DeviceStore`; what `Database` builds when more than one GPU is visible -- the
    replacement of the reference's thread fan-out (_align.py:145-172).
    """

    def __init__(self, residues, offsets, lengths, alphabet_length, devices):
        lib = load_library()
        self._lib = lib
        self._handle = ctypes.c_void_p(None)
        self.devices = list(devices)
        n = len(lengths)
        lengths = as_int32_array(lengths)
        devs = (ctypes.c_int * len(self.devices))(*self.devices)
        if len(self.devices) > 1:
            self._ensure_nccl()
        check(
            lib.pyopal_cuda_mdb_create(devs, len(self.devices), _addr(residues), _addr(offsets), _addr(lengths), n,
                                       alphabet_length, ctypes.byref(self._handle))
        )
        self.size = n
        self.alphabet_length = alphabet_length
        self.device = self.devices[0]

    @classmethod
    def _wrap(cls, handle, size, alphabet_length, devices):
        self = cls.__new__(cls)
        self._lib = load_library()
        self._handle = handle
        self.devices = list(devices)
        self.device = self.devices[0]
        self.size = size
        self.alphabet_length = alphabet_length
        return self

    @staticmethod
    def _ensure_nccl():
        """Bind NCCL (only handles that span several GPUs need it).

        Order: a copy this process has already mapped (e.g. by an earlier ``import torch``), else the one shipped
        with the ``nvidia-nccl-cu12`` wheel (found by path, not imported), else the system's.  The wheel goes before
        the system copy on purpose: libraries are matched by soname, so whichever ``libnccl.so.2`` is mapped first
        is the one every later ``import`` in the process gets, and an older system NCCL under a newer PyTorch makes
        ``import torch`` fail with an undefined symbol.
        """
        lib = load_library()
        if lib.pyopal_cuda_nccl_loaded():
            return
        import sys

        for base in sys.path:
            cand = os.path.join(base, "nvidia", "nccl", "lib", "libnccl.so.2")
            if os.path.exists(cand) and lib.pyopal_cuda_nccl_load(cand.encode()) == 0:
                return
        lib.pyopal_cuda_nccl_version()  # default search: PYOPAL_NCCL_LIB, libnccl.so.2 on the loader's path

    def close(self):
        handle = getattr(self, "_handle", None)
        if handle:
            self._handle = None
            self._lib.pyopal_cuda_mdb_destroy(handle)

    __del__ = close

    @property
    def handle(self):
        return self._handle

    @property
    def world(self):
        return self._lib.pyopal_cuda_mdb_world(self._handle)

    def local_residues(self, local_rank=0):
        return self._lib.pyopal_cuda_mdb_local_residues(self._handle, local_rank)

    def stats(self, local_rank=0):
        """Statistics of the last search on one shard (same keys as `DeviceStore.stats`)."""
        shard = self._lib.pyopal_cuda_mdb_shard(self._handle, local_rank)
        buf = (ctypes.c_int64 * 12)()
        check(self._lib.pyopal_cuda_last_stats(shard, buf, 12))
        out = dict(zip(_STAT_KEYS, buf))
        out["launches"] = self._lib.pyopal_cuda_mdb_last_launches(self._handle)
        return out

    def device_bytes(self, local_rank=0):
        return self._lib.pyopal_cuda_db_device_bytes(self._lib.pyopal_cuda_mdb_shard(self._handle, local_rank))

    def search(self, query, gap_open, gap_extend, int_matrix, search_type, algorithm, overflow=OPAL_OVERFLOW_BUCKETS,
               start=0, end=None, lengths=None):
        end = self.size if end is None or end > self.size else end
        n = max(end - start, 0)
        scores, eq, et, sq, st, ops, oo, ol = _result_arrays(n, search_type, len(query), lengths, start)
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        check(
            self._lib.pyopal_cuda_mdb_search(
                self._handle, _addr(query), len(query), gap_open, gap_extend, _addr(matrix), alen, search_type,
                algorithm, overflow, start, end, _addr(scores), _addr(eq), _addr(et), _addr(sq), _addr(st),
                _addr(ops), _addr(oo), _addr(ol),
            )
        )
        return scores, eq, et, sq, st, ops, oo, ol

    def search_batch(self, queries, gap_open, gap_extend, int_matrix, algorithm, start=0, end=None, to_host=True, out=None):
        """Score-mode search of several encoded queries, stream-ordered on every GPU; one ``array('i')`` of
        nq * (end - start) scores, or ``None`` with ``to_host=False`` (results stay on the devices: timing).
        ``out``: a writable int32 buffer of that size to receive the scores instead of a fresh array (a caller that
        searches at a high rate reuses one: allocating and zero-filling tens of MB per call costs milliseconds)."""
        end = self.size if end is None or end > self.size else end
        blob, qo, ql = _pack_queries(queries)
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        need = len(queries) * max(end - start, 0)
        if not to_host:
            scores = None
        elif out is not None:
            view = memoryview(out)
            if view.readonly or view.nbytes != 4 * need or not view.c_contiguous:
                raise ValueError("out must be a writable contiguous buffer of nq * (end - start) int32")
            scores = out
        else:
            scores = array.array("i", bytes(4 * need))
        check(
            self._lib.pyopal_cuda_mdb_search_batch(
                self._handle, _addr(blob), _addr(qo), _addr(ql), len(queries), gap_open, gap_extend, _addr(matrix),
                alen, algorithm, start, end, _addr(scores),
            )
        )
        return scores

    def search_batch_topk(self, queries, gap_open, gap_extend, int_matrix, algorithm, k):
        """The ``k`` best (score, database index) of every query: two ``array('i')`` of nq * k entries."""
        k = max(1, min(int(k), self.size))
        blob, qo, ql = _pack_queries(queries)
        matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
        alen = int(round(len(matrix) ** 0.5))
        vals = array.array("i", bytes(4 * len(queries) * k))
        idx = array.array("i", bytes(4 * len(queries) * k))
        check(
            self._lib.pyopal_cuda_mdb_search_batch_topk(
                self._handle, _addr(blob), _addr(qo), _addr(ql), len(queries), gap_open, gap_extend, _addr(matrix),
                alen, algorithm, k, _addr(vals), _addr(idx),
            )
        )
        return vals, idx

    def search_topk(self, query, gap_open, gap_extend, int_matrix, algorithm, k, start=0, end=None):
        end = self.size if end is None or end > self.size else end
        if (start, end) != (0, self.size):  # a slice: select on the host (the batched top-k runs over the whole database)
            sc = self.search(query, gap_open, gap_extend, int_matrix, OPAL_SEARCH_SCORE, algorithm, start=start, end=end)[0]
            order = sorted(range(len(sc)), key=lambda i: (-sc[i], i))[: max(0, int(k))]
            return array.array("i", [sc[i] for i in order]), array.array("i", [start + i for i in order])
        k = max(0, min(int(k), self.size))
        if k == 0:
            return array.array("i"), array.array("i")
        return self.search_batch_topk([query], gap_open, gap_extend, int_matrix, algorithm, k)

    def time_begin(self):
        check(self._lib.pyopal_cuda_mdb_time_begin(self._handle))

    def time_end(self):
        ms = ctypes.c_double(0)
        check(self._lib.pyopal_cuda_mdb_time_end(self._handle, ctypes.byref(ms)))
        return ms.value


def nccl_unique_id():
    """128 bytes identifying a new communicator: created on rank 0, handed to every rank by the caller."""
    MultiDeviceStore._ensure_nccl()
    buf = ctypes.create_string_buffer(128)
    check(load_library().pyopal_cuda_nccl_unique_id(buf))
    return buf.raw


def deal(lengths, world):
    """owner[i] = rank of target i under the cell-balanced snake deal (host code of csrc/multi.cu)."""
    lens = as_int32_array(lengths)
    owner = array.array("i", bytes(4 * len(lens)))
    check(load_library().pyopal_cuda_deal(_addr(lens) if len(lens) else None, len(lens), world,
                                          _addr(owner) if len(lens) else None))
    return owner


_STAT_KEYS = ("launches", "targets16", "targets32", "escalated", "device_us", "G", "K", "passes",
              "k32_us", "k16_us", "rescore_us", "profile_in_smem")


def _pack_queries(queries):
    ql = array.array("i", map(len, queries))
    qo, _ = offsets_from_lengths(ql)
    return b"".join(bytes(q) for q in queries), qo, ql


def _result_arrays(n, search_type, qlen, lengths, start):
    scores = array.array("i", bytes(4 * n))
    eq = et = sq = st = ops = oo = ol = None
    if search_type >= OPAL_SEARCH_SCORE_END:
        eq = array.array("i", bytes(4 * n))
        et = array.array("i", bytes(4 * n))
    if search_type == OPAL_SEARCH_ALIGNMENT:
        sq = array.array("i", bytes(4 * n))
        st = array.array("i", bytes(4 * n))
        ol = array.array("i", bytes(4 * n))
        oo = array.array("q", bytes(8 * n))
        total = 0
        for k in range(n):
            oo[k] = total
            total += qlen + lengths[start + k]
        ops = bytearray(total + 1)
    return scores, eq, et, sq, st, ops, oo, ol


def _search_generic(encoded, sequences, lengths, mode, overflow, algorithm, gap_open, gap_extend, matrix, alen):
    """Any ``BaseDatabase`` subclass: the C function with the reference's own signature (opal.pxd:40-54)."""
    lib = load_library()
    n = len(sequences)
    results_raw = (OpalSearchResult * n)()
    for r in results_raw:
        lib.opalInitSearchResult(ctypes.byref(r))
    ptrs = (ctypes.c_void_p * n)(*[ctypes.addressof(results_raw[i]) for i in range(n)])
    seqs = (ctypes.c_void_p * n)(*[_addr(s) if len(s) else None for s in sequences])
    lens = array.array("i", lengths)
    rc = lib.opalSearchDatabase(
        _addr(encoded), len(encoded), seqs, n, _addr(lens), gap_open, gap_extend, _addr(matrix), alen, ptrs,
        mode, algorithm, overflow,
    )
    check(rc)
    return results_raw


def searchCUDA(encoded, database, mode, overflow, algorithm, gap_open, gap_extend, int_matrix, start=0, end=UINT32_MAX):
    """``search${SIMD}`` of pyx.in:16-108 for the CUDA backend."""
    from ..lib import ScoreResult, EndResult, FullResult, Database  # late: lib imports this module

    if not isinstance(encoded, bytes):
        raise TypeError("expected bytes for the encoded query")
    matrix = int_matrix if isinstance(int_matrix, array.array) else array.array("i", int_matrix)
    alen = int(len(matrix) ** 0.5)  # pyx.in:41
    if mode == OPAL_SEARCH_SCORE:
        result_type = ScoreResult
    elif mode == OPAL_SEARCH_SCORE_END:
        result_type = EndResult
    elif mode == OPAL_SEARCH_ALIGNMENT:
        result_type = FullResult
    else:
        raise ValueError("invalid search mode")

    dbsize = database.get_size()
    if end > dbsize:
        end = dbsize
    if start > end:  # the reference wraps around here; raise instead
        raise IndexError("database slice start is past the end")
    size = 0 if end == 0 else end - start  # pyx.in:62
    if size == 0:
        return []
    qlen = len(encoded)
    lengths = database.get_lengths()

    if isinstance(database, Database):
        store = database._device_store()
        scores, eq, et, sq, st, ops, oo, ol = store.search(
            encoded, gap_open, gap_extend, matrix, mode, algorithm, overflow, start, end, lengths
        )
    else:
        sequences = database.get_sequences()
        raw = _search_generic(
            encoded, sequences[start:end], lengths[start:end], mode, overflow, algorithm, gap_open, gap_extend,
            matrix, alen,
        )
        scores = [r.score for r in raw]
        eq = [r.endLocationQuery for r in raw]
        et = [r.endLocationTarget for r in raw]
        sq = [r.startLocationQuery for r in raw]
        st = [r.startLocationTarget for r in raw]
        ol = [r.alignmentLength for r in raw]
        oo = None
        ops = [bytes(bytearray(r.alignment[: r.alignmentLength])) if r.alignmentLength else b"" for r in raw]
        libc = ctypes.CDLL(None)
        libc.free.argtypes = [ctypes.c_void_p]
        for r in raw:
            if r.alignment:
                libc.free(ctypes.cast(r.alignment, ctypes.c_void_p))

    if oo is not None or mode != OPAL_SEARCH_ALIGNMENT:
        # arrays straight from the C-ABI: the list is built natively (csrc/_results.c), the
        # equivalent of the reference's Cython loop at pyx.in:65-72 / 95-99
        from .. import _results

        lens = None
        if mode == OPAL_SEARCH_ALIGNMENT:
            lens = lengths if isinstance(lengths, array.array) else array.array("i", lengths)
        as_i32 = lambda a: a if a is None or isinstance(a, array.array) else array.array("i", a)
        return _results.make_results(result_type, mode, start, size, as_i32(scores), as_i32(eq), as_i32(et),
                                     as_i32(sq), as_i32(st), qlen, lens, ops, oo, as_i32(ol))
    # generic path in full mode: one bytes object per alignment already
    new = result_type.__new__
    results = [None] * size
    for i in range(size):
        r = new(result_type)
        r._target_index = start + i
        r._score = scores[i]
        r._query_end = eq[i]
        r._target_end = et[i]
        r._query_start = sq[i]
        r._target_start = st[i]
        r._query_length = qlen  # pyx.in:95-99
        r._target_length = lengths[start + i]
        r._alignment = ops[i]
        results[i] = r
    return results
