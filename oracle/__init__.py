"""oracle -- TEST INFRASTRUCTURE ONLY.

ctypes front end of ``oracle/liboracle.so`` (oracle/opal_oracle.c), the scalar
CPU restatement of the Opal database search that PyOpal binds at
/root/reference/src/pyopal/opal.pxd:40-54.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package; the product
(``pyopal_b200``) never does.

Parity status: pinned by the reference's golden vectors for NW / SW (scores,
end / start cells, one traceback tie -- tests/test_oracle_golden.py); HW / OV
values and all other tie-breaks are "parity unpinned" (the Opal sources are
absent from /root/reference; see oracle/opal_oracle.c).
"""

import array
import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

MODES = {"nw": 0, "hw": 1, "ov": 2, "sw": 3}
SEARCH = {"score": 0, "end": 1, "full": 2}


def build(force=False):
    """Compile liboracle.so with gcc (oracle/Makefile)."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("opal_oracle.c", "opal_simd.c", "opal_oracle.h")]
    if force or not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(f) for f in srcs):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = ctypes.CDLL(so)
        vp, ci = ctypes.c_void_p, ctypes.c_int
        L.oracle_search_flat.restype = ci
        L.oracle_search_flat.argtypes = [vp, ci, vp, vp, vp, ci, ci, ci, vp, ci, ci, ci] + [vp] * 8 + [ci]
        L.oracle_sw_scores_simd.restype = ci
        L.oracle_sw_scores_simd.argtypes = [vp, ci, vp, vp, vp, ci, ci, ci, vp, ci, vp, ci]
        _LIB = L
    return _LIB


def _addr(a):
    if a is None:
        return None
    if isinstance(a, bytes):
        return ctypes.cast(ctypes.c_char_p(a), ctypes.c_void_p).value
    if isinstance(a, array.array):
        return a.buffer_info()[0]
    return ctypes.addressof(ctypes.c_char.from_buffer(a))


def search(query, targets, matrix, alphabet_length, gap_open, gap_extend, algorithm="sw", mode="score", threads=1):
    """Search encoded ``query`` (bytes) against encoded ``targets`` (list of bytes).

    ``matrix`` is the row-major int matrix (query letter x target letter).
    Returns a dict of lists: score, and for end/full: query_end, target_end,
    query_start, target_start, alignment (bytes of ops M=0 D=1 I=2 X=3).
    """
    n = len(targets)
    lengths = array.array("i", map(len, targets))
    offsets = array.array("q", bytes(8 * n))
    total = 0
    for i, ln in enumerate(lengths):
        offsets[i] = total
        total += ln
    residues = b"".join(targets) + b"\0"
    return search_flat(query, residues, offsets, lengths, matrix, alphabet_length, gap_open, gap_extend,
                       algorithm, mode, threads)


def search_flat(query, residues, offsets, lengths, matrix, alphabet_length, gap_open, gap_extend,
                algorithm="sw", mode="score", threads=1):
    n = len(lengths)
    m = matrix if isinstance(matrix, array.array) else array.array("i", matrix)
    st = SEARCH[mode]
    scores = array.array("i", bytes(4 * n))
    eq = et = sq = stt = ol = oo = ops = None
    if st >= 1:
        eq = array.array("i", bytes(4 * n))
        et = array.array("i", bytes(4 * n))
    if st == 2:
        sq = array.array("i", bytes(4 * n))
        stt = array.array("i", bytes(4 * n))
        ol = array.array("i", bytes(4 * n))
        oo = array.array("q", bytes(8 * n))
        total = 0
        for k in range(n):
            oo[k] = total
            total += len(query) + lengths[k]
        ops = bytearray(total + 1)
    rc = lib().oracle_search_flat(
        _addr(query), len(query), _addr(residues), _addr(offsets), _addr(lengths), n, gap_open, gap_extend,
        _addr(m), alphabet_length, st, MODES[algorithm], _addr(scores), _addr(eq), _addr(et), _addr(sq),
        _addr(stt), _addr(ops), _addr(oo), _addr(ol), threads,
    )
    if rc == 1:
        raise OverflowError("oracle: score overflow")
    if rc != 0:
        raise RuntimeError(f"oracle failed (code={rc})")
    out = {"score": list(scores)}
    if st >= 1:
        out["query_end"] = list(eq)
        out["target_end"] = list(et)
    if st == 2:
        out["query_start"] = list(sq)
        out["target_start"] = list(stt)
        out["alignment"] = [bytes(ops[oo[k] : oo[k] + ol[k]]) for k in range(n)]
    return out


def sw_scores_simd(query, residues, offsets, lengths, matrix, alphabet_length, gap_open, gap_extend, threads=1):
    """Smith-Waterman scores with the inter-sequence SIMD baseline (opal_simd.c): the CPU arm of bench.py."""
    n = len(lengths)
    m = matrix if isinstance(matrix, array.array) else array.array("i", matrix)
    scores = array.array("i", bytes(4 * n))
    rc = lib().oracle_sw_scores_simd(_addr(query), len(query), _addr(residues), _addr(offsets), _addr(lengths), n,
                                     gap_open, gap_extend, _addr(m), alphabet_length, _addr(scores), threads)
    if rc != 0:
        raise RuntimeError(f"oracle simd failed (code={rc})")
    return scores
