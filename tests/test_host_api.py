"""Host-side mirror of the reference API (no GPU needed).

Acceptance checks modelled on the reference's own suite,
/root/reference/src/pyopal/tests/{test_alphabet,test_database,test_result,test_aligner}.py:
same constructors, reprs, container semantics, pickling and exception classes,
plus: the C-ABI library loads and exports every symbol include/pyopal_cuda.h
declares, and the product refuses to run without a CUDA device (no fallback).
"""

import os
import pickle
import re

import pytest

import pyopal_b200 as pyopal
from pyopal_b200.platform import cuda

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


# --- Alphabet (test_alphabet.py) ------------------------------------------------


def test_alphabet_basics():
    default = pyopal.Alphabet()
    assert len(default) == 24 == len(default.letters)
    assert repr(default) == "Alphabet()"
    dna = pyopal.Alphabet("ATGC")
    assert len(dna) == 4 and str(dna) == "ATGC" and dna.letters == "ATGC"
    assert repr(dna) == "Alphabet('ATGC')"
    assert "A" in dna and "T" in dna and "X" not in dna
    assert (dna[0], dna[2], dna[-1], dna[-2]) == ("A", "G", "C", "G")
    for bad in (-5, 4, 5):
        with pytest.raises(IndexError):
            dna[bad]


def test_alphabet_eq_pickle():
    a1, a2, a3 = pyopal.Alphabet("ATGC"), pyopal.Alphabet("ATGC"), pyopal.Alphabet("TCGA")
    assert a1 == a1 and a1 == a1.letters and a1 == a2 and a1 != a3 and a1 != 10
    again = pickle.loads(pickle.dumps(a1))
    assert again == a1 and again.letters == "ATGC"


@pytest.mark.parametrize("letters", ["AAAA", "AtgC", "A[]C"])
def test_alphabet_rejects(letters):
    with pytest.raises(ValueError):
        pyopal.Alphabet(letters)


def test_alphabet_encode_decode():
    dna = pyopal.Alphabet("ATGC")
    assert dna.encode("ATGC") == bytes([0, 1, 2, 3]) == dna.encode(b"ATGC")
    assert dna.encode("AAAAA") == bytes(5)
    for kind in (bytes, bytearray, lambda b: memoryview(bytearray(b))):
        assert dna.decode(kind([0, 1, 2, 3])) == "ATGC"
        assert dna.decode(kind([0, 0, 0, 0, 0])) == "AAAAA"
    # lib.pyx:304-306
    assert pyopal.Alphabet("ACGT").encode("GATACA") == b"\x02\x00\x03\x00\x01\x00"
    # unknown letters map to the '*' wildcard of the default alphabet; non-letters are rejected (lib.pyx:264-272)
    protein = pyopal.Alphabet()
    assert protein.encode("AJ") == bytes([0, protein.letters.index("*")])
    with pytest.raises(ValueError):
        protein.encode("A-C")
    with pytest.raises(ValueError):
        dna.encode("ATXG")


# --- Database (test_database.py) -------------------------------------------------


def test_database_container():
    seqs = ["ATGC", "ATTTAC", "TTACCG"]
    db = pyopal.Database(seqs)
    assert all(s in db for s in seqs) and "TAACCG" not in db and "AAAA" not in db
    with pytest.raises(TypeError):
        1 in db
    assert db.lengths == [4, 6, 6] and db.total_length == 16 and len(db) == 3
    assert [db[i] for i in range(3)] == seqs and [db[-1], db[-2], db[-3]] == seqs[::-1]
    assert list(pyopal.Database([s.encode("ascii") for s in seqs])) == seqs
    for bad in (3, -4, -8):
        with pytest.raises(IndexError):
            db[bad]


def test_database_slices_reverse_pickle():
    seqs = ["ATGC", "ATTC", "TTCG", "TTAT", "AAAC"]
    db = pyopal.Database(seqs)
    assert list(db[:2]) == seqs[:2]
    assert list(db[1:4:2]) == seqs[1:4:2]
    assert list(db[1::-1]) == seqs[1::-1]
    assert list(pickle.loads(pickle.dumps(db))) == seqs
    db.reverse()
    assert list(db) == seqs[::-1]
    empty = pyopal.Database()
    empty.reverse()
    assert len(empty) == 0


def test_database_mutation():
    db = pyopal.Database(["ATGC", "ATTC"])
    db.insert(1, "TTCC")
    db.insert(-10, "TTTT")
    db.insert(10, "AAAA")
    assert list(db) == ["TTTT", "ATGC", "TTCC", "ATTC", "AAAA"]
    db = pyopal.Database(["ATGC", "ATTC", "TTCG"])
    del db[1]
    del db[-2]
    assert list(db) == ["TTCG"]
    del db[0]
    assert list(db) == []
    for bad in (0, -1):
        with pytest.raises(IndexError):
            del db[bad]
    db = pyopal.Database(["ATGC", "ATTC", "TTCG"])
    db[2] = "AAAT"
    assert list(db) == ["ATGC", "ATTC", "AAAT"]
    for bad in (-8, 5):
        with pytest.raises(IndexError):
            db[bad] = "TCGA"
    db.extend(["CCCC", "GG"])
    db.append("T")
    assert db.lengths == [4, 4, 4, 4, 2, 1]
    db.clear()
    assert len(db) == 0 and db.total_length == 0


def test_database_mask_extract():
    seqs = ["ATGC", "ATTC", "TTCG", "TTAT"]
    db = pyopal.Database(seqs)
    assert list(db.mask([True, False, False, True])) == ["ATGC", "TTAT"]
    assert list(db.extract([2, 0])) == ["TTCG", "ATGC"]
    with pytest.raises(IndexError):
        db.mask([True])
    with pytest.raises(IndexError):
        db.extract([9])


# --- Results (test_result.py) ------------------------------------------------------


def test_score_result():
    r = pyopal.ScoreResult(10, score=30)
    assert (r.target_index, r.score) == (10, 30)
    assert repr(pyopal.ScoreResult(target_index=10, score=30)) == "ScoreResult(10, score=30)"
    assert pickle.loads(pickle.dumps(r)) == r
    assert r == pyopal.ScoreResult(10, 30) and r != pyopal.ScoreResult(12, 50) and r != 12


def test_end_result():
    r = pyopal.EndResult(2, score=30, query_end=10, target_end=20)
    assert (r.target_index, r.score, r.query_end, r.target_end) == (2, 30, 10, 20)
    assert repr(pyopal.EndResult(10, 30, 10, 20)) == "EndResult(10, score=30, query_end=10, target_end=20)"
    assert pickle.loads(pickle.dumps(r)) == r
    assert r != pyopal.EndResult(10, score=35, query_end=20, target_end=60) and r != 12


def test_full_result():
    args = dict(score=30, query_end=10, target_end=20, query_start=0, target_start=10, query_length=100,
                target_length=100, alignment="M" * 10)
    r = pyopal.FullResult(10, **args)
    assert (r.query_start, r.target_start, r.query_length, r.target_length) == (0, 10, 100, 100)
    assert r.alignment == "M" * 10 and r.cigar() == "10M" and r.identity() == 1.0
    again = pickle.loads(pickle.dumps(r))
    assert again == r and again.alignment == "M" * 10
    other = dict(args, score=48, target_start=30)
    assert r != pyopal.FullResult(2, **other) and r != 12
    with pytest.raises(TypeError):
        pyopal.FullResult(10, **dict(args, alignment=b"MMMM"))
    mixed = pyopal.FullResult(0, **dict(args, alignment="IMMMXMIM", query_start=0, target_start=0, query_end=5,
                                        target_end=7, query_length=6, target_length=8))
    assert mixed.cigar() == "1D5M1D1M"  # lib.pyx:1006-1010: Opal INS -> CIGAR D
    assert mixed.coverage("query") == 1.0 and mixed.coverage("target") == 7 / 8  # lib.pyx:1076-1082
    with pytest.raises(ValueError):
        mixed.coverage("nothing")


# --- Aligner (test_aligner.py:10-21, lib.pyx:1161-1238, 1342-1370) -------------------


def test_aligner_constructor():
    m = pyopal.ScoringMatrix.from_name("BLOSUM62")
    assert pyopal.Aligner("BLOSUM62").scoring_matrix == m
    assert pyopal.Aligner(m).scoring_matrix == m
    with pytest.raises(TypeError):
        pyopal.Aligner(1)
    default = pyopal.Aligner()
    assert (default.gap_open, default.gap_extend) == (3, 1)  # lib.pyx:1154-1155
    assert default.scoring_matrix == pyopal.ScoringMatrix.from_name("BLOSUM50")  # lib.pyx:1153
    assert default.alphabet == pyopal.Alphabet()
    assert pickle.loads(pickle.dumps(pyopal.Aligner("BLOSUM62", 11, 1))) == pyopal.Aligner("BLOSUM62", 11, 1)
    with pytest.raises(ValueError):
        pyopal.Aligner(pyopal.ScoringMatrix([[1.5, 0], [0, 1.5]], "AC"))  # not integer (lib.pyx:1230-1231)


def test_aligner_argument_validation():
    aligner = pyopal.Aligner()
    db = pyopal.Database(["AACCGCTG"])
    for kw in ({"mode": "nope"}, {"overflow": "nope"}, {"algorithm": "nope"}):
        with pytest.raises(ValueError):
            aligner.align("ACCTCG", db, **kw)
    with pytest.raises(ValueError):  # alphabet mismatch (lib.pyx:1357-1358)
        aligner.align("ACCTCG", pyopal.Database(["ACGT"], pyopal.Alphabet("ACGT")))
    with pytest.raises(IndexError):  # lib.pyx:1367-1368
        aligner.align("ACCTCG", db, start=1, end=0)
    with pytest.raises(ValueError):
        aligner.align("AC-T", db)


# --- the boundary -----------------------------------------------------------------------


def test_library_exports_every_declared_symbol():
    lib = cuda.load_library()
    header = open(os.path.join(ROOT, "include", "pyopal_cuda.h")).read()
    declared = set(re.findall(r"\b((?:opal|pyopal_cuda_|pyopal_host_)\w+)\s*\(", header))
    listed = set(open(os.path.join(ROOT, "include", "exports.txt")).read().split())
    assert declared == listed, declared ^ listed
    for name in sorted(declared):
        assert getattr(lib, name) is not None


def test_batch_units_pair_queries_by_length():
    # the order / pairing in which a batch is searched (csrc/api.cu batch_units): host code every rank repeats, so
    # it must be a pure function of the query lengths
    import array
    import ctypes

    lib = cuda.load_library()
    lib.pyopal_host_batch_units.restype = ctypes.c_int
    lib.pyopal_host_batch_units.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]

    def units(lengths, pair):
        ql = array.array("i", lengths)
        out = array.array("i", [-7] * (2 * max(len(ql), 1)))
        n = lib.pyopal_host_batch_units(ql.buffer_info()[0] if len(ql) else None, len(ql), int(pair), out.buffer_info()[0])
        return [(out[2 * i], out[2 * i + 1]) for i in range(n)]

    assert units([], True) == []
    assert units([300], True) == [(0, -1)]
    assert units([100, 300, 1000, 2000, 5000], False) == [(i, -1) for i in range(5)]      # caller's order, one by one
    # a dual pass costs the longer query twice over: only queries within 0.7 of each other are paired
    assert units([100, 300, 1000, 2000, 5000], True) == [(4, -1), (3, -1), (2, -1), (1, -1), (0, -1)]
    assert units([300, 290, 1000, 800, 100], True) == [(2, 3), (0, 1), (4, -1)]
    assert units([50, 0, 70, 0], True) == [(2, 0), (1, -1), (3, -1)]                       # empty queries stay single
    got = units([5, 9, 9, 1, 7, 3, 8], True)
    assert got == [(1, 2), (6, 4), (0, -1), (5, -1), (3, -1)]
    assert sorted(q for u in got for q in u if q >= 0) == list(range(7))                  # every query exactly once


def test_constants_match_header():
    header = open(os.path.join(ROOT, "include", "pyopal_cuda.h")).read()
    for name, value in re.findall(r"#define\s+((?:OPAL|PYOPAL_CUDA)_\w+)\s+(\d+)", header):
        assert getattr(cuda, name) == int(value), name
    # pinned by lib.pyx:991-996
    assert (cuda.OPAL_ALIGN_MATCH, cuda.OPAL_ALIGN_DEL, cuda.OPAL_ALIGN_INS, cuda.OPAL_ALIGN_MISMATCH) == (0, 1, 2, 3)


def test_result_helpers_without_device():
    lib = cuda.load_library()
    r = cuda.OpalSearchResult()
    lib.opalInitSearchResult(r)
    assert (r.scoreSet, r.endLocationQuery, r.startLocationTarget, r.alignmentLength) == (0, -1, -1, 0)
    assert lib.opalSearchResultIsEmpty(r) == 1
    lib.opalSearchResultSetScore(r, 42)
    assert lib.opalSearchResultIsEmpty(r) == 0 and r.score == 42


def test_no_cpu_fallback():
    """Without a CUDA device the search raises RuntimeError (pyx.in:102-103), it never computes on the CPU."""
    if cuda.device_count() > 0:
        pytest.skip("a CUDA device is visible")
    with pytest.raises(RuntimeError):
        pyopal.Aligner().align("ACCTCG", pyopal.Database(["AACCGCTG"]))
    import sys

    assert "oracle" not in sys.modules or True  # the product never imports the oracle:
    import subprocess

    out = subprocess.run(
        [sys.executable, "-c", "import sys, pyopal_b200, pyopal_b200.platform.cuda; print('oracle' in sys.modules)"],
        cwd=ROOT, capture_output=True, text=True, check=True,
    )
    assert out.stdout.strip() == "False"


def test_native_result_storage():
    # csrc/_results.c: the C base of the result classes and the bulk list constructor (pyx.in:65-72, 95-99)
    import array

    from pyopal_b200 import _results

    assert issubclass(pyopal.ScoreResult, _results.ResultBase)
    blank = pyopal.EndResult.__new__(pyopal.EndResult)
    assert (blank._target_index, blank._score, blank._query_end, blank._target_end) == (-1, None, -1, -1)  # lib.pyx:787-795
    assert not hasattr(blank, "__dict__")
    scores = array.array("i", [7, -3, 0])
    res = _results.make_results(pyopal.ScoreResult, 0, 10, 3, scores, None, None, None, None, 5, None, None, None, None)
    assert [(r.target_index, r.score) for r in res] == [(10, 7), (11, -3), (12, 0)]
    assert res[0] == pyopal.ScoreResult(10, 7) and repr(res[1]) == "ScoreResult(11, score=-3)"
    ops = bytearray(b"\x00\x00\x03\x01\x02\x00")
    full = _results.make_results(
        pyopal.FullResult, 2, 1, 2, array.array("i", [5, 6]), array.array("i", [1, 2]), array.array("i", [3, 4]),
        array.array("i", [0, 0]), array.array("i", [0, 1]), 7, array.array("i", [9, 8, 7]), ops,
        array.array("q", [0, 3]), array.array("i", [3, 3]))
    assert full[0] == pyopal.FullResult(1, 5, 1, 3, 0, 0, 7, 8, "MMX")
    assert full[1].alignment == "DIM" and full[1].target_length == 7 and full[1].query_length == 7
    with pytest.raises(ValueError):
        _results.make_results(pyopal.ScoreResult, 0, 0, 4, scores, None, None, None, None, 5, None, None, None, None)
    with pytest.raises(TypeError):
        _results.make_results(int, 0, 0, 3, scores, None, None, None, None, 5, None, None, None, None)


def test_database_bulk_ingest_and_persistence(tmp_path):
    # additive constructors: packed buffers, FASTA, dump / load
    import io

    seqs = ["ACCTCG", "AACCGCTG", "", "TTATTACG"]
    ref = pyopal.Database(seqs)
    packed = pyopal.Database.from_packed("".join(seqs).encode(), [len(s) for s in seqs], encoded=False)
    assert list(packed) == seqs and packed.lengths == ref.lengths
    enc = b"".join(ref.get_sequences())
    assert list(pyopal.Database.from_packed(enc, ref.lengths)) == seqs
    with pytest.raises(ValueError):
        pyopal.Database.from_packed(enc, [1, 2, 3])
    with pytest.raises(ValueError):
        pyopal.Database.from_packed(bytes([200]), [1])
    with pytest.raises(ValueError):
        pyopal.Database.from_packed(b"AC1T", [4], encoded=False)
    fasta = b">sp|P1 first\nACCT\nCG\n>second\r\nAACCGCTG\r\n>empty\n>last\nTTATTACG"
    db, names = pyopal.Database.from_fasta(io.BytesIO(fasta))
    assert list(db) == seqs and names == ["sp|P1 first", "second", "empty", "last"]
    path = tmp_path / "db.fa"
    path.write_bytes(fasta)
    assert list(pyopal.Database.from_fasta(str(path))[0]) == seqs
    dna = pyopal.Database(["ACGT", "GATTACA"], pyopal.Alphabet("ACGT"))
    dna.dump(str(tmp_path / "dna.bin"))
    back = pyopal.Database.load(str(tmp_path / "dna.bin"))
    assert list(back) == ["ACGT", "GATTACA"] and back.alphabet == dna.alphabet
    buf = io.BytesIO()
    ref.dump(buf)
    assert list(pyopal.Database.load(io.BytesIO(buf.getvalue()))) == seqs
    with pytest.raises(ValueError):
        pyopal.Database.load(io.BytesIO(b"nonsense"))


def test_cython_binding_loads_and_reports_errors():
    # platform/_cuda_cy.pyx: the Cython `with nogil` binding of pyopal_cuda_search (pyx.in:76-91 equivalent).
    # Without a GPU only the argument / error path can be exercised: a null handle must come back as a
    # non-zero return code with a message, never a crash.
    import array

    from pyopal_b200.platform import _cuda_cy, cuda

    assert cuda._cy is _cuda_cy
    assert _cuda_cy.device_count() == cuda.device_count()
    matrix = array.array("i", [1] * (24 * 24))
    scores = array.array("i", [0] * 4)
    rc = _cuda_cy.search(0, b"\x00\x01", 11, 1, matrix, 24, 0, 3, 0, 0, 4, scores)
    assert rc != 0 and _cuda_cy.last_error()
    with pytest.raises(RuntimeError):
        cuda.check(rc)
