"""Pin the oracle to every golden value the reference's tests hold for the path.

Sources: tests/test_aligner.py:42-79,93-131,
tests/test_align.py:9-37, doctests _align.py:106-111, lib.pyx:1006-1010,
lib.pyx:1076-1082 -- all under /root/reference/src/pyopal/.
"""

import json
import os

import pytest

import pyopal_b200 as pyopal

OPS = b"MDIX" + bytes(252)
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _enc(s):
    return pyopal.Alphabet().encode(s)


def _matrix(name):
    return pyopal.Aligner(name)._int_matrix


def test_nw_vector(oracle):
    r = oracle.search(_enc("ACCTCG"), [_enc("AACCGCTG")], _matrix("BLOSUM50"), 24, 3, 1, "nw", "full")
    assert r["score"] == [44]
    assert (r["query_end"], r["target_end"]) == ([5], [7])
    assert (r["query_start"], r["target_start"]) == ([0], [0])
    assert r["alignment"][0].translate(OPS) == b"IMMMXMIM"  # cigar 1D5M1D1M


def test_sw_vector(oracle):
    r = oracle.search(_enc("ACCTCG"), [_enc("AACCGCTG")], _matrix("BLOSUM50"), 24, 3, 1, "sw", "full")
    assert r["score"] == [47]
    assert (r["query_end"], r["target_end"]) == ([5], [7])
    assert (r["query_start"], r["target_start"]) == ([0], [1])
    assert r["alignment"][0].translate(OPS) == b"MMMXMIM"


def test_align_doctest_scores(oracle):
    targets = ["AACCGCTG", "ATGCGCT", "TTATTACG"]
    r = oracle.search(_enc("ACCTG"), [_enc(t) for t in targets], _matrix("BLOSUM50"), 24, 2, 1, "sw", "score")
    assert r["score"] == [41, 31, 23]


def test_test_align_vector(oracle):
    targets = ["AACCGCTG", "AACCGCTA", "AACCGCTC", "AACCGCTT"]
    for threads in (1, 2):
        r = oracle.search(_enc("ACCTCG"), [_enc(t) for t in targets], _matrix("BLOSUM50"), 24, 3, 1, "nw", "full",
                          threads=threads)
        assert r["score"][0] == 44 and r["query_end"][0] == 5 and r["target_end"][0] == 7
        assert r["query_start"][0] == 0 and r["target_start"][0] == 0


def test_hw_ov_selfconsistency(oracle):
    # not pinned by the reference: brute-force value of the survey
    for algo in ("hw", "ov"):
        r = oracle.search(_enc("ACCTCG"), [_enc("AACCGCTG")], _matrix("BLOSUM50"), 24, 3, 1, algo, "end")
        assert r["score"] == [47] and r["query_end"] == [5] and r["target_end"] == [7]


def test_mode_ordering(oracle):
    """sw >= ov >= hw >= nw for any pair (each relaxes the previous one's constraints)."""
    import numpy as np
    import synthetic

    rng = np.random.Generator(np.random.PCG64(11))
    m = _matrix("BLOSUM62")
    for _ in range(50):
        q = synthetic.random_residues(rng, rng.integers(1, 60)).tobytes()
        t = synthetic.random_residues(rng, rng.integers(1, 60)).tobytes()
        s = {a: oracle.search(q, [t], m, 24, 11, 1, a, "score")["score"][0] for a in ("nw", "hw", "ov", "sw")}
        assert s["sw"] >= s["ov"] >= s["hw"] >= s["nw"]


def test_full_mode_alignment_consistency(oracle):
    """The op string re-scores to the reported score and spans [start, end]."""
    import numpy as np
    import synthetic

    rng = np.random.Generator(np.random.PCG64(12))
    M = _matrix("BLOSUM62")
    go, ge = 11, 1
    for algo in ("nw", "hw", "ov", "sw"):
        for _ in range(40):
            q = synthetic.random_residues(rng, rng.integers(1, 80)).tobytes()
            t = synthetic.random_residues(rng, rng.integers(1, 80)).tobytes()
            r = oracle.search(q, [t], M, 24, go, ge, algo, "full")
            ops = r["alignment"][0]
            if not ops:
                assert algo == "sw" and r["score"][0] == 0
                continue
            i, j, score, prev = r["query_start"][0], r["target_start"][0], 0, None
            for op in ops:
                if op in (0, 3):
                    assert (q[i] == t[j]) == (op == 0)
                    score += M[q[i] * 24 + t[j]]
                    i += 1
                    j += 1
                elif op == 2:
                    score -= ge if prev == 2 else go
                    j += 1
                else:
                    score -= ge if prev == 1 else go
                    i += 1
                prev = op
            assert (i - 1, j - 1) == (r["query_end"][0], r["target_end"][0])
            assert score == r["score"][0], (algo, q, t)


FIXTURES = sorted(f for f in os.listdir(GOLDEN) if f.startswith("oracle_") and f.endswith(".json"))


@pytest.mark.parametrize("name", FIXTURES)
def test_committed_fixtures(oracle, name):
    """Regression fixtures under tests/golden/ (made by tests/golden/make_fixtures.py) still reproduce."""
    with open(os.path.join(GOLDEN, name)) as f:
        fx = json.load(f)
    q = bytes(fx["query"])
    targets = [bytes(t) for t in fx["targets"]]
    M = _matrix(fx["matrix"])
    for case in fx["cases"]:
        r = oracle.search(q, targets, M, 24, fx["gap_open"], fx["gap_extend"], case["algorithm"], "full")
        assert r["score"] == case["score"]
        assert r["query_end"] == case["query_end"] and r["target_end"] == case["target_end"]
        assert r["query_start"] == case["query_start"] and r["target_start"] == case["target_start"]
        assert [a.translate(OPS).decode() for a in r["alignment"]] == case["alignment"]


def test_reference_vectors_file(oracle):
    """tests/golden/reference_vectors.json: every value the reference's own tests pin."""
    with open(os.path.join(GOLDEN, "reference_vectors.json")) as f:
        ref = json.load(f)
    assert ref["alphabet"] == pyopal.Alphabet().letters
    assert len(ref["vectors"]) == 4
    for v in ref["vectors"]:
        r = oracle.search(_enc(v["query"]), [_enc(t) for t in v["targets"]], _matrix(v["matrix"]), 24,
                          v["gap_open"], v["gap_extend"], v["algorithm"], "full")
        n = 1 if v.get("first_only") else len(v["targets"])
        for key in ("score", "query_end", "target_end", "query_start", "target_start"):
            if key in v:
                assert r[key][:n] == v[key], (v["source"], key)
        if "alignment" in v:
            assert [a.translate(OPS).decode() for a in r["alignment"]][:n] == v["alignment"]


def test_simd_baseline_equals_scalar_oracle(oracle):
    # oracle/opal_simd.c (the CPU arm of bench.py) against the scalar restatement, overflow included
    import array

    import numpy as np

    import pyopal_b200 as pyopal
    import synthetic

    M = pyopal.Aligner("BLOSUM62")._int_matrix
    res, off, ln = synthetic.database(19, 700, 5.3, 0.7, 1, 3000)
    buf = res.tobytes() + b"\0"
    o, l = array.array("q", off.tolist()), array.array("i", ln.tolist())
    for lq, go, ge in ((1, 11, 1), (45, 11, 1), (300, 3, 1), (611, 5, 2)):
        q = synthetic.random_residues(np.random.Generator(np.random.PCG64(lq)), lq).tobytes()
        got = list(oracle.sw_scores_simd(q, buf, o, l, M, 24, go, ge, threads=4))
        assert got == oracle.search_flat(q, buf, o, l, M, 24, go, ge, "sw", "score", 4)["score"], (lq, go, ge)
    base = synthetic.random_residues(np.random.Generator(np.random.PCG64(1)), 9000)
    buf = base.tobytes() + base[:100].tobytes() + b"\0"
    o, l = array.array("q", [0, 9000]), array.array("i", [9000, 100])
    q = base[:8000].tobytes()
    got = list(oracle.sw_scores_simd(q, buf, o, l, M, 24, 11, 1, threads=2))
    assert got == oracle.search_flat(q, buf, o, l, M, 24, 11, 1, "sw", "score", 2)["score"] and got[0] > 32767
