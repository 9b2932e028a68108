"""GPU parity: end and full (start + traceback) modes through the C-ABI vs the oracle (bit-exact).

Reference behaviour under test: searchType OPAL_SEARCH_SCORE_END / OPAL_SEARCH_ALIGNMENT of
opalSearchDatabase (/root/reference/src/pyopal/opal.pxd:40-54), read back through EndResult /
FullResult (lib.pyx:848-1119); golden values from tests/test_aligner.py:42-79,93-131,
tests/test_align.py:9-37 and the doctests lib.pyx:1006-1010,1076-1082.
"""

import numpy as np
import pytest

import pyopal_b200 as pyopal
import synthetic

from test_gpu_score import ALGOS, make_db, query_str, rand_query, rand_seqs

pytestmark = pytest.mark.gpu


def check_mode(oracle, aligner, query, seqs, algo, mode, db=None, **kw):
    db = db or make_db(seqs, aligner.alphabet)
    got = aligner.align(query_str(aligner, query), db, algorithm=algo, mode=mode, **kw)
    s, e = kw.get("start", 0), min(kw.get("end", len(seqs)), len(seqs))
    want = oracle.search(query, seqs[s:e], aligner._int_matrix, len(aligner.alphabet), aligner.gap_open,
                         aligner.gap_extend, algo, mode, threads=8)
    assert len(got) == len(want["score"])
    for k, r in enumerate(got):
        ctx = (algo, mode, k, len(query), len(seqs[s + k]))
        assert r.target_index == s + k
        assert r.score == want["score"][k], ctx
        assert (r._query_end, r._target_end) == (want["query_end"][k], want["target_end"][k]), ctx
        if mode == "full":
            assert (r._query_start, r._target_start) == (want["query_start"][k], want["target_start"][k]), ctx
            assert bytes(r._alignment or b"") == want["alignment"][k], ctx
            assert r.query_length == len(query) and r.target_length == len(seqs[s + k])
    return db


def test_golden_vectors(cuda):
    # test_aligner.py:42-79 (nw) and :93-131 (sw); lib.pyx:1006-1010 (cigar), :1076-1082 (coverage)
    aligner = pyopal.Aligner()
    db = pyopal.Database(["AACCGCTG"])
    r = aligner.align("ACCTCG", db, algorithm="nw", mode="end")[0]
    assert isinstance(r, pyopal.EndResult) and (r.score, r.query_end, r.target_end) == (44, 5, 7)
    r = aligner.align("ACCTCG", db, algorithm="nw", mode="full")[0]
    assert isinstance(r, pyopal.FullResult)
    assert (r.score, r.query_start, r.target_start, r.query_end, r.target_end) == (44, 0, 0, 5, 7)
    assert r.cigar() == "1D5M1D1M" and r.alignment == "IMMMXMIM"
    assert r.coverage("query") == 1.0 and r.coverage("target") == 7 / 8
    r = aligner.align("ACCTCG", db, algorithm="sw", mode="end")[0]
    assert (r.score, r.query_end, r.target_end) == (47, 5, 7)
    r = aligner.align("ACCTCG", db, algorithm="sw", mode="full")[0]
    assert (r.score, r.query_start, r.target_start, r.query_end, r.target_end) == (47, 0, 1, 5, 7)
    assert r.alignment == "MMMXMIM"
    assert r.coverage("query") == 1.0 and r.coverage("target") == 7 / 8


def test_align_threads_golden(cuda):
    # test_align.py:9-37
    targets = ["AACCGCTG", "AACCGCTA", "AACCGCTC", "AACCGCTT"]
    for threads in (1, 2):
        res = list(pyopal.align("ACCTCG", targets, mode="full", algorithm="nw", ordered=True, threads=threads))
        assert [r.target_index for r in res] == [0, 1, 2, 3]
        r = res[0]
        assert (r.score, r.query_start, r.target_start, r.query_end, r.target_end) == (44, 0, 0, 5, 7)


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("mode", ("end", "full"))
@pytest.mark.parametrize("lq", [1, 9, 100, 300])
def test_random_db(cuda, oracle, algo, mode, lq):
    seqs = rand_seqs(300 + lq, 400, 4.8, 0.7, 1, 2500)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    check_mode(oracle, aligner, rand_query(lq + 1, lq), seqs, algo, mode)


@pytest.mark.parametrize("algo", ALGOS)
def test_homologs_with_gaps(cuda, oracle, algo):
    # mutated copies (10 % substitutions, 2 % indels): real alignments with gap runs and traceback ties
    rng = np.random.Generator(np.random.PCG64(401))
    base = synthetic.random_residues(rng, 700)
    seqs = [synthetic.mutate(rng, base, 0.15, 0.06).tobytes() for _ in range(40)]
    seqs += [base[100:450].tobytes(), base.tobytes(), synthetic.mutate(rng, base[200:600], 0.3, 0.1).tobytes()]
    for aligner in (pyopal.Aligner("BLOSUM62", 11, 1), pyopal.Aligner()):
        check_mode(oracle, aligner, base[50:650].tobytes(), seqs, algo, "full")


@pytest.mark.parametrize("algo", ALGOS)
def test_default_scoring_many_ties(cuda, oracle, algo):
    # BLOSUM50 / 3 / 1 (lib.pyx:1153-1155) on a 4-letter alphabet subset: many co-optimal paths
    rng = np.random.Generator(np.random.PCG64(411))
    seqs = [rng.choice([0, 4, 7, 16], size=int(n)).astype(np.uint8).tobytes() for n in rng.integers(1, 120, size=300)]
    aligner = pyopal.Aligner()
    for lq in (5, 40, 90):
        q = rng.choice([0, 4, 7, 16], size=lq).astype(np.uint8).tobytes()
        check_mode(oracle, aligner, q, seqs, algo, "full")


@pytest.mark.parametrize("algo", ALGOS)
def test_multi_pass_long_query(cuda, oracle, algo):
    # 1100 and 2300 rows: 2 and 3 passes of the widest pair tile (32 lanes x 32 rows)
    rng = np.random.Generator(np.random.PCG64(421))
    base = synthetic.random_residues(rng, 2300)
    seqs = rand_seqs(422, 30, 5.5, 0.8, 1, 3000) + [synthetic.mutate(rng, base).tobytes(), base[500:1800].tobytes()]
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    for lq in (1100, 2300):
        check_mode(oracle, aligner, base[:lq].tobytes(), seqs, algo, "full")


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("mode", ("end", "full"))
def test_empty_and_tiny_targets(cuda, oracle, algo, mode):
    seqs = [b"", b"\x00", b"\x01\x02", b"", rand_query(1, 50), b"\x05" * 3, b""] + rand_seqs(431, 40, 2.0, 0.8, 1, 30)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    for lq in (1, 5, 64):
        check_mode(oracle, aligner, rand_query(430 + lq, lq), seqs, algo, mode)


def test_sw_zero_score_has_no_alignment(cuda, oracle):
    # 'W' against 'P' only: every substitution is negative -> score 0, no end, no alignment
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    seqs = [bytes([14]) * 20, bytes([14]) * 3]
    db = check_mode(oracle, aligner, bytes([17]) * 10, seqs, "sw", "full")
    r = aligner.align("W" * 10, db, mode="full")[0]
    assert r.score == 0 and r._query_end == -1 and r._alignment in (None, b"") and r.cigar() is None


def test_slices_and_scores_agree(cuda, oracle):
    seqs = rand_seqs(441, 301, 5.0, 0.6, 1, 1200)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    q = rand_query(442, 120)
    db = make_db(seqs, aligner.alphabet)
    for s, e in ((0, 10), (3, 4), (150, 301), (100, 10_000)):
        check_mode(oracle, aligner, q, seqs, "sw", "full", db=db, start=s, end=e)
        check_mode(oracle, aligner, q, seqs, "hw", "end", db=db, start=s, end=e)
    for algo in ALGOS:
        score = [r.score for r in aligner.align(query_str(aligner, q), db, algorithm=algo)]
        end = [r.score for r in aligner.align(query_str(aligner, q), db, algorithm=algo, mode="end")]
        assert score == end


def test_generic_entry_point_full(cuda, oracle):
    """opalSearchDatabase's own signature in ALIGNMENT mode (malloc()ed alignment, opal.pxd:26-54)."""

    class ListDatabase(pyopal.BaseDatabase):
        def __init__(self, seqs, alphabet=None):
            super().__init__(alphabet=alphabet)
            self._seqs = list(seqs)

        def get_sequences(self):
            return self._seqs

        def get_lengths(self):
            return [len(s) for s in self._seqs]

        def get_size(self):
            return len(self._seqs)

    seqs = rand_seqs(451, 40, 4.5, 0.6, 1, 500)
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    q = rand_query(452, 70)
    db = ListDatabase(seqs, aligner.alphabet)
    for algo in ALGOS:
        res = aligner.align(query_str(aligner, q), db, algorithm=algo, mode="full")
        want = oracle.search(q, seqs, aligner._int_matrix, 24, 11, 1, algo, "full")
        assert [r.score for r in res] == want["score"]
        assert [r._query_start for r in res] == want["query_start"]
        assert [bytes(r._alignment or b"") for r in res] == want["alignment"]


def test_top_hits_workflow_config4(cuda, oracle):
    """C4: score pass -> top hits -> extract -> end + full on the hits (docs/examples/abc.ipynb cells 23-24)."""
    residues, offsets, lengths = synthetic.database(461, 3000, 5.3, 0.6, 20, 2500)
    seqs = synthetic.as_sequences((residues, offsets, lengths))
    rng = np.random.Generator(np.random.PCG64(462))
    q = synthetic.mutate(rng, np.frombuffer(seqs[77], dtype=np.uint8)).tobytes()
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = make_db(seqs, aligner.alphabet)
    scores = list(aligner.align_scores(query_str(aligner, q), db))
    top = sorted(range(len(scores)), key=lambda i: (-scores[i], i))[:100]
    assert top[0] == 77
    hits = db.extract(top)
    hit_seqs = [seqs[i] for i in top]
    check_mode(oracle, aligner, q, hit_seqs, "sw", "end", db=hits)
    check_mode(oracle, aligner, q, hit_seqs, "sw", "full", db=hits)


GOLDEN = __import__("os").path.join(__import__("os").path.dirname(__file__), "golden")


@pytest.mark.parametrize("name", sorted(f for f in __import__("os").listdir(GOLDEN) if f.startswith("oracle_")))
def test_committed_fixtures_on_gpu(cuda, name):
    """The committed fixtures (tests/golden/, no oracle call) through the C-ABI."""
    import json

    with open(__import__("os").path.join(GOLDEN, name)) as f:
        fx = json.load(f)
    aligner = pyopal.Aligner(fx["matrix"], fx["gap_open"], fx["gap_extend"])
    q = bytes(fx["query"])
    db = make_db([bytes(t) for t in fx["targets"]], aligner.alphabet)
    for case in fx["cases"]:
        res = aligner.align(query_str(aligner, q), db, algorithm=case["algorithm"], mode="full")
        assert [r.score for r in res] == case["score"]
        assert [r._query_end for r in res] == case["query_end"] and [r._target_end for r in res] == case["target_end"]
        assert [r._query_start for r in res] == case["query_start"]
        assert [r._target_start for r in res] == case["target_start"]
        assert [r.alignment for r in res] == case["alignment"]
        score_only = aligner.align(query_str(aligner, q), db, algorithm=case["algorithm"])
        assert [r.score for r in score_only] == case["score"]


def test_reference_vectors_on_gpu(cuda):
    """tests/golden/reference_vectors.json: the reference's own pinned values through the public API."""
    import json

    with open(__import__("os").path.join(GOLDEN, "reference_vectors.json")) as f:
        ref = json.load(f)
    for v in ref["vectors"]:
        aligner = pyopal.Aligner(v["matrix"], v["gap_open"], v["gap_extend"])
        res = aligner.align(v["query"], pyopal.Database(v["targets"]), algorithm=v["algorithm"], mode="full")
        n = 1 if v.get("first_only") else len(v["targets"])
        got = {
            "score": [r.score for r in res], "query_end": [r._query_end for r in res],
            "target_end": [r._target_end for r in res], "query_start": [r._query_start for r in res],
            "target_start": [r._target_start for r in res], "alignment": [r.alignment for r in res],
            "cigar": [r.cigar() for r in res], "coverage_query": [r.coverage("query") for r in res],
            "coverage_target": [r.coverage("target") for r in res],
        }
        for key, want in v.items():
            if key in got:
                assert got[key][:n] == want, (v["source"], key)


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("pair_k", ["4", "10", "12", "22", "32", "4x4", "8x4", "8x1"])
def test_band_teams_match_oracle(cuda, oracle, monkeypatch, algo, pair_k):
    # pairdp.cu walks a long query as bands of 32 K rows on several warps at once (pass p one chunk of columns
    # behind pass p - 1, results merged by the band that finishes last); the band height is normally picked by a
    # cost model, here it is forced: 700 rows = 6 bands of K = 4, 3 of K = 10,
    # 2 of K = 12, 1 of K = 22 or 32.  Homologs give the
    # reverse pass something to cut short, and equal scores in different bands exercise the merge order.
    # "KxC": C columns per step (the latency-bound shape takes four)
    k, _, c = pair_k.partition("x")
    monkeypatch.setenv("PYOPAL_CUDA_PAIR_K", k)
    monkeypatch.setenv("PYOPAL_CUDA_PAIR_C", c or "1")
    rng = np.random.Generator(np.random.PCG64(977 + int(k) + 100 * int(c or 1)))
    qarr = synthetic.random_residues(rng, 700)
    query = qarr.tobytes()
    seqs = rand_seqs(978, 60, 5.5, 0.8, 1, 3000)
    for k in range(0, 60, 5):
        a = int(rng.integers(0, 500))
        piece = synthetic.mutate(rng, qarr[a : a + int(rng.integers(60, 200))])
        t = np.frombuffer(seqs[k], dtype=np.uint8)
        pos = int(rng.integers(0, len(t) + 1))
        seqs[k] = np.concatenate([t[:pos], piece, t[pos:]]).astype(np.uint8).tobytes()
    seqs[7] = seqs[3]  # duplicates: identical candidates
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    db = None
    for mode in ("end", "full"):
        db = check_mode(oracle, aligner, query, seqs, algo, mode, db=db)


def test_band_teams_repeats_tie_across_bands(cuda, oracle, monkeypatch):
    # a periodic query against a periodic target: the same maximum in every band and many columns, so the first one
    # in (column, row) order must win whichever band finishes last
    monkeypatch.setenv("PYOPAL_CUDA_PAIR_K", "4")
    monkeypatch.setenv("PYOPAL_CUDA_PAIR_C", "4")
    unit = np.array([0, 4, 7, 10, 13, 16, 19, 3], dtype=np.uint8)
    query = np.tile(unit, 64).tobytes()  # 512 rows = 4 bands
    seqs = [np.tile(unit, r).tobytes() for r in (1, 2, 5, 40, 64, 100)] + [np.tile(unit[::-1], 30).tobytes()]
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    for algo in ALGOS:
        db = None
        for mode in ("end", "full"):
            db = check_mode(oracle, aligner, query, seqs, algo, mode, db=db)
