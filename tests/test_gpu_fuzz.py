"""GPU parity fuzz: random scoring matrices, gap costs, alphabets, lengths and modes vs the oracle (bit-exact).

The fixed-shape tests pin the paths the benchmark takes; this one walks the corners between them: asymmetric
matrices (the matrix is indexed [query letter][target letter], pyx.in:86 -- the transposed giant kernel must not
swap them), gap_extend > gap_open and zero gap costs (32-bit rung, long F chain), alphabets of 2..26 letters, scores
at the edge of the int8 profile, queries of 1..700 residues against targets of 0..9000 with planted homologs.
"""

import os

import numpy as np
import pytest

import pyopal_b200 as pyopal

from test_gpu_score import ALGOS, make_db
from test_gpu_endfull import check_mode

pytestmark = pytest.mark.gpu

# more cases on demand: PYOPAL_FUZZ_N=600 python -m pytest tests/test_gpu_fuzz.py
N_SCORE = int(os.environ.get("PYOPAL_FUZZ_N", "60"))
N_FULL = max(24, N_SCORE * 2 // 5)



def _case(seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    A = int(rng.choice([2, 4, 5, 20, 24, 27, 32]))
    letters = "".join(chr(ord("A") + i) for i in range(min(A, 26))) + "".join("abcdef"[i] for i in range(max(0, A - 26)))
    # Alphabet only accepts uppercase ASCII letters: cap at 26
    A = min(A, 26)
    letters = letters[:A]
    style = int(rng.integers(0, 4))
    if style == 0:    # BLOSUM-like symmetric, small
        lo, hi, diag = -4, 3, (4, 11)
    elif style == 1:  # asymmetric
        lo, hi, diag = -6, 5, (2, 9)
    elif style == 2:  # near the int8 edge once gap_open is folded in
        lo, hi, diag = -90, 40, (60, 100)
    else:             # beyond int8: wide-profile kernel
        lo, hi, diag = -300, 200, (250, 400)
    M = rng.integers(lo, hi + 1, size=(A, A))
    if style == 0:
        M = np.minimum(M, M.T)
    M[np.arange(A), np.arange(A)] = rng.integers(diag[0], diag[1] + 1, size=A)
    gaps = [(11, 1), (3, 1), (1, 1), (0, 0), (2, 5), (5, 0), (0, 3), (20, 2), (40, 7)]
    go, ge = gaps[int(rng.integers(0, len(gaps)))]
    if style == 3:
        go, ge = go * 9, ge * 9
    lq = int(rng.choice([1, 2, 17, 64, 130, 300, 700]))
    n = int(rng.choice([1, 7, 60, 300]))
    lens = np.clip(np.rint(rng.lognormal(4.5, 1.0, size=n)), 0, 3000).astype(int)
    if n >= 60:
        lens[int(rng.integers(0, n))] = int(rng.choice([1500, 4200, 9000]))  # a "giant" for small stores
        lens[int(rng.integers(0, n))] = 0
    seqs = [rng.integers(0, A, size=int(l), dtype=np.uint8) for l in lens]
    query = rng.integers(0, A, size=lq, dtype=np.uint8)
    if n >= 7 and lq >= 17:  # plant the query (mutated) into a target so long diagonals exist
        k = int(rng.integers(0, n))
        noisy = query.copy()
        flip = rng.random(lq) < 0.15
        noisy[flip] = rng.integers(0, A, size=int(flip.sum()), dtype=np.uint8)
        seqs[k] = np.concatenate([seqs[k][: len(seqs[k]) // 2], noisy, seqs[k][len(seqs[k]) // 2:]])
    return letters, M.tolist(), go, ge, query.tobytes(), [s.tobytes() for s in seqs]


@pytest.mark.parametrize("seed", range(N_SCORE))
def test_fuzz_score(cuda, oracle, seed):
    letters, rows, go, ge, query, seqs = _case(1000 + seed)
    alphabet = pyopal.Alphabet(letters)
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, letters), go, ge)
    db = make_db(seqs, alphabet)
    flat = [int(v) for row in rows for v in row]
    for algo in ALGOS:
        got = list(aligner.align_scores(alphabet.decode(query), db, algorithm=algo))
        want = oracle.search(query, seqs, flat, len(letters), go, ge, algo, "score", threads=8)["score"]
        bad = [i for i, (a, b) in enumerate(zip(got, want)) if a != b]
        assert not bad, (seed, algo, go, ge, len(letters), len(query),
                         [(i, got[i], want[i], len(seqs[i])) for i in bad[:4]])


@pytest.mark.parametrize("seed", range(N_FULL))
def test_fuzz_end_full(cuda, oracle, seed):
    letters, rows, go, ge, query, seqs = _case(5000 + seed)
    seqs = [s[:1200] for s in seqs[:40]]  # the oracle keeps full matrices for the traceback
    query = query[:300]
    alphabet = pyopal.Alphabet(letters)
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, letters), go, ge)
    flat = [int(v) for row in rows for v in row]
    for algo in ALGOS:
        for mode in ("end", "full"):
            if mode == "full" and ge > go and algo != "nw":
                # gap_extend > gap_open: re-opening a gap beats extending it inside the matrix but the borders are
                # literal, so the anchored reverse pass may not reproduce the forward score.  Both
                # the oracle (code 98) and the library (PYOPAL_CUDA_ERR_INTERNAL) refuse such a pair instead of
                # returning an inconsistent alignment; they must refuse the same inputs.
                try:
                    oracle.search(query, seqs, flat, len(letters), go, ge, algo, mode, threads=8)
                    oracle_ok = True
                except RuntimeError:
                    oracle_ok = False
                if not oracle_ok:
                    with pytest.raises(RuntimeError):
                        aligner.align(alphabet.decode(query), make_db(seqs, alphabet), algorithm=algo, mode=mode)
                    continue
            check_mode(oracle, aligner, query, seqs, algo, mode, db=make_db(seqs, alphabet))


@pytest.mark.parametrize("seed", range(max(30, N_SCORE // 2)))
def test_fuzz_forced_tiles(cuda, oracle, monkeypatch, seed):
    # the same cases under a FORCED bulk tile (PYOPAL_CUDA_PLAN=logG,K): narrow tiles turn a 300-row query into
    # dozens of passes (boundary rows through the L2 buffer, padding rows above row 0), wide ones leave most lanes
    # of a group on padding; slices and the transposed-giant shapes are drawn as well
    rng = np.random.Generator(np.random.PCG64(9000 + seed))
    letters, rows, go, ge, query, seqs = _case(2000 + seed)
    monkeypatch.setenv("PYOPAL_CUDA_PLAN", f"{int(rng.integers(2, 6))},{4 * int(rng.integers(1, 11))}")
    if rng.random() < 0.5:
        monkeypatch.setenv("PYOPAL_CUDA_TGIANT", f"{int(rng.integers(1, 9))},{int(rng.choice([16, 32]))},{int(rng.choice([0, 700, 2000]))}")
    alphabet = pyopal.Alphabet(letters)
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, letters), go, ge)
    db = make_db(seqs, alphabet)
    flat = [int(v) for row in rows for v in row]
    lo = int(rng.integers(0, len(seqs)))
    hi = int(rng.integers(lo, len(seqs) + 1))
    for algo in ALGOS:
        want = oracle.search(query, seqs, flat, len(letters), go, ge, algo, "score", threads=8)["score"]
        got = list(aligner.align_scores(alphabet.decode(query), db, algorithm=algo))
        assert got == want, (seed, algo, go, ge, os.environ["PYOPAL_CUDA_PLAN"])
        part = list(aligner.align_scores(alphabet.decode(query), db, algorithm=algo, start=lo, end=hi))
        assert part == want[lo:hi], (seed, algo, lo, hi)
        k = int(rng.integers(1, len(seqs) + 2))
        hits = aligner.align_topk(alphabet.decode(query), db, k, algorithm=algo)
        order = sorted(range(len(want)), key=lambda i: (-want[i], i))[:k]
        assert [(h.target_index, h.score) for h in hits] == [(i, want[i]) for i in order], (seed, algo, k)


def test_sw_overflow_threshold_is_exact(cuda, oracle):
    # scores spread densely around the 16 -> 32 bit escalation threshold (65535 - 129 - bias): pairs where the low
    # target, the high target, both or neither overflow must all come back exact
    import synthetic

    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    W, C = synthetic.ALPHABET.index("W"), synthetic.ALPHABET.index("C")  # 11 and 9 on the diagonal
    rng = np.random.Generator(np.random.PCG64(77))
    query = np.full(6300, W, dtype=np.uint8)
    query[rng.choice(6300, 930, replace=False)] = C
    seqs = []
    for n in range(5950, 6300, 3):  # identical prefixes: score = sum of the diagonal, 63.7k .. 67.4k in steps of ~30
        seqs.append(query[:n].tobytes())
    for n in rng.integers(6040, 6140, size=40):  # and prefixes with a few mismatches: every residue class
        t = query[: int(n)].copy()
        t[rng.choice(int(n), int(rng.integers(1, 6)), replace=False)] = synthetic.ALPHABET.index("A")
        seqs.append(t.tobytes())
    db = make_db(seqs, aligner.alphabet)
    got = list(aligner.align_scores(aligner.alphabet.decode(query.tobytes()), db))
    want = oracle.search(query.tobytes(), seqs, aligner._int_matrix, 24, 11, 1, "sw", "score", threads=8)["score"]
    assert got == want
    assert min(want) < 64500 and max(want) > 65800
    st = db._device_store().stats()
    assert st["escalated"] > 0 or st["targets32"] > 0


@pytest.mark.parametrize("algo", ("nw", "hw", "ov"))
@pytest.mark.parametrize("gaps", [(11, 4), (3, 2), (20, 8)])
def test_static_16bit_bound_edge(cuda, oracle, algo, gaps):
    # NW / HW / OV: targets whose length straddles the static bound up to which the 16-bit rung is exact
    # (|score| ~ open + len * ext close to 32768 - slack); both sides of the cut must be exact
    import synthetic

    go, ge = gaps
    aligner = pyopal.Aligner("BLOSUM62", go, ge)
    rng = np.random.Generator(np.random.PCG64(88))
    cut = (32768 - ge - 512 - 3 * go) // ge
    lens = list(range(cut - 420, cut + 60, 23)) + [cut - 130 + d for d in range(-3, 4)] + [200, 5, 900]
    seqs = [synthetic.random_residues(rng, n).tobytes() for n in lens]
    for lq in (100, 260):
        q = synthetic.random_residues(rng, lq).tobytes()
        got = list(aligner.align_scores(aligner.alphabet.decode(q), make_db(seqs, aligner.alphabet), algorithm=algo))
        want = oracle.search(q, seqs, aligner._int_matrix, 24, go, ge, algo, "score", threads=8)["score"]
        assert got == want, (algo, gaps, lq, [(i, a, b, lens[i]) for i, (a, b) in enumerate(zip(got, want)) if a != b][:5])


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("mode", ("score", "end", "full"))
def test_degenerate_inputs(cuda, oracle, algo, mode):
    # empty query, empty / one-residue targets, a database of only empty targets: closed forms on the host side
    # of the library (api.cu degenerate_score, pairdp.cu) against the oracle's m == 0 || n == 0 branches
    aligner = pyopal.Aligner("BLOSUM62", 11, 1)
    rng = np.random.Generator(np.random.PCG64(123))
    some = [b"", b"\x03", b"", rng.integers(0, 20, 40, dtype=np.uint8).tobytes(), b"\x00\x00", b""]
    for seqs in (some, [b"", b""], [b"\x05"]):
        for query in (b"", b"\x07", rng.integers(0, 20, 33, dtype=np.uint8).tobytes()):
            if mode == "score":
                got = list(aligner.align_scores(aligner.alphabet.decode(query), make_db(seqs, aligner.alphabet), algorithm=algo))
                want = oracle.search(query, seqs, aligner._int_matrix, 24, 11, 1, algo, "score", threads=2)["score"]
                assert got == want, (algo, len(query), [len(s) for s in seqs])
            else:
                check_mode(oracle, aligner, query, seqs, algo, mode)


@pytest.mark.parametrize("seed", range(6))
def test_fuzz_end_full_long(cuda, oracle, seed):
    # end / full on longer pairs: several passes of the pair kernel (more than 32 x K query rows), long gaps,
    # direction matrices of a few MB per pair
    rng = np.random.Generator(np.random.PCG64(7000 + seed))
    aligner = pyopal.Aligner(*[("BLOSUM62", 11, 1), ("BLOSUM50", 3, 1), ("BLOSUM62", 5, 5)][seed % 3])
    lq = int(rng.choice([900, 1700, 2600]))
    query = rng.integers(0, 20, lq, dtype=np.uint8)
    seqs = []
    for _ in range(6):
        n = int(rng.integers(300, 2400))
        t = rng.integers(0, 20, n, dtype=np.uint8)
        if rng.random() < 0.7:  # a homologous stretch with an indel
            a = int(rng.integers(0, lq // 2))
            piece = np.delete(query[a: a + lq // 3], slice(40, 40 + int(rng.integers(1, 30))))
            b = int(rng.integers(0, max(1, n - len(piece))))
            t = np.concatenate([t[:b], piece, t[b:]])
        seqs.append(t.tobytes())
    for algo in ALGOS:
        for mode in ("end", "full"):
            check_mode(oracle, aligner, query.tobytes(), seqs, algo, mode)


@pytest.mark.parametrize("seed", range(max(24, N_SCORE // 3)))
def test_fuzz_forced_stream(cuda, oracle, monkeypatch, seed):
    # the continuous-wavefront kernel (csrc/stream.cuh) forced for every whole-database score search
    # (PYOPAL_CUDA_STREAM=1; by default it only takes wide, deep tiles) under random tiles: several passes with the
    # boundary rows handed from a group's last lane to its first, segments shorter than a group, more groups than
    # pairs, every algorithm, matrices with negative folded scores (the packed-adder path)
    rng = np.random.Generator(np.random.PCG64(7000 + seed))
    letters, rows, go, ge, query, seqs = _case(3000 + seed)
    monkeypatch.setenv("PYOPAL_CUDA_STREAM", "1")
    monkeypatch.setenv("PYOPAL_CUDA_PLAN", f"{int(rng.integers(2, 6))},{4 * int(rng.integers(1, 11))}")
    alphabet = pyopal.Alphabet(letters)
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, letters), go, ge)
    db = make_db(seqs, alphabet)
    flat = [int(v) for row in rows for v in row]
    for algo in ALGOS:
        want = oracle.search(query, seqs, flat, len(letters), go, ge, algo, "score", threads=8)["score"]
        got = list(aligner.align_scores(alphabet.decode(query), db, algorithm=algo))
        bad = [i for i, (a, b) in enumerate(zip(got, want)) if a != b]
        assert not bad, (seed, algo, go, ge, os.environ["PYOPAL_CUDA_PLAN"], [(i, got[i], want[i], len(seqs[i])) for i in bad[:4]])


@pytest.mark.parametrize("seed", range(max(24, N_SCORE // 3)))
def test_fuzz_dual_query_batches(cuda, oracle, monkeypatch, seed):
    # batched Smith-Waterman searches: queries are paired by length and each pair goes down as ONE pass of
    # dual_kernel (csrc/dual.cuh: two queries in the halves of a register against one target, pre-packed profile),
    # forced here on small databases (PYOPAL_CUDA_DUAL=1; by default only databases without giants take it) under
    # random tiles -- queries of different length in one pair (the shorter one padded above row 0), several passes,
    # queries left single, matrices with negative folded scores, giants of either query beside the dual launch,
    # top-k of every query
    rng = np.random.Generator(np.random.PCG64(11000 + seed))
    letters, rows, go, ge, query, seqs = _case(4000 + seed)
    if ge > go:
        ge = go   # the 16-bit rung needs ext <= open; the single-query path covers the other case
    monkeypatch.setenv("PYOPAL_CUDA_DUAL", "1")
    if rng.random() < 0.7:
        monkeypatch.setenv("PYOPAL_CUDA_PLAN", f"{int(rng.integers(2, 6))},{4 * int(rng.integers(1, 11))}")
    alphabet = pyopal.Alphabet(letters)
    aligner = pyopal.Aligner(pyopal.ScoringMatrix(rows, letters), go, ge)
    db = make_db(seqs, alphabet)
    flat = [int(v) for row in rows for v in row]
    A = len(letters)
    nq = int(rng.integers(2, 6))
    queries = [query]
    for _ in range(nq - 1):
        if rng.random() < 0.7:   # within 0.7 of a query already drawn: the two are paired (csrc/api.cu batch_units)
            lq = max(1, int(len(queries[int(rng.integers(0, len(queries)))]) * rng.uniform(0.72, 1.0)))
        else:
            lq = int(rng.choice([1, 7, 60, 300, 700, 1400]))
        q = rng.integers(0, A, size=lq, dtype=np.uint8)
        if rng.random() < 0.5 and len(seqs):   # a homolog of some target: real scores, possibly an overflow
            src = np.frombuffer(seqs[int(rng.integers(0, len(seqs)))], dtype=np.uint8)
            if len(src):
                q = np.resize(src, lq).copy()
        queries.append(q.tobytes())
    got = aligner.align_batch([alphabet.decode(q) for q in queries], db)
    want = [oracle.search(q, seqs, flat, A, go, ge, "sw", "score", threads=8)["score"] for q in queries]
    for qi, (g, w) in enumerate(zip(got, want)):
        bad = [i for i, (x, y) in enumerate(zip(g, w)) if x != y]
        assert not bad, (seed, qi, len(queries[qi]), go, ge, os.environ.get("PYOPAL_CUDA_PLAN"), [(i, g[i], w[i], len(seqs[i])) for i in bad[:4]])
    k = int(rng.integers(1, len(seqs) + 1))
    tops = aligner.align_batch_topk([alphabet.decode(q) for q in queries], db, k)
    for hits, w in zip(tops, want):
        order = sorted(range(len(w)), key=lambda i: (-w[i], i))[:k]
        assert [(h.target_index, h.score) for h in hits] == [(i, w[i]) for i in order]
