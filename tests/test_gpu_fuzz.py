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
