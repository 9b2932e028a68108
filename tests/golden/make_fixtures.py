#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/.

Two kinds of file:

* ``reference_vectors.json`` -- every known-answer value the reference's own
  tests and doctests hold for the search path, transcribed by hand with the
  file:line each comes from.  The reference itself cannot be
  imported in this environment (its Opal submodule and the `scoring_matrices`
  dependency are absent), so these are the only reference-produced outputs
  that exist; the oracle must reproduce all of them (tests/test_oracle_golden.py)
  and so must the GPU path (tests/test_gpu_endfull.py).

* ``oracle_*.json`` -- regression vectors produced by the oracle
  (oracle/opal_oracle.c) on seeded random inputs, all four algorithms in full
  mode.  They pin the oracle against accidental change and let the GPU parity
  tests run against committed data; they are NOT reference outputs.

Run from the repo root:  python tests/golden/make_fixtures.py
"""

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import oracle  # noqa: E402
import pyopal_b200 as pyopal  # noqa: E402
import synthetic  # noqa: E402

OPS = b"MDIX" + bytes(252)


def reference_vectors():
    src = "/root/reference/src/pyopal/"
    return {
        "alphabet": "ARNDCQEGHILKMFPSTWYVBZX*",
        "vectors": [
            {"source": src + "tests/test_aligner.py:42-79; lib.pyx:1006-1010,1076-1082", "matrix": "BLOSUM50",
             "gap_open": 3, "gap_extend": 1, "algorithm": "nw", "query": "ACCTCG", "targets": ["AACCGCTG"],
             "score": [44], "query_end": [5], "target_end": [7], "query_start": [0], "target_start": [0],
             "alignment": ["IMMMXMIM"], "cigar": ["1D5M1D1M"], "coverage_query": [1.0], "coverage_target": [0.875]},
            {"source": src + "tests/test_aligner.py:93-131", "matrix": "BLOSUM50", "gap_open": 3, "gap_extend": 1,
             "algorithm": "sw", "query": "ACCTCG", "targets": ["AACCGCTG"], "score": [47], "query_end": [5],
             "target_end": [7], "query_start": [0], "target_start": [1], "coverage_query": [1.0],
             "coverage_target": [0.875]},
            {"source": src + "tests/test_align.py:9-37 (first result only is pinned)", "matrix": "BLOSUM50",
             "gap_open": 3, "gap_extend": 1, "algorithm": "nw", "query": "ACCTCG",
             "targets": ["AACCGCTG", "AACCGCTA", "AACCGCTC", "AACCGCTT"], "first_only": True, "score": [44],
             "query_end": [5], "target_end": [7], "query_start": [0], "target_start": [0]},
            {"source": src + "_align.py:106-111", "matrix": "BLOSUM50", "gap_open": 2, "gap_extend": 1,
             "algorithm": "sw", "query": "ACCTG", "targets": ["AACCGCTG", "ATGCGCT", "TTATTACG"],
             "score": [41, 31, 23]},
        ],
    }


def oracle_fixture(name, seed, matrix, go, ge, lq, n, mu, sigma, hi, homologs=0, letters=None):
    rng = np.random.Generator(np.random.PCG64(seed))
    aligner = pyopal.Aligner(matrix, go, ge)

    def rand(k):
        if letters is None:
            return synthetic.random_residues(rng, k)
        return rng.choice(letters, size=int(k)).astype(np.uint8)

    q = rand(lq)
    lens = synthetic.random_lengths(rng, n, mu, sigma, 0, hi)
    targets = [rand(k).tobytes() for k in lens]
    for _ in range(homologs):
        a = int(rng.integers(0, max(1, lq // 3)))
        targets.append(synthetic.mutate(rng, q[a:], 0.15, 0.08).tobytes())
    cases = []
    for algo in ("nw", "hw", "ov", "sw"):
        r = oracle.search(q.tobytes(), targets, aligner._int_matrix, 24, go, ge, algo, "full", threads=8)
        r["alignment"] = [a.translate(OPS).decode() for a in r["alignment"]]
        r["algorithm"] = algo
        cases.append(r)
    fx = {"generator": "tests/golden/make_fixtures.py (oracle regression vectors, not reference outputs)",
          "matrix": matrix, "gap_open": go, "gap_extend": ge, "query": list(q.tobytes()),
          "targets": [list(t) for t in targets], "cases": cases}
    with open(os.path.join(HERE, name), "w") as f:
        json.dump(fx, f, separators=(",", ":"))


def main():
    oracle.build()
    with open(os.path.join(HERE, "reference_vectors.json"), "w") as f:
        json.dump(reference_vectors(), f, indent=1)
    oracle_fixture("oracle_blosum62_11_1.json", 1001, "BLOSUM62", 11, 1, 60, 24, 4.0, 0.7, 300, homologs=6)
    oracle_fixture("oracle_blosum50_3_1.json", 1002, "BLOSUM50", 3, 1, 45, 24, 3.8, 0.7, 250, homologs=6)
    oracle_fixture("oracle_ties_small_alphabet.json", 1003, "BLOSUM50", 3, 1, 30, 40, 3.0, 0.6, 80, homologs=4,
                   letters=[0, 4, 7, 16])


if __name__ == "__main__":
    main()
