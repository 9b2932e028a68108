"""Small end-to-end check (all algorithms, score/end/full, giants + bulk) for compute-sanitizer."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import oracle, synthetic, pyopal_b200 as pyopal
oracle.build()
rng = np.random.Generator(np.random.PCG64(5))
seqs = synthetic.as_sequences(synthetic.database(11, 300, 5.0, 0.6, 1, 1500, giants=(5000, 4000, 3000)))
al = pyopal.Aligner("BLOSUM62", 11, 1)
db = pyopal.Database.__new__(pyopal.Database)
db.alphabet = al.alphabet; db._sequences = seqs; db._lengths = [len(s) for s in seqs]
for lq in (50, 700):
    q = synthetic.random_residues(rng, lq).tobytes()
    for algo in ("sw", "nw", "hw", "ov"):
        got = list(al.align_scores(al.alphabet.decode(q), db, algorithm=algo))
        want = oracle.search(q, seqs, al._int_matrix, 24, 11, 1, algo, "score", threads=8)["score"]
        assert got == want, (lq, algo)
    print("score ok", lq, db._device_store().stats()["targets32"], flush=True)
small = pyopal.Database.__new__(pyopal.Database)
small.alphabet = al.alphabet; small._sequences = seqs[:40]; small._lengths = [len(s) for s in seqs[:40]]
q = synthetic.random_residues(rng, 120).tobytes()
for algo in ("sw", "nw"):
    res = al.align(al.alphabet.decode(q), small, mode="full", algorithm=algo)
    want = oracle.search(q, seqs[:40], al._int_matrix, 24, 11, 1, algo, "full", threads=8)
    assert [r.score for r in res] == want["score"]
    assert [bytes(r._alignment) for r in res] == want["alignment"]
print("full ok")
