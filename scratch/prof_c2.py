"""One search of config 2 per query length (after a warm-up each): the short command the ncu captures under
profiles/ are taken from."""
import argparse, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthetic
import pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda

ap = argparse.ArgumentParser()
ap.add_argument("lengths", nargs="*", type=int, default=[1000])
ap.add_argument("--end", action="store_true")
ap.add_argument("--shard", type=int, default=1, help="search rank 0's 1/N shard of the deal")
ap.add_argument("--config1", action="store_true")
ap.add_argument("--reps", type=int, default=2)
a = ap.parse_args()
if a.config1:
    q1, db = synthetic.config1()
    queries = {300: q1.tobytes()}
else:
    db = synthetic.config2()
    queries = {len(q): q.tobytes() for q in synthetic.config2_queries(db)}
if a.shard > 1:
    db = synthetic.pack_subset(db, np.asarray(sharding.shard_indices(db[2], a.shard, 0)))
M = pyopal.Aligner("BLOSUM62", 11, 1)._int_matrix
store = cuda.DeviceStore(db[0], db[1], np.ascontiguousarray(db[2]), 24, device=0)
res = int(db[2].sum())
for L in a.lengths:
    q = queries[L]
    for _ in range(a.reps):
        store.search(q, 11, 1, M, 1 if a.end else 0, 3, lengths=db[2])
    st = store.stats()
    print(f"Lq={L} end={a.end} dev {st['device_us']} us k16 {st['k16_us']} us k32 {st['k32_us']} us G={st['G']} K={st['K']} "
          f"passes={st['passes']} {L * res / max(st['k16_us'], 1) / 1e3:.1f} GCUPS(k16)", flush=True)
