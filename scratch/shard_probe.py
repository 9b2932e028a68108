"""One rank's shard of config 2 at world size W on one GPU: per-query device times."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda
W = int(sys.argv[1]) if len(sys.argv) > 1 else 8
db = synthetic.config2()
res, off, ln = db
queries = [q.tobytes() for q in synthetic.config2_queries(db)]
mine = sharding.shard_indices(ln, W, 0) if W > 1 else np.arange(len(ln))
sr, so, sl = sharding.pack_shard(res, off, ln, mine) if W > 1 else db
store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
tot = 0.0
for q in queries:
    best = None
    for _ in range(3):
        store.search(q, 11, 1, al._int_matrix, 0, 3)
        st = store.stats()
        best = st if best is None or st["device_us"] < best["device_us"] else best
    tot += best["device_us"]
    print(f"W={W} Lq={len(q)} device_us={best['device_us']} k16_us={best['k16_us']} side_us={best['k32_us']} targets32={best['targets32']} "
          f"gcups={len(q) * int(sl.sum()) / best['device_us'] / 1e3:.0f}", flush=True)
cells = sum(len(q) for q in queries) * int(ln.sum())
print(f"W={W} step {tot/1e3:.2f} ms -> whole-job {cells / tot / 1e3:.0f} GCUPS if every rank matches rank 0", flush=True)
