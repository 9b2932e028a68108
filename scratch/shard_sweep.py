"""Tile sweep on rank 0's shard of config 2 at world size W (PYOPAL_CUDA_PLAN override)."""
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
mine = sharding.shard_indices(ln, W, 0)
sr, so, sl = sharding.pack_shard(res, off, ln, mine)
store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
for qi in (1, 2, 3, 4):
    q = queries[qi]
    ideal = len(q) * int(sl.sum()) / 5.08e6  # us at the single-GPU kernel rate
    for plan in ["", "2,36", "3,32", "3,36", "3,40", "4,32", "4,36", "4,40", "5,32", "5,36", "5,40"]:
        if plan: os.environ["PYOPAL_CUDA_PLAN"] = plan
        else: os.environ.pop("PYOPAL_CUDA_PLAN", None)
        best = None
        for _ in range(2):
            store.search(q, 11, 1, al._int_matrix, 0, 3)
            st = store.stats()
            best = st if best is None or st["device_us"] < best["device_us"] else best
        print(f"W={W} Lq={len(q)} plan={plan or 'auto':6s} G={best['G']} K={best['K']} passes={best['passes']} device_us={best['device_us']} "
              f"k16_us={best['k16_us']} side_us={best['k32_us']} t32={best['targets32']} ideal_us={ideal:.0f}", flush=True)
