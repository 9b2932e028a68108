"""One two-query batch (a dual search) on one block of config 5: the short command the ncu capture of dual_kernel
under profiles/ is taken from."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthetic
import pyopal_b200 as pyopal
from pyopal_b200.platform import cuda

sr, so, sl, _ = synthetic.config5_shard(0, 1, 1, 781_250)
M = pyopal.Aligner("BLOSUM62", 11, 1)._int_matrix
qs = sorted((q.tobytes() for q in synthetic.config5_queries(64)), key=len, reverse=True)
pair = [qs[20], qs[21]]
store = cuda.DeviceStore(sr, so, np.ascontiguousarray(sl), 24, device=0)
res = int(sl.sum())
for _ in range(2):
    store.search_batch(pair, 11, 1, M, 3)
    st = store.stats()
    print(f"Lq={len(pair[0])}+{len(pair[1])} k16 {st['k16_us']} us G={st['G']} K={st['K']} passes={st['passes']} dual={st['profile_in_smem'] == 2} "
          f"{(len(pair[0]) + len(pair[1])) * res / max(st['k16_us'], 1) / 1e3:.1f} GCUPS(k16)", flush=True)
