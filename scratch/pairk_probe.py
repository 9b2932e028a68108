"""End-mode time of the top-1000 hits (config 4) as a function of the forced band height PYOPAL_CUDA_PAIR_K."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda
al = pyopal.Aligner("BLOSUM62", 11, 1)
M = al._int_matrix
db2 = synthetic.config2()
res2, off2, ln2 = db2
queries = [q.tobytes() for q in synthetic.config2_queries(db2)]
store2 = cuda.DeviceStore(res2, off2, np.ascontiguousarray(ln2), 24, device=0)
for qi in (1, 2, 4):
    q = queries[qi]
    sc = np.frombuffer(store2.search(q, 11, 1, M, 0, 3)[0], dtype=np.int32)
    top = np.sort(np.argsort(-sc, kind="stable")[:1000])
    sr, so, sl = sharding.pack_shard(res2, off2, ln2, top)
    sub = cuda.DeviceStore(sr, so, sl, 24, device=0)
    for k in ("", "4", "4x4", "8", "8x4", "12", "16", "32"):
        kk, _, cc = k.partition("x")
        if kk: os.environ["PYOPAL_CUDA_PAIR_K"] = kk; os.environ["PYOPAL_CUDA_PAIR_C"] = cc or "1"
        else: os.environ.pop("PYOPAL_CUDA_PAIR_K", None); os.environ.pop("PYOPAL_CUDA_PAIR_C", None)
        for mode in (1, 2):
            best = 1e9
            for _ in range(3):
                t = time.perf_counter(); sub.search(q, 11, 1, M, mode, 3, lengths=sl); best = min(best, time.perf_counter() - t)
            st = sub.stats()
            print(f"Lq={len(q)} forcedK={k or 'model'} mode={mode} K={st['K']} C={st['G']} bands={st['passes']} wall {best*1e3:.2f} ms", flush=True)
