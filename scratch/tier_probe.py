"""Per-query timing of the score search on config 2 (whole database or a 1/N shard) under the current
PYOPAL_CUDA_* switches: wall clock of the blocking call, k16 (bulk) / k32 (side streams) event times."""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthetic
import pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda

shard_of = int(sys.argv[1]) if len(sys.argv) > 1 else 1
mode = int(sys.argv[2]) if len(sys.argv) > 2 else 0   # 0 score, 1 end
db = synthetic.config2()
queries = [q.tobytes() for q in synthetic.config2_queries(db)]
if shard_of > 1:
    mine = np.asarray(sharding.shard_indices(db[2], shard_of, 0))
    db = synthetic.pack_subset(db, mine)
res = int(db[2].sum())
M = pyopal.Aligner("BLOSUM62", 11, 1)._int_matrix
store = cuda.DeviceStore(db[0], db[1], np.ascontiguousarray(db[2]), 24, device=0)
tot = 0.0
for q in queries:
    store.search(q, 11, 1, M, mode, 3, lengths=db[2])
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); store.search(q, 11, 1, M, mode, 3, lengths=db[2]); dt = time.perf_counter() - t0
        best = min(best, dt)
    st = store.stats()
    tot += best
    print(f"Lq={len(q):5d} wall {best*1e3:8.3f} ms  dev {st['device_us']/1e3:8.3f}  k16 {st['k16_us']/1e3:8.3f}  k32/side {st['k32_us']/1e3:8.3f}  G={st['G']} K={st['K']} passes={st['passes']}  {len(q)*res/best/1e9:8.1f} GCUPS", flush=True)
print(f"TOTAL {tot*1e3:.2f} ms  {sum(len(q) for q in queries)*res/tot/1e9:.1f} GCUPS  env={ {k:v for k,v in os.environ.items() if k.startswith('PYOPAL_CUDA')} }", flush=True)
if mode == 0:
    t0 = time.perf_counter(); store.search_batch(queries, 11, 1, M, 3); store.search_batch(queries, 11, 1, M, 3); dt = (time.perf_counter() - t0) / 2
    print(f"BATCH {dt*1e3:.2f} ms  {sum(len(q) for q in queries)*res/dt/1e9:.1f} GCUPS", flush=True)
