"""End / full mode timings: whole C1 database in end mode, top-1000 hits in end + full (config 4)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda
q300, db = synthetic.config1()
res, off, ln = db
store = cuda.DeviceStore(res, off, np.ascontiguousarray(ln), 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
M = al._int_matrix
cells = 300 * int(ln.sum())
for algo, name in ((3, "sw"), (0, "nw")):
    for mode, mname in ((0, "score"), (1, "end")):
        best = 1e9
        for _ in range(2):
            t = time.perf_counter(); out = store.search(q300.tobytes(), 11, 1, M, mode, algo, lengths=ln); dt = time.perf_counter() - t
            best = min(best, dt)
        st = store.stats()
        print(f"C1 {name} {mname}: wall {best*1e3:.1f} ms device {st['device_us']/1e3:.1f} ms -> {cells/best/1e9:.0f} GCUPS launches={st['launches']}", flush=True)
# config 4: top 1000 hits of the 300-aa and the 1000-aa query on the C2 database
db2 = synthetic.config2()
res2, off2, ln2 = db2
queries = [q.tobytes() for q in synthetic.config2_queries(db2)]
store2 = cuda.DeviceStore(res2, off2, np.ascontiguousarray(ln2), 24, device=0)
for qi in (1, 2, 4):
    q = queries[qi]
    sc = np.frombuffer(store2.search(q, 11, 1, M, 0, 3)[0], dtype=np.int32)
    top = np.sort(np.argsort(-sc, kind="stable")[:1000])
    sr, so, sl = sharding.pack_shard(res2, off2, ln2, top)
    t = time.perf_counter(); sub = cuda.DeviceStore(sr, so, sl, 24, device=0); tb = time.perf_counter() - t
    for mode, mname in ((1, "end"), (2, "full")):
        best = 1e9
        for _ in range(2):
            t = time.perf_counter(); out = sub.search(q, 11, 1, M, mode, 3, lengths=sl); dt = time.perf_counter() - t
            best = min(best, dt)
        st = sub.stats()
        print(f"C4 Lq={len(q)} top1000 ({int(sl.sum())} residues, longest {int(sl.max())}) {mname}: wall {best*1e3:.1f} ms device {st['device_us']/1e3:.1f} ms "
              f"-> {1000/best:.0f} alignments/s, store build {tb*1e3:.1f} ms", flush=True)
