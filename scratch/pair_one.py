"""One end-mode search of the top-1000 hits of the 300-aa query (config 4), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200 import sharding
from pyopal_b200.platform import cuda
al = pyopal.Aligner("BLOSUM62", 11, 1)
M = al._int_matrix
db2 = synthetic.config2()
res2, off2, ln2 = db2
q = synthetic.config2_queries(db2)[1].tobytes()
store2 = cuda.DeviceStore(res2, off2, np.ascontiguousarray(ln2), 24, device=0)
sc = np.frombuffer(store2.search(q, 11, 1, M, 0, 3)[0], dtype=np.int32)
top = np.sort(np.argsort(-sc, kind="stable")[:1000])
sr, so, sl = sharding.pack_shard(res2, off2, ln2, top)
sub = cuda.DeviceStore(sr, so, sl, 24, device=0)
sub.search(q, 11, 1, M, 1, 3, lengths=sl)
print(sub.stats())
