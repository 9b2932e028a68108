import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
lq = int(sys.argv[1]) if len(sys.argv) > 1 else 100
rng = np.random.Generator(np.random.PCG64(3))
lengths = np.linspace(10_000, 35_213, 32).astype(np.int32)
res = synthetic.random_residues(rng, int(lengths.sum()))
off = np.zeros(len(lengths), dtype=np.int64); off[1:] = np.cumsum(lengths[:-1])
store = cuda.DeviceStore(res, off, lengths, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
q = synthetic.random_residues(np.random.Generator(np.random.PCG64(lq)), lq).tobytes()
for _ in range(2):
    store.search(q, 11, 1, al._int_matrix, 0, 3)
    print(store.stats())
