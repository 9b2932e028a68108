"""One C1-shaped search (300-aa query x 100k proteins) repeated: the ncu target."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
lq = int(sys.argv[1]) if len(sys.argv) > 1 else 300
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
q, db = synthetic.config1()
if lq != 300:
    import numpy as np
    q = synthetic.random_residues(np.random.Generator(np.random.PCG64(2)), lq)
res, off, ln = db
store = cuda.DeviceStore(res, off, ln, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
for _ in range(reps):
    store.search(q.tobytes(), 11, 1, al._int_matrix, 0, 3)
    print(store.stats())
