"""One whole-database end-mode search of config 1 (100k proteins, 300-aa query), for ncu."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
q300, db = synthetic.config1()
res, off, ln = db
store = cuda.DeviceStore(res, off, np.ascontiguousarray(ln), 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
store.search(q300.tobytes(), 11, 1, al._int_matrix, 1, 3, lengths=ln)
print(store.stats())
