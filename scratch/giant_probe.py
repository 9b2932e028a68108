"""Giants only: 32 targets of 10000..35213 aa; device time of one search per query length."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
rng = np.random.Generator(np.random.PCG64(3))
lengths = np.linspace(10_000, 35_213, 32).astype(np.int32)
res = synthetic.random_residues(rng, int(lengths.sum()))
off = np.zeros(len(lengths), dtype=np.int64); off[1:] = np.cumsum(lengths[:-1])
store = cuda.DeviceStore(res, off, lengths, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
for lq in (100, 300, 1000, 2000, 5000):
    q = synthetic.random_residues(np.random.Generator(np.random.PCG64(lq)), lq).tobytes()
    for algo in (3, 0):
        best = None
        for _ in range(3):
            store.search(q, 11, 1, al._int_matrix, 0, algo)
            st = store.stats()
            best = st if best is None or st["device_us"] < best["device_us"] else best
        steps = 35213
        print(f"Lq={lq} algo={algo} device_us={best['device_us']} side_us={best['k32_us']} k16_us={best['k16_us']} launches={best['launches']} "
              f"-> {best['device_us'] * 1e3 / steps:.0f} ns/column of the longest target", flush=True)
