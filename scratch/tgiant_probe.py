"""Giants alone: 32 targets of 10000..35213 aa, per-query device time with the transposed giant kernel
(PYOPAL_CUDA_TGIANT=warps,K) and with the query-over-lanes giant kernel (PYOPAL_CUDA_NO_TGIANT=1)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
rng = np.random.Generator(np.random.PCG64(11))
lens = np.linspace(10_000, 35_213, 32).astype(np.int32)
res = synthetic.random_residues(rng, int(lens.sum()))
_, off, _ = synthetic.pack(lens, res)
store = cuda.DeviceStore(res, off, lens, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1)
qs = [synthetic.random_residues(rng, n).tobytes() for n in (100, 300, 1000, 2000, 5000)]
ref = {}
for cfg in ["old", "1,32", "2,32", "4,32", "4,16", "8,16", "6,32"]:
    os.environ.pop("PYOPAL_CUDA_NO_TGIANT", None); os.environ.pop("PYOPAL_CUDA_TGIANT", None)
    if cfg == "old": os.environ["PYOPAL_CUDA_NO_TGIANT"] = "1"
    else: os.environ["PYOPAL_CUDA_TGIANT"] = cfg
    out = []
    for q in qs:
        best = None
        for _ in range(3):
            sc = store.search(q, 11, 1, al._int_matrix, 0, 3)[0]
            st = store.stats()
            best = st["device_us"] if best is None else min(best, st["device_us"])
        key = len(q)
        if cfg == "old": ref[key] = list(sc)
        ok = list(sc) == ref[key]
        out.append(f"Lq={len(q)}: {best/1e3:.2f} ms{'' if ok else ' MISMATCH'}")
    print(cfg, st["targets32"], " | ".join(out), flush=True)
