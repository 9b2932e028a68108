"""Sweep the bulk tile (logG, K, threads) on config 2 for a few query lengths."""
import os, sys, subprocess, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if len(sys.argv) > 1 and sys.argv[1] == "worker":
    import numpy as np
    import synthetic, pyopal_b200 as pyopal
    from pyopal_b200.platform import cuda
    db = synthetic.config2()
    res, off, ln = db
    queries = [q.tobytes() for q in synthetic.config2_queries(db)]
    store = cuda.DeviceStore(res, off, np.ascontiguousarray(ln), 24, device=0)
    al = pyopal.Aligner("BLOSUM62", 11, 1)
    plans = json.loads(sys.argv[2])
    for qi in (1, 2, 4):
        q = queries[qi]
        for plan in plans:
            if plan: os.environ["PYOPAL_CUDA_PLAN"] = plan
            else: os.environ.pop("PYOPAL_CUDA_PLAN", None)
            best = None
            try:
                for _ in range(2):
                    store.search(q, 11, 1, al._int_matrix, 0, 3)
                    st = store.stats()
                    best = st if best is None or st["device_us"] < best["device_us"] else best
                print(f"Lq={len(q)} plan={plan or 'auto':10s} G={best['G']} K={best['K']} passes={best['passes']} device_us={best['device_us']} "
                      f"k16_us={best['k16_us']} gcups={len(q) * int(ln.sum()) / best['device_us'] / 1e3:.0f}", flush=True)
            except Exception as e:
                print(f"Lq={len(q)} plan={plan} failed: {e}", flush=True)
else:
    plans = ["", "2,36,224", "2,36,256", "2,36,192", "2,32,256", "2,32,288", "2,40,224", "2,40,256", "3,36,224", "3,36,256", "3,40,224", "3,40,256"]
    subprocess.check_call([sys.executable, __file__, "worker", json.dumps(plans)])
