import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
db = synthetic.config2(570000)
queries = synthetic.config2_queries(db)
res, off, ln = db
store = cuda.DeviceStore(res, off, ln, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1); M = al._int_matrix
out = torch.zeros(len(ln) + 8, dtype=torch.int32, device="cuda")
def timed(stream_obj, q, reps=3):
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream_obj):
            e0.record()
            store.search_scores_device(q, 11, 1, M, 3, out.data_ptr(), stream_obj.cuda_stream)
            e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts)
s_default = torch.cuda.default_stream()
s_nb = torch.cuda.Stream()
for q in queries:
    qb = q.tobytes()
    a = timed(s_default, qb); st = store.stats()
    b = timed(s_nb, qb)
    t0 = time.perf_counter(); store.search(qb, 11, 1, M, 0, 3); c = (time.perf_counter() - t0) * 1e3
    st2 = store.stats()
    print(len(qb), "default %.2f ms  nonblocking %.2f ms  host-api wall %.2f ms dev %.2f ms  k16 %.2f ms rescue %.2f esc %d G%d K%d p%d" % (a, b, c, st2["device_us"]/1e3, st["k16_us"]/1e3, st["rescore_us"]/1e3, st2["escalated"], st["G"], st["K"], st["passes"]), flush=True)
