import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import synthetic, pyopal_b200 as pyopal
from pyopal_b200.platform import cuda
import bench
db = synthetic.config2(570000)
queries = [q.tobytes() for q in synthetic.config2_queries(db)]
res, off, ln = db
store = cuda.DeviceStore(res, off, ln, 24, device=0)
al = pyopal.Aligner("BLOSUM62", 11, 1); M = al._int_matrix
outs = [torch.zeros(len(ln) + 8, dtype=torch.int32, device="cuda") for _ in queries]
stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
def step():
    for qi, q in enumerate(queries):
        store.search_scores_device(q, 11, 1, M, 3, outs[qi].data_ptr(), stream.cuda_stream)
def run(label, sampler_period=None, steps=3):
    step(); torch.cuda.synchronize()
    s = None
    if sampler_period:
        props = torch.cuda.get_device_properties(0)
        s = bench.ClockSampler("%08x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id), sampler_period)
        s.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    e0.record()
    for _ in range(steps): step()
    e1.record(); torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / steps * 1e3
    r = s.finish() if s else None
    print(label, "%.1f ms/step (events)  %.1f wall" % (e0.elapsed_time(e1) / steps, wall), r, flush=True)
run("no sampler")
run("nvml 0.2s", 0.2)
run("nvml 1.0s", 1.0)
run("no sampler again")
for qi, q in enumerate(queries):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): store.search_scores_device(q, 11, 1, M, 3, outs[qi].data_ptr(), stream.cuda_stream)
    e1.record(); torch.cuda.synchronize()
    print(len(q), "back-to-back %.2f ms" % (e0.elapsed_time(e1) / 3), flush=True)
for qi, q in enumerate(queries):
    for _ in range(2):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        store.search_scores_device(q, 11, 1, M, 3, outs[qi].data_ptr(), stream.cuda_stream)
        e1.record(); torch.cuda.synchronize()
        st = store.stats()
    print(len(q), "single %.2f ms" % (e0.elapsed_time(e1)), st["k32_us"], st["k16_us"], st["rescore_us"], flush=True)
os.environ["PYOPAL_CUDA_NO_TAIL"] = "1"
run("no tail split")
