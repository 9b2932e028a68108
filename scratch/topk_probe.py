"""Device top-k (csrc/topk.cu) on config-5-sized score arrays: time per call against torch.topk, and GB/s."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import synthetic
from pyopal_b200.platform import cuda
res, off, ln = synthetic.database(1, 1000, 5.0, 0.5, 10, 500)
store = cuda.DeviceStore(res, off, ln, 24, device=0)
st = torch.cuda.current_stream().cuda_stream
for n in (570_000, 6_250_000, 50_000_000):
    g = torch.Generator(device="cuda").manual_seed(1)
    # SW-like scores: small, heavily tied
    scores = (torch.randn(n, generator=g, device="cuda") * 8 + 40).clamp_(0, 3000).to(torch.int32)
    for k in (1000,):
        vals = torch.empty(k, dtype=torch.int32, device="cuda"); idx = torch.empty(k, dtype=torch.int32, device="cuda")
        for fn, name in ((lambda: store.topk_device(scores.data_ptr(), n, k, vals.data_ptr(), idx.data_ptr(), st), "topk.cu"),
                         (lambda: torch.topk(scores, k), "torch.topk")):
            for _ in range(3): fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(); e0.record()
            for _ in range(10): fn()
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / 10
            print(f"n={n} k={k} {name}: {ms*1e3:.0f} us, {20.0*n/ms/1e6:.0f} GB/s at 20 B per target", flush=True)
        tv, ti = torch.topk(scores, k)
        assert sorted(vals.tolist(), reverse=True) == tv.tolist()
