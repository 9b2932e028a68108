#!/bin/bash
# 2-GPU validation of the final code: NCCL tests, config 2 at N = 2, a small config 5 at N = 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/test_multi_gpu.log 2>&1; echo "test_multi_gpu rc=$?"; tail -3 gpurun_out/test_multi_gpu.log
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r2_bench_n2.json 2> gpurun_out/r2_bench_n2.err; echo "bench n2 rc=$?"
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --workload config5 --blocks-per-gpu 2 --steps 1 --warmup 1 --verify > gpurun_out/c5_n2_small.json 2> gpurun_out/c5_n2_small.err; echo "c5 n2 rc=$?"
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_n1_node2.json 2>/dev/null; echo "bench n1 rc=$?"
python - <<'PY'
import json
for f in ("r2_bench_n2","c5_n2_small","r2_bench_n1_node2"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json").read().strip().splitlines() if l.startswith('{"metric')][-1])
        print(f, round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), d.get("verified"), d.get("checksums"), d.get("top_hits_checksum"))
    except Exception as e: print(f,"ERR",e)
PY
