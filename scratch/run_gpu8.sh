#!/bin/bash
# 8-GPU pass: NCCL tests, config 2 at N = 8 and N = 1 on the same node (add config 5 with: --workload config5)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary8.txt
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/test_multi_gpu.log 2>&1; echo "test_multi_gpu rc=$?" >> gpurun_out/summary8.txt
tail -3 gpurun_out/test_multi_gpu.log >> gpurun_out/summary8.txt
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > gpurun_out/r2_bench_n8.json 2> gpurun_out/r2_bench_n8.err; echo "bench n8 rc=$?" >> gpurun_out/summary8.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-verify > gpurun_out/r2_bench_n1_node8.json 2> gpurun_out/r2_bench_n1_node8.err; echo "bench n1 rc=$?" >> gpurun_out/summary8.txt
cat gpurun_out/summary8.txt
python - <<'PY'
import json
for f in ("r2_bench_n8","r2_bench_n1_node8"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json").read().strip().splitlines() if l.startswith('{"metric')][-1])
        print(f, round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), d.get("verified"), d["checksums"]["5000"])
        for k in d["kernels"]: print("   ", k["Lq"], k["G"], k["K"], k["passes"], k["k16_us"], k["k32_us"])
    except Exception as e: print(f,"ERR",e)
PY
