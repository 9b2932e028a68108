#!/bin/bash
# 8-GPU pass: NCCL tests, config 2 at N = 8 and N = 1 on the same node, config 5 at N = 8
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary8.txt
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/test_multi_gpu.log 2>&1; echo "test_multi_gpu rc=$?" >> gpurun_out/summary8.txt
tail -3 gpurun_out/test_multi_gpu.log >> gpurun_out/summary8.txt
for n in 8 4; do
NCCL_DEBUG=WARN timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/r2_bench_n$n.json 2> gpurun_out/r2_bench_n$n.err; echo "bench n$n rc=$?" >> gpurun_out/summary8.txt
done
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_n1_node8.json 2> gpurun_out/r2_bench_n1_node8.err; echo "bench n1 rc=$?" >> gpurun_out/summary8.txt
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29520 bench.py --gpus 8 --workload config5 --steps 1 --warmup 1 > gpurun_out/r2_bench_config5_n8.json 2> gpurun_out/r2_bench_config5_n8.err; echo "bench c5 n8 rc=$?" >> gpurun_out/summary8.txt
cat gpurun_out/summary8.txt
