#!/bin/bash
# 2-GPU pass: whole -m gpu suite (incl. the NCCL tests), bench at N=1 and N=2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary2.txt
nvidia-smi -L > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/summary2.txt
tail -4 gpurun_out/pytest_gpu.log >> gpurun_out/summary2.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?" >> gpurun_out/summary2.txt
NCCL_DEBUG=INFO timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?" >> gpurun_out/summary2.txt
cat gpurun_out/summary2.txt
