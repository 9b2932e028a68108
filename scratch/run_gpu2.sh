#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== whole, default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 1 2>&1
echo "== whole, NO_MID"; PYOPAL_CUDA_NO_MID=1 python scratch/tier_probe.py 1 2>&1
echo "== 1/8, default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 8 2>&1
echo "== 1/8, NO_MID"; PYOPAL_CUDA_NO_MID=1 python scratch/tier_probe.py 8 2>&1
echo "== 1/8, MID=0.25"; PYOPAL_CUDA_MID=0.25 python scratch/tier_probe.py 8 2>&1
echo "== 1/8, MID=1.0"; PYOPAL_CUDA_MID=1.0 python scratch/tier_probe.py 8 2>&1
echo "== whole END default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 1 1 2>&1
echo "== whole END, NO_END16"; PYOPAL_CUDA_NO_END16=1 python scratch/tier_probe.py 1 1 2>&1
} > gpurun_out/tier_probe.log
timeout 300 python -m pytest tests/test_gpu_score.py -x -q -m gpu > gpurun_out/test_gpu_score.log 2>&1; tail -3 gpurun_out/test_gpu_score.log
grep -E "TOTAL|BATCH|==" gpurun_out/tier_probe.log
