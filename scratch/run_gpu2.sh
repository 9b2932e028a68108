#!/bin/bash
# 2-GPU validation: NCCL tests (one process per GPU, one process driving both)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multi_gpu.py -x -q -m gpu > gpurun_out/test_multi_gpu.log 2>&1; echo "test_multi_gpu rc=$?"; tail -15 gpurun_out/test_multi_gpu.log
