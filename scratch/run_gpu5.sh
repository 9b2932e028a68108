#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
P="python scratch/prof_c2.py 5000 --shard 8 --reps 2"
PYOPAL_CUDA_STREAM=1 $P > gpurun_out/plain_s.log 2>&1 && PYOPAL_CUDA_STREAM=1 ncu --set full --clock-control none -k regex:stream_kernel -s 1 -c 1 -o gpurun_out/r2_stream_shard8_lq5000 $P > gpurun_out/ncu_s.log 2>&1; echo "ncu stream rc=$?"; cat gpurun_out/plain_s.log
$P > gpurun_out/plain_c.log 2>&1 && ncu --set full --clock-control none -k regex:search_kernel -s 2 -c 1 -o gpurun_out/r2_classic_shard8_lq5000 $P > gpurun_out/ncu_c.log 2>&1; echo "ncu classic rc=$?"; cat gpurun_out/plain_c.log
