#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== whole default (U8 residue loads)"; python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 3
for plan in "2,28,640" "2,24,640" "2,20,640"; do
  echo "== PLAN=$plan"; PYOPAL_CUDA_PLAN=$plan python scratch/prof_c2.py 1000 2000 --reps 3
done
echo "== 1/8 default"; python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 3 --shard 8
for plan in "2,28,512" "3,16,512" "2,16,512" "3,28,512"; do
  echo "== 1/8 PLAN=$plan"; PYOPAL_CUDA_PLAN=$plan python scratch/prof_c2.py 100 300 --reps 3 --shard 8
done
echo "== c1 default"; python scratch/prof_c2.py 300 --reps 3 --config1
echo "== c1 end"; python scratch/prof_c2.py 300 --reps 3 --config1 --end
} > gpurun_out/sweep3.log 2>&1
cat gpurun_out/sweep3.log
