#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in test_gpu_fuzz test_gpu_fullsize test_gpu_score; do
  timeout 900 python -m pytest tests/$f.py -x -q -m gpu > gpurun_out/$f.log 2>&1; echo "$f rc=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$f.log >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
{
echo "== whole default"; PYOPAL_CUDA_DEBUG=1 python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 3 2>&1 | grep -v "after " | uniq
echo "== 1/8 default"; PYOPAL_CUDA_DEBUG=1 python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 3 --shard 8 2>&1 | grep -v "after " | uniq
echo "== 1/8 NO_STREAM 512"; PYOPAL_CUDA_NO_STREAM=1 PYOPAL_CUDA_PLAN=5,32,512 python scratch/prof_c2.py 1000 2000 5000 --reps 3 --shard 8
echo "== 1/8 stream 5,32,512"; PYOPAL_CUDA_PLAN=5,32,512 python scratch/prof_c2.py 5000 --reps 3 --shard 8
echo "== 1/4 default"; PYOPAL_CUDA_DEBUG=1 python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 3 --shard 4 2>&1 | grep -v "after " | uniq
echo "== 1/4 STREAM forced"; PYOPAL_CUDA_STREAM=1 python scratch/prof_c2.py 1000 2000 5000 --reps 3 --shard 4
} > gpurun_out/stream_sweep.log 2>&1
cat gpurun_out/stream_sweep.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --workload config1 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"
