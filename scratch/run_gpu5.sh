#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{
echo "== default all"; python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 2
echo "== default all END"; python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 2 --end
echo "== 1/4"; PYOPAL_CUDA_DEBUG=1 python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 2 --shard 4 2>&1 | grep -v "after " | uniq
echo "== 1/8"; python scratch/prof_c2.py 100 300 1000 2000 5000 --reps 2 --shard 8
echo "== c1"; python scratch/prof_c2.py 300 --reps 3 --config1
} > gpurun_out/cliff.log 2>&1
cat gpurun_out/cliff.log
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
timeout 300 python bench.py --workload config1 --no-cpu-baseline > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?"
