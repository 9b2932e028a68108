#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
{
echo "== whole, default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 1 2>&1
echo "== 1/8, default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 8 2>&1
echo "== 1/8, LANES=2"; PYOPAL_CUDA_LANES=2 python scratch/tier_probe.py 8 2>&1
echo "== whole END default"; PYOPAL_CUDA_DEBUG=1 python scratch/tier_probe.py 1 1 2>&1
} > gpurun_out/tier_probe.log
grep -E "TOTAL|BATCH|==" gpurun_out/tier_probe.log
for f in test_gpu_score test_gpu_endfull test_gpu_fuzz test_gpu_fullsize; do
  timeout 900 python -m pytest tests/$f.py -x -q -m gpu > gpurun_out/$f.log 2>&1; echo "$f rc=$?" >> gpurun_out/summary.txt
  tail -2 gpurun_out/$f.log >> gpurun_out/summary.txt
done
timeout 600 python bench.py --steps 5 --warmup 3 --cpu-seconds 4 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --workload config1 --steps 20 --warmup 5 --cpu-seconds 3 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --workload config1 --mode end --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_c1_end.json 2> gpurun_out/bench_c1_end.err; echo "bench c1 end rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
