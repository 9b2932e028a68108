#!/bin/bash
# first GPU pass of round 2: smoke, quick parity subsets (stop at first failure per file), short bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
for f in test_gpu_score test_gpu_endfull test_gpu_fuzz test_gpu_fullsize test_multi_gpu; do
  timeout 900 python -m pytest tests/$f.py -x -q -m gpu > gpurun_out/$f.log 2>&1; echo "$f rc=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$f.log >> gpurun_out/summary.txt
done
timeout 600 python bench.py --steps 3 --warmup 2 --cpu-seconds 4 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
