#!/bin/bash
# GPU pass: smoke, parity suites, bench lines
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
nvidia-smi -L > gpurun_out/smi.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
for f in test_gpu_score test_gpu_endfull test_gpu_fuzz test_gpu_fullsize test_multi_gpu; do
  timeout 900 python -m pytest tests/$f.py -x -q -m gpu > gpurun_out/$f.log 2>&1; echo "$f rc=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$f.log >> gpurun_out/summary.txt
done
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/summary.txt
timeout 300 python bench.py --workload config1 > gpurun_out/bench_c1.json 2> gpurun_out/bench_c1.err; echo "bench c1 rc=$?" >> gpurun_out/summary.txt
timeout 300 python bench.py --workload config1 --mode end --no-cpu-baseline > gpurun_out/bench_c1_end.json 2> gpurun_out/bench_c1_end.err; echo "bench c1 end rc=$?" >> gpurun_out/summary.txt
timeout 600 python bench.py --mode end --no-cpu-baseline > gpurun_out/bench_c2_end.json 2> gpurun_out/bench_c2_end.err; echo "bench c2 end rc=$?" >> gpurun_out/summary.txt
cat gpurun_out/summary.txt
