#!/bin/bash
# GPU pass: whole -m gpu suite, then the bench lines that changed (kept under profiles/ as r2_bench_*.json)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/pytest_gpu.log >> gpurun_out/summary.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?" >> gpurun_out/summary.txt; }
run r2_bench_n1
run r2_bench_config1 --workload config1
run r2_bench_config5_n1 --workload config5 --steps 1 --warmup 1 --verify --no-cpu-baseline
cat gpurun_out/summary.txt
