#!/bin/bash
# GPU pass: whole -m gpu suite, then every driver-runnable bench line (kept under profiles/ as r2_bench_*.json)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/summary.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/summary.txt
tail -3 gpurun_out/pytest_gpu.log >> gpurun_out/summary.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?" >> gpurun_out/summary.txt; }
run r2_bench_n1
run r2_bench_reference_arm --impl reference --steps 2 --warmup 1
run r2_bench_config2_end --mode end --no-cpu-baseline
run r2_bench_config1 --workload config1
run r2_bench_config1_end --workload config1 --mode end --no-cpu-baseline
run r2_bench_config3_nw --algorithm nw --no-cpu-baseline
run r2_bench_config3_hw --algorithm hw --no-cpu-baseline
run r2_bench_config3_ov --algorithm ov --no-cpu-baseline
run r2_bench_config4 --workload config4 --steps 3 --warmup 3
run r2_bench_config5_n1 --workload config5 --steps 1 --warmup 1 --verify --no-cpu-baseline
cat gpurun_out/summary.txt
