#!/bin/bash
# final profile pass: -m gpu suite, launch list of the bench command, full captures of the two score kernels
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_config2.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
P="python scratch/prof_c2.py 1000"
$P > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 2 -c 1 -o gpurun_out/r2_bulk_lq1000_final $P > gpurun_out/ncu_prof.log 2>&1
echo "bulk capture rc=$?"; cat gpurun_out/plain_prof.log
P2="python scratch/prof_c2.py 1000 --shard 8"
$P2 > gpurun_out/plain_prof2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:stream_kernel -s 1 -c 1 -o gpurun_out/r2_stream_shard8_lq1000 $P2 > gpurun_out/ncu_prof2.log 2>&1
echo "stream capture rc=$?"; cat gpurun_out/plain_prof2.log
