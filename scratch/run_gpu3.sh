#!/bin/bash
# ncu pass: launch list of the bench command, full capture of the bulk kernel (1000-aa query), END kernel (config 1)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_config2.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
P="python scratch/prof_c2.py 1000"
$P > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 2 -c 1 -o gpurun_out/r2_bulk_lq1000 $P > gpurun_out/ncu_prof.log 2>&1
echo "full capture rc=$?"; cat gpurun_out/plain_prof.log
P2="python scratch/prof_c2.py 300 --config1 --end"
$P2 > gpurun_out/plain_prof2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 1 -c 1 -o gpurun_out/r2_end_c1_lq300 $P2 > gpurun_out/ncu_prof2.log 2>&1
echo "end capture rc=$?"; cat gpurun_out/plain_prof2.log
ls -la gpurun_out/*.ncu-rep
