#!/bin/bash
# profile pass: -m gpu suite, bench, launch list of the bench command, full capture of the bulk kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r2_bench_n1.json 2> gpurun_out/r2_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --workload config5 --steps 1 --warmup 1 --no-cpu-baseline --blocks-per-gpu 2 > gpurun_out/c5_small.json 2> gpurun_out/c5_small.err; echo "c5 rc=$?"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-verify"
$B > gpurun_out/plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_bench_config2.csv $B > gpurun_out/ncu_bench.log 2>&1
echo "launch list rc=$?"
P="python scratch/prof_c2.py 1000"
$P > gpurun_out/plain_prof.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:search_kernel -s 2 -c 1 -o gpurun_out/r2_bulk_lq1000_final $P > gpurun_out/ncu_prof.log 2>&1
echo "bulk capture rc=$?"; cat gpurun_out/plain_prof.log
python - <<'PY'
import json
for f in ("r2_bench_n1","c5_small"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json").read().strip().splitlines() if l.startswith("{")][-1])
        print(f, round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "frac", round(d["roofline"]["frac"],4), d.get("verified"), d.get("top_hits_checksum"))
        for k in d.get("kernels",[]): print("   ",k)
    except Exception as e: print(f,"ERR",e)
PY
