#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
rm -f gpurun_out/summary.txt
for f in test_gpu_fuzz test_gpu_score test_gpu_fullsize; do
  timeout 900 python -m pytest tests/$f.py -x -q -m gpu > gpurun_out/$f.log 2>&1; echo "$f rc=$?" >> gpurun_out/summary.txt
  tail -3 gpurun_out/$f.log >> gpurun_out/summary.txt
done
cat gpurun_out/summary.txt
run() { name=$1; shift; timeout 900 python bench.py "$@" > gpurun_out/$name.json 2> gpurun_out/$name.err; echo "$name rc=$?"; }
run r2_bench_n1
run r2_bench_config1 --workload config1
python - <<'PY'
import json
for f in ("r2_bench_n1","r2_bench_config1"):
    try:
        d=json.loads([l for l in open(f"gpurun_out/{f}.json").read().strip().splitlines() if l.startswith("{")][-1])
        print(f, round(d["value"],1), round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), "frac", round(d["roofline"]["frac"],4), d.get("verified"), d.get("batch"), d["checksums"]["300"], (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e: print(f,"ERR",e)
PY
