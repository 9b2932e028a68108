#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_score.py -x -q -m gpu -k "giant or batch" > gpurun_out/test_gpu_score.log 2>&1; echo "score rc=$?"; tail -2 gpurun_out/test_gpu_score.log
timeout 600 python -m pytest tests/test_gpu_fuzz.py -x -q -m gpu -k "dual" > gpurun_out/test_gpu_fuzz.log 2>&1; echo "fuzz rc=$?"; tail -2 gpurun_out/test_gpu_fuzz.log
PYOPAL_CUDA_DEBUG=1 timeout 900 python bench.py --workload config5 --steps 1 --warmup 1 --no-cpu-baseline --blocks-per-gpu 2 > gpurun_out/c5_dual.json 2> gpurun_out/c5_dual.err; echo "c5 dual rc=$?"
grep "dual" gpurun_out/c5_dual.err | sort | uniq -c | sort -rn | head -40
for plan in "3,40" "3,28" "2,28" "3,36" "4,20"; do
PYOPAL_CUDA_PLAN=$plan timeout 900 python bench.py --workload config5 --steps 1 --warmup 1 --no-cpu-baseline --blocks-per-gpu 2 > gpurun_out/c5_dual_$plan.json 2> /dev/null
done
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/c5_dual*.json")):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d["value"], d["ms_per_step"], d["top_hits_checksum"])
    except Exception as e: print(f,"ERR",e)
PY
