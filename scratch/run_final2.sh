mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/t_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "rc=$?" >> gpurun_out/smoke.log
timeout 500 python scratch/endfull_probe.py > gpurun_out/endfull.log 2>&1
PYOPAL_CUDA_PAIR_K=16 timeout 500 python scratch/endfull_probe.py > gpurun_out/endfull_k16.log 2>&1
