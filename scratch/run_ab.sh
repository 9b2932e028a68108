mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/t_gpu.log
timeout 300 python scratch/shard_probe.py 1 > gpurun_out/ab_new.log 2>&1
timeout 300 python scratch/_old/scratch/shard_probe.py 1 > gpurun_out/ab_old.log 2>&1
timeout 300 python scratch/shard_probe.py 8 > gpurun_out/ab_new8.log 2>&1
timeout 300 python scratch/_old/scratch/shard_probe.py 8 > gpurun_out/ab_old8.log 2>&1
