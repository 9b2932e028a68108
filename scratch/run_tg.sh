mkdir -p gpurun_out
timeout 300 python scratch/shard_probe.py 8 > gpurun_out/shard8.log 2>&1
timeout 300 python scratch/shard_probe.py 1 > gpurun_out/shard1.log 2>&1
timeout 300 python scratch/shard_probe.py 2 > gpurun_out/shard2.log 2>&1
timeout 300 python scratch/shard_probe.py 4 > gpurun_out/shard4.log 2>&1
