mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/t_gpu.log
timeout 900 python bench.py --workload config5 --steps 1 --warmup 1 > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "rc=$?" >> gpurun_out/bench_c5_n1.err
timeout 300 python scratch/shard_probe.py 8 > gpurun_out/shard8.log 2>&1
nvidia-smi --query-gpu=memory.used,memory.total --format=csv > gpurun_out/mem.txt
