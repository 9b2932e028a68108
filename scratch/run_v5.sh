mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "topk or sharded_database" > gpurun_out/t_topk.log 2>&1; echo "rc=$?" >> gpurun_out/t_topk.log
timeout 900 python bench.py --workload config5 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c5_n1.json 2> gpurun_out/bench_c5_n1.err; echo "rc=$?" >> gpurun_out/bench_c5_n1.err
