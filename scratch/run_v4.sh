mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_fullsize.py -x -q -k "sharded_database or config5" > gpurun_out/t_sdb.log 2>&1; echo "rc=$?" >> gpurun_out/t_sdb.log
N=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?" >> gpurun_out/bench_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --workload config5 --gpus $N --steps 1 --warmup 1 > gpurun_out/bench_c5_n$N.json 2> gpurun_out/bench_c5_n$N.err; echo "rc=$?" >> gpurun_out/bench_c5_n$N.err
