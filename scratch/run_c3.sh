mkdir -p gpurun_out
for a in nw hw ov; do
  timeout 300 python bench.py --algorithm $a --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c3_$a.json 2> gpurun_out/bench_c3_$a.err
done
timeout 600 python bench.py > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "rc=$?" >> gpurun_out/bench_n1.err
