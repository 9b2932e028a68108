mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "concurrent or config5_block or golden" > gpurun_out/t_new.log 2>&1; echo "rc=$?" >> gpurun_out/t_new.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tgiant -c 3 -o gpurun_out/full_tgiant python bench.py --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_tg.log 2>&1
