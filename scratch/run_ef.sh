mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_endfull.py tests/test_gpu_fuzz.py -x -q > gpurun_out/t_ef.log 2>&1; echo "rc=$?" >> gpurun_out/t_ef.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -x -q -k "config4 or whole_database_end" > gpurun_out/t_c4.log 2>&1; echo "rc=$?" >> gpurun_out/t_c4.log
timeout 500 python scratch/endfull_probe.py > gpurun_out/endfull.log 2>&1
