mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_fullsize.py -x -q -k "whole_database_end" --durations=8 > gpurun_out/t_c4.log 2>&1; echo "rc=$?" >> gpurun_out/t_c4.log
