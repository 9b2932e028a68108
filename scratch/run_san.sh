mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 7 --log-file gpurun_out/memcheck.log python -m pytest tests/test_gpu_score.py tests/test_gpu_endfull.py -x -q -k "transposed_giants and sw or giant_teams and nw or test_random_db and 300 or topk or overflow or static_32 or golden or top_hits or homologs or batch" > gpurun_out/t_san.log 2>&1; echo "rc=$?" >> gpurun_out/t_san.log
tail -5 gpurun_out/memcheck.log
