mkdir -p gpurun_out
for w in 8 1; do
PYOPAL_CUDA_DEBUG=1 timeout 300 python scratch/shard_probe.py $w > gpurun_out/shard$w.log 2> gpurun_out/shard$w.err
PYOPAL_CUDA_ONE_TIER=1 PYOPAL_CUDA_DEBUG=1 timeout 300 python scratch/shard_probe.py $w > gpurun_out/shard${w}_one.log 2> gpurun_out/shard${w}_one.err
done
