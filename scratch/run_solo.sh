mkdir -p gpurun_out
for w in 1 2 4 8; do
timeout 300 python scratch/shard_probe.py $w 2>/dev/null | tail -1 >> gpurun_out/solo_sweep3.log
done
