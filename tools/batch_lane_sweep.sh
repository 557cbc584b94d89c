# usage (on the GPU box): bash tools/batch_lane_sweep.sh  -- templates/s of the whole loop against batch size and number of lanes
# (small batches keep the per-batch working set inside the 126 MB L2; lanes overlap the kernels of different batches)
for b in 2 3 4 8 16 64; do
  for s in 1 2 3; do
    echo "batch $b lanes $s"; BRP_STREAMS=$s python tools/stage_times.py --steps 2 --warmup 2 --batch $b --templates 2048
  done
done
