set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
bash tools/batch_lane_sweep.sh > gpurun_out/r2_batch_lane_sweep.log 2>&1
python bench.py --steps 1 --warmup 1 --templates 128 --no-cpu-baseline > gpurun_out/plain.log 2>&1 && \
ncu --metrics regex:'^lts__t_sectors.*\.sum$',regex:'^lts__t_requests.*\.sum$',regex:'^l1tex__m_.*sectors.*\.sum$',dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:'fft_pass' -s 4 -c 4 --csv --log-file gpurun_out/r2_l2_breakdown.csv python bench.py --steps 1 --warmup 1 --templates 128 --no-cpu-baseline > gpurun_out/ncu.log 2>&1
echo done
