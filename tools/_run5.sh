set -x
CMD="python bench.py --steps 1 --warmup 1 --templates 128 --no-cpu-baseline --no-other-configs"
$CMD > gpurun_out/plain5.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'fft_pass1_tma|fft_pass2_tma|harmonic_summing|resample_kernel' -s 8 -c 4 -o gpurun_out/r2_prof_a $CMD > gpurun_out/ncu5.log 2>&1
tail -3 gpurun_out/ncu5.log
