# usage (on the GPU box): bash tools/r2_capture.sh  -- the round-2 evidence set: full GPU test suite, the legacy-path rate, the bench
# line, then the ncu launch list and one --set full capture of the same short bench command (profiles/README.md)
set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -6
tools/bin/legacy_rate 2000 2>&1 | tee gpurun_out/r2_legacy_rate.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_v6.json 2> gpurun_out/r2_bench_v6.err; tail -c 600 gpurun_out/r2_bench_v6.json
CMD="python bench.py --steps 1 --warmup 1 --templates 256 --no-cpu-baseline --no-other-configs"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_v6.csv $CMD > gpurun_out/ncu_l.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'fft_pass1|fft_pass2_tma|harmonic_summing|resample_kernel' -s 16 -c 4 -o gpurun_out/r2_ncu_full_v6 $CMD > gpurun_out/ncu_f.log 2>&1
tail -2 gpurun_out/ncu_f.log
