set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -12
python tests/legacy_speed_probe.py 600 2>&1 | tail -2
python bench.py --steps 3 --warmup 3 > gpurun_out/r2_bench_a.json 2> gpurun_out/r2_bench_a.err; tail -c 3000 gpurun_out/r2_bench_a.json; tail -5 gpurun_out/r2_bench_a.err
