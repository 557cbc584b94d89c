set -x
python -m pytest tests/test_gpu_fft.py tests/test_gpu_fullsize.py tests/test_gpu_search.py -x -q 2>&1 | tail -8
python tools/stage_times.py --steps 2 --warmup 2
BRP_FFT_BLOCKED=0 python tools/stage_times.py --steps 2 --warmup 2
python tools/stage_times.py --steps 2 --warmup 2
BRP_FFT_BLOCKED=0 python tools/stage_times.py --steps 2 --warmup 2
python -m pytest tests/test_gpu_realbank.py tests/test_gpu_cli.py tests/test_gpu_whiten.py tests/test_gpu_resample.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -15
