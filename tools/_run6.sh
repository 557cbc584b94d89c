set -x
python -m pytest tests/test_gpu_hs.py tests/test_gpu_search.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -8
python tools/stage_times.py --steps 2 --warmup 2
BRP_HS_V1=1 python tools/stage_times.py --steps 2 --warmup 2
python tools/stage_times.py --steps 2 --warmup 2
python tools/stage_times.py --steps 2 --warmup 2 --padding 3 --templates 2048
BRP_HS_V1=1 python tools/stage_times.py --steps 2 --warmup 2 --padding 3 --templates 2048
