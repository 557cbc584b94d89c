# usage (on the GPU box): bash tools/resamp_sweep.sh -- rebuilds with different occupancy targets for the resampler
set -e
for mb in 8 6 5 4; do
  BRP_EXTRA_NVCC_FLAGS="-DBRP_RESAMP_MIN_BLOCKS=$mb" python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
  python bench.py --no-cpu-baseline --steps 1 --warmup 1 --templates 1024 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('min_blocks $mb', round(d['value']), {k:round(v['ms_per_batch']*1e3,1) for k,v in d['roofline']['stages'].items()})"
done
