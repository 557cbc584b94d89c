# usage (on the GPU box): bash tools/hs_sweep.sh  -- rebuilds the library with several ring depths / tile sizes of the
# harmonic-summing kernel and prints the stage times of a short bench run for each
set -e
for cfg in "2 16 64" "2 8 64" "2 12 96" "2 6 192" "2 10 96"; do
  set -- $cfg
  BRP_EXTRA_NVCC_FLAGS="-DBRP_HS_STAGES=$1 -DBRP_HS_PER_THREAD=$2 -DBRP_HS_THREADS=$3" python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
  python -m pytest tests/test_gpu_hs.py -q -x 2>&1 | tail -1
  python bench.py --no-cpu-baseline --steps 1 --warmup 1 --templates 512 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('stages $1 per_thread $2 threads $3', round(d['value']), {k:round(v['ms_per_batch']*1e3,1) for k,v in d['roofline']['stages'].items()})"
done
