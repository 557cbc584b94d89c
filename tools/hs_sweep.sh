# usage (on the GPU box): bash tools/hs_sweep.sh  -- rebuilds the library with several ring depths / tile sizes of the
# harmonic-summing kernel (stages, bins per thread, consumer threads) and prints the stage times of a short bench run for each
set -e
for cfg in "2 8 128" "1 8 128" "1 8 64" "1 4 128" "2 4 128" "1 16 64" "2 8 128"; do
  set -- $cfg
  BRP_EXTRA_NVCC_FLAGS="-DBRP_HS_STAGES=$1 -DBRP_HS_PER_THREAD=$2 -DBRP_HS_THREADS=$3" python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
  python -m pytest tests/test_gpu_hs.py -q -x 2>&1 | tail -1
  echo "stages $1 per_thread $2 threads $3"; python tools/stage_times.py --steps 1 --warmup 1 --templates 1024
done
python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
