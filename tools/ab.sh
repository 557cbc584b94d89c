# usage (on the GPU box): bash tools/ab.sh "<nvcc flags A>" "<nvcc flags B>"  -- builds both variants and prints stage times, twice each
set -e
for round in 1 2; do
  for fl in "$1" "$2"; do
    BRP_EXTRA_NVCC_FLAGS="$fl" python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
    echo "flags [$fl]"; python tools/stage_times.py --steps 1 --warmup 1 --templates 1024
  done
done
