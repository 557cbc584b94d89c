# A/B of BRP_FFT_P1_TB_SMEM on one box: two builds, stage times at -P 1 and -P 3
for v in 0 1 0 1; do
  BRP_EXTRA_NVCC_FLAGS="-DBRP_FFT_P1_TB_SMEM=$v" python boinc-app-eah-brp_b200/build.py --force > /dev/null 2>&1
  echo "TB_SMEM=$v"
  python tools/stage_times.py --steps 1 --warmup 1 --templates 1024 | tail -1
  python tools/stage_times.py --steps 1 --warmup 1 --templates 1024 --padding 3 | tail -1
done
