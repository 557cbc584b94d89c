# usage (on an N-GPU box): bash tools/r2_multi.sh N  -- strong-scaling bench over N GPUs (one process per GPU, NCCL) and the C-side
# split of the bank over the N devices of one process (brp_b200_search --devices), checked against the one-device run
set -x
N=${1:-2}
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
python -m pytest tests/test_gpu_cli.py -x -q -k "several_devices" 2>&1 | tail -2
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 \
   > gpurun_out/r2_bench_${N}gpu.json 2> gpurun_out/r2_bench_${N}gpu.err
tail -c 1500 gpurun_out/r2_bench_${N}gpu.json; tail -3 gpurun_out/r2_bench_${N}gpu.err
# the same split from C: wall clock of the whole command (ingest + whitening + search of the real work-unit, full real bank)
D=tests/golden/realwu
for devs in 0 0-$((N-1)); do
  boinc-app-eah-brp_b200/brp_b200_search -i $D/real_wu_1099.bin4 -t $D/stochastic_full.bank -l $D/real_wu.zap \
     -o gpurun_out/c_$devs.cand -A 0.08 -P 3.0 -f 400.0 -W -z --devices $devs 2>&1 | tail -2
done
cmp gpurun_out/c_0.cand gpurun_out/c_0-$((N-1)).cand && echo "candidate files identical"
