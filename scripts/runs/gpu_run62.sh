cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_allreduce.py tests/test_gpu_norm.py -q -m gpu -k "packet or multi_gpu" > gpurun_out/t_ll62.log 2>&1; echo "LL tests rc=$?"; tail -6 gpurun_out/t_ll62.log | cut -c1-300
for mode in "" "--unfused-allreduce"; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline $mode > gpurun_out/bench62_tp2$mode.log 2>&1; echo "tp2 $mode rc=$?"; tail -1 gpurun_out/bench62_tp2$mode.log | cut -c1-260
done
