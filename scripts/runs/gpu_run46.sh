set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_ar46.log 2>&1; echo "ar tests rc=$?"; tail -3 gpurun_out/t_ar46.log
for mode in "" "--nccl-allreduce"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 $mode > gpurun_out/bench_tp2$mode.log 2>&1; echo "tp2 $mode rc=$?"; tail -1 gpurun_out/bench_tp2$mode.log | cut -c1-900
done
