set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_norm.py tests/test_gpu_allreduce.py -q -m gpu -k multi_gpu > gpurun_out/t_mgpu59.log 2>&1; echo "multi-gpu tests rc=$?"; tail -4 gpurun_out/t_mgpu59.log | cut -c1-300
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29519 scripts/norm_bench.py > gpurun_out/norm_bench59_tp8.log 2>&1; echo "rc=$?"; grep "^{" gpurun_out/norm_bench59_tp8.log
for mode in "" "--nccl-allreduce"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 $mode > gpurun_out/bench59_tp8$mode.log 2>&1; echo "tp8 $mode rc=$?"; tail -1 gpurun_out/bench59_tp8$mode.log | cut -c1-330
done
timeout 100 python scripts/norm_bench.py > gpurun_out/norm_bench59.log 2>&1; tail -5 gpurun_out/norm_bench59.log | cut -c1-140
