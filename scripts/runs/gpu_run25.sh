set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_ar.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/t_ar.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/bench_tp2.log 2>&1; echo "bench rc=$?"; tail -2 gpurun_out/bench_tp2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --nccl-allreduce > gpurun_out/bench_tp2_nccl.log 2>&1; echo "bench rc=$?"; tail -2 gpurun_out/bench_tp2_nccl.log
