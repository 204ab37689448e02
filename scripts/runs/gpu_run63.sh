cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench63_tp8.log 2>&1; echo "tp8 rc=$?"; tail -1 gpurun_out/bench63_tp8.log | cut -c1-900
