cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 60 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench65_tp2.log 2>&1; echo "tp2 default rc=$?"; tail -1 gpurun_out/bench65_tp2.log | cut -c1-700
