set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python scripts/norm_bench.py 4096x8192 > gpurun_out/norm56_once.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:add_rmsnorm -s 3 -c 1 -o gpurun_out/prof_r1e_add_rmsnorm_4096x8192 -f python scripts/norm_bench.py 4096x8192 > gpurun_out/ncu56_norm.log 2>&1; echo "ncu norm rc=$?"
timeout 600 python -m pytest tests -q -m gpu > gpurun_out/t_all56.log 2>&1; echo "all tests rc=$?"; tail -3 gpurun_out/t_all56.log
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench56.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench56.log | cut -c1-300
for mode in "" "--nccl-allreduce"; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 10 --warmup 3 $mode > gpurun_out/bench56_tp2$mode.log 2>&1; echo "tp2 $mode rc=$?"; tail -1 gpurun_out/bench56_tp2$mode.log | cut -c1-400
done
