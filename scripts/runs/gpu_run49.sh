set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_norm.py tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_norm49.log 2>&1; echo "norm tests rc=$?"; tail -5 gpurun_out/t_norm49.log
timeout 200 python scripts/norm_bench.py > gpurun_out/norm_bench49.log 2>&1; echo "rc=$?"; cat gpurun_out/norm_bench49.log | tail -6
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 scripts/norm_bench.py > gpurun_out/norm_bench49_tp2.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/norm_bench49_tp2.log
timeout 400 python -m pytest tests -x -q -m gpu > gpurun_out/t_all49.log 2>&1; echo "all tests rc=$?"; tail -3 gpurun_out/t_all49.log
