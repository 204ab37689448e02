set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python scripts/mlp_bench.py --model llama2-7b --layers 16 > gpurun_out/mlp_bench51.log 2>&1; echo "rc=$?"
timeout 300 python scripts/mlp_bench.py --model llama2-70b-tp8 --layers 32 >> gpurun_out/mlp_bench51.log 2>&1; echo "rc=$?"; cat gpurun_out/mlp_bench51.log
timeout 300 python -m pytest tests/test_gpu_norm.py -q -m gpu -k multi_gpu > gpurun_out/t_norm51.log 2>&1; echo "norm mgpu rc=$?"; tail -3 gpurun_out/t_norm51.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 scripts/norm_bench.py > gpurun_out/norm_bench51_tp2.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/norm_bench51_tp2.log
