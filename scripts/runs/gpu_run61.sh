cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_norm.py -q -m gpu -k "packets or multi_gpu" > gpurun_out/t_ll61.log 2>&1; echo "LL tests rc=$?"; tail -12 gpurun_out/t_ll61.log | cut -c1-300
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519 scripts/norm_bench.py > gpurun_out/norm_bench61_tp2.log 2>&1; echo "rc=$?"; grep "^{" gpurun_out/norm_bench61_tp2.log || tail -5 gpurun_out/norm_bench61_tp2.log
