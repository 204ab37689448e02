cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_norm.py tests/test_gpu_allreduce.py -q -m gpu -k multi_gpu > gpurun_out/t_mgpu60.log 2>&1; echo "multi-gpu tests rc=$?"; tail -3 gpurun_out/t_mgpu60.log | cut -c1-300
