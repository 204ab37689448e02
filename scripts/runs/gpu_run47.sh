cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_moe.py -x -q -m gpu > gpurun_out/t_moe47.log 2>&1; echo "moe tests rc=$?"; tail -25 gpurun_out/t_moe47.log
timeout 200 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t_gemm47.log 2>&1; echo "gemm tests rc=$?"; tail -2 gpurun_out/t_gemm47.log
