set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_activation.py tests/test_gpu_norm.py -q -m gpu > gpurun_out/t_act50.log 2>&1; echo "act tests rc=$?"; tail -30 gpurun_out/t_act50.log | cut -c1-300
timeout 400 python -m pytest tests -x -q -m gpu > gpurun_out/t_all50.log 2>&1; echo "all tests rc=$?"; tail -3 gpurun_out/t_all50.log
