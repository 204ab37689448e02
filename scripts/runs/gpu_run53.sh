set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_moe.py -q -m gpu > gpurun_out/t_moe53.log 2>&1; echo "moe tests rc=$?"; tail -25 gpurun_out/t_moe53.log | cut -c1-250
