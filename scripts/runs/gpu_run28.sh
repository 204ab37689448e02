set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_ar.log 2>&1; echo "tests rc=$?"
tail -25 gpurun_out/t_ar.log
