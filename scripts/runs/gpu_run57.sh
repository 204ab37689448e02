set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_norm.py -q -m gpu > gpurun_out/t_norm57.log 2>&1; echo "norm tests rc=$?"; tail -5 gpurun_out/t_norm57.log | cut -c1-250
timeout 200 python scripts/norm_bench.py > gpurun_out/norm_bench57.log 2>&1; echo "rc=$?"; cat gpurun_out/norm_bench57.log | tail -6
