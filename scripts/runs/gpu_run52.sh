set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_checkpoint.py -q -m gpu > gpurun_out/t_ckpt52.log 2>&1; echo "ckpt tests rc=$?"; tail -25 gpurun_out/t_ckpt52.log | cut -c1-250
