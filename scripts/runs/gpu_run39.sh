cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t39.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t39.log
timeout 200 python scripts/kernel_bench.py --m 16,1 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,16384x22016 --iters 10 > gpurun_out/kb39.log 2>&1
cat gpurun_out/kb39.log | cut -c1-150
