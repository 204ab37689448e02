set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_golden.py tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 1,16 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
timeout 300 python scripts/kernel_bench.py --m 1,16 --shapes 16384x22016 --min-mb 700 --iters 10
