set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu -k "small" > gpurun_out/t_gemm_small.log 2>&1; echo "gemm small rc=$?"
tail -15 gpurun_out/t_gemm_small.log
timeout 900 python -m pytest tests/test_gpu_gemm.py -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
tail -25 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 16 > gpurun_out/kb16.log 2>&1; cat gpurun_out/kb16.log
timeout 300 python scripts/kernel_bench.py --m 1,64 > gpurun_out/kb1_64.log 2>&1; cat gpurun_out/kb1_64.log
timeout 300 python scripts/kernel_bench.py --m 512,2048 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 64 > gpurun_out/kb_prefill.log 2>&1; cat gpurun_out/kb_prefill.log
