set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 32,64,128,256 > gpurun_out/kb_mid.log 2>&1; cat gpurun_out/kb_mid.log
timeout 300 python scripts/kernel_bench.py --m 512,2048,8192 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 64 > gpurun_out/kb_prefill.log 2>&1; cat gpurun_out/kb_prefill.log
