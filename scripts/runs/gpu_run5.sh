set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py -q -x -m gpu > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
tail -4 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 16 > gpurun_out/kb16.log 2>&1; cat gpurun_out/kb16.log
timeout 300 python scripts/kernel_bench.py --m 1,64 > gpurun_out/kb1_64.log 2>&1; cat gpurun_out/kb1_64.log
timeout 300 python scripts/kernel_bench.py --m 512,2048,8192 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 64 > gpurun_out/kb_prefill.log 2>&1; cat gpurun_out/kb_prefill.log
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 22016 > gpurun_out/trace_22016.log 2>&1; tail -3 gpurun_out/trace_22016.log
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 4096 > gpurun_out/trace_4096.log 2>&1; tail -3 gpurun_out/trace_4096.log
