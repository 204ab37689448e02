set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
tail -8 gpurun_out/t_gemm.log
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 22016 > gpurun_out/trace_22016.log 2>&1; tail -3 gpurun_out/trace_22016.log
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 4096 > gpurun_out/trace_4096.log 2>&1; tail -3 gpurun_out/trace_4096.log
timeout 300 python scripts/trace_gemm.py --m 64 --k 4096 --n 4096 > gpurun_out/trace_4096_m64.log 2>&1; tail -3 gpurun_out/trace_4096_m64.log
B200_W4A16_FLAGS=1 timeout 300 python scripts/kernel_bench.py --m 16 > gpurun_out/kb16_noevict.log 2>&1; cat gpurun_out/kb16_noevict.log
