set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 1,16 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 --launches 4 > gpurun_out/td_4096.log 2>&1; cat gpurun_out/td_4096.log
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 22016 --launches 4 > gpurun_out/td_22016.log 2>&1; cat gpurun_out/td_22016.log
