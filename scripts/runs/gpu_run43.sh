cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t43.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t43.log
timeout 90 python scripts/kernel_bench.py --m 16,1 --shapes 4096x4096,4096x12288,4096x22016,11008x4096,16384x22016 --iters 10 > gpurun_out/kb43.log 2>&1
cat gpurun_out/kb43.log | cut -c1-150
timeout 90 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 --launches 6 > gpurun_out/trace43_4096.log 2>&1
grep -A14 "CTAs by exit" gpurun_out/trace43_4096.log
