set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/t_gemm.log
for f in 0 2 6; do
B200_W4A16_FLAGS=$f timeout 300 python scripts/kernel_bench.py --m 16 --shapes 4096x22016,4096x4096 > gpurun_out/kb_f$f.log 2>&1; echo "flags=$f"; cat gpurun_out/kb_f$f.log
done
timeout 300 python scripts/kernel_bench.py --m 1,16 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 22016 --launches 3 > gpurun_out/td_22016.log 2>&1; cat gpurun_out/td_22016.log
