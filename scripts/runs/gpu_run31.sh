set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
B200_W4A16_KERNEL=tcgen05 timeout 300 python scripts/trace_gemm.py --m 64 --k 4096 --n 22016 --ctas 73 > gpurun_out/trace_tc64.log 2>&1; head -50 gpurun_out/trace_tc64.log
