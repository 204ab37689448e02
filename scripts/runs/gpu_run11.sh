set -x
nvidia-smi -L
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 60 scripts/probes/hmma_probe > gpurun_out/hmma_probe.txt 2>&1; cat gpurun_out/hmma_probe.txt
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/t_gemm.log
timeout 300 python scripts/kernel_bench.py --m 1,8,16 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
B200_W4A16_KERNEL=tcgen05 timeout 300 python scripts/kernel_bench.py --m 16 > gpurun_out/kb_decode_tc.log 2>&1; cat gpurun_out/kb_decode_tc.log
