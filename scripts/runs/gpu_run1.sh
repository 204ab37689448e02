set -x
nvidia-smi -L
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_aux.py -x -q -m gpu > gpurun_out/t_aux.log 2>&1; echo "aux rc=$?"
tail -15 gpurun_out/t_aux.log
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu -k "small" > gpurun_out/t_gemm_small.log 2>&1; echo "gemm small rc=$?"
tail -30 gpurun_out/t_gemm_small.log
timeout 900 python -m pytest tests/test_gpu_gemm.py -q -m gpu > gpurun_out/t_gemm.log 2>&1; echo "gemm rc=$?"
tail -40 gpurun_out/t_gemm.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"
tail -5 gpurun_out/bench1.log
