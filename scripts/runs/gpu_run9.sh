set -x
nvidia-smi -L
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"
tail -25 gpurun_out/t_all.log
timeout 300 python scripts/kernel_bench.py --m 1,16,64 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
timeout 300 python scripts/kernel_bench.py --m 512,2048,8192 --shapes 5120x15360,13824x5120 --iters 5 --min-mb 64 > gpurun_out/kb_prefill.log 2>&1; cat gpurun_out/kb_prefill.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench1.log 2>&1; echo "bench rc=$?"
tail -3 gpurun_out/bench1.log
