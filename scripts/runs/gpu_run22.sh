set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/t_all.log
timeout 300 python scripts/kernel_bench.py --m 1,8,16 > gpurun_out/kb_decode.log 2>&1; cat gpurun_out/kb_decode.log
timeout 300 python scripts/kernel_bench.py --m 32,64,128,256 > gpurun_out/kb_mid.log 2>&1; cat gpurun_out/kb_mid.log
