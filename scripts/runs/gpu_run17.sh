set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 22016 --launches 3 > gpurun_out/td_22016.log 2>&1; cat gpurun_out/td_22016.log
