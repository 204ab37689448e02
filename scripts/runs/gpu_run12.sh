set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > gpurun_out/td_4096.log 2>&1; cat gpurun_out/td_4096.log
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 22016 > gpurun_out/td_22016.log 2>&1; cat gpurun_out/td_22016.log
timeout 120 python scripts/trace_decode.py --m 1 --k 4096 --n 22016 > gpurun_out/td_22016_m1.log 2>&1; cat gpurun_out/td_22016_m1.log
