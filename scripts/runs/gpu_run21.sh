set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 120 python scripts/trace_decode.py --m 16 --k 16384 --n 22016 --launches 2 > gpurun_out/td_big.log 2>&1; head -4 gpurun_out/td_big.log; grep -A29 "CTA 74:" gpurun_out/td_big.log
