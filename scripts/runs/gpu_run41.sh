cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python scripts/trace_decode.py --m 16 --k 16384 --n 22016 --launches 4 > gpurun_out/trace41.log 2>&1
head -12 gpurun_out/trace41.log | cut -c1-330
