cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 90 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 --launches 6 > gpurun_out/trace44_4096.log 2>&1
grep -A14 "CTAs by exit" gpurun_out/trace44_4096.log
