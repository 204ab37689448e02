set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for f in 0 2 4 6 1; do
B200_W4A16_FLAGS=$f timeout 300 python scripts/kernel_bench.py --m 16 --shapes 4096x22016,4096x4096 > gpurun_out/kb_f$f.log 2>&1; echo "flags=$f"; cat gpurun_out/kb_f$f.log
done
