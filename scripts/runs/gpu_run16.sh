set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for f in 8 10; do
for m in 1 16; do
B200_W4A16_FLAGS=$f timeout 300 python scripts/kernel_bench.py --m $m --shapes 4096x22016 > gpurun_out/kb_f$f.log 2>&1; echo "flags=$f"; cat gpurun_out/kb_f$f.log
done; done
