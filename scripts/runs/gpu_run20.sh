set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for f in 38 34; do
B200_W4A16_FLAGS=$f timeout 300 python scripts/kernel_bench.py --m 16 --shapes 16384x22016 --min-mb 700 --iters 10 > gpurun_out/kb_big_f$f.log 2>&1; echo "flags=$f"; cat gpurun_out/kb_big_f$f.log
done
