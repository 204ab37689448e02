cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for v in base diag0g2 diag0g4; do
  if [ $v = base ]; then L=""; else L="libb200_w4a16_$v.so"; fi
  echo "== $v" >> gpurun_out/diag35.log
  B200_W4A16_LIB=$L timeout 200 python scripts/kernel_bench.py --m 16,1 --shapes 4096x4096,4096x22016,11008x4096,16384x22016 --iters 10 >> gpurun_out/diag35.log 2>&1
done
cat gpurun_out/diag35.log | cut -c1-150
