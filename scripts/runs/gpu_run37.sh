cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for v in base diag48 diag128 diag256 diag384 diag432 diag2; do
  if [ $v = base ]; then L=""; else L="libb200_w4a16_$v.so"; fi
  echo "== $v" >> gpurun_out/diag37.log
  B200_W4A16_LIB=$L timeout 200 python scripts/kernel_bench.py --m 16 --shapes 16384x22016 --iters 10 >> gpurun_out/diag37.log 2>&1
done
cat gpurun_out/diag37.log | cut -c1-150
