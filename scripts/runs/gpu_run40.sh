cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for v in diag2 diag1536 diag2560 diag4608 diag7680; do
  L="libb200_w4a16_$v.so"
  echo "== $v" >> gpurun_out/diag40.log
  B200_W4A16_LIB=$L timeout 200 python scripts/kernel_bench.py --m 16 --shapes 16384x22016 --iters 10 >> gpurun_out/diag40.log 2>&1
done
cat gpurun_out/diag40.log | cut -c1-150
