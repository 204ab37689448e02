cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for v in base diag2 diag7; do
  if [ $v = base ]; then L=""; else L="libb200_w4a16_$v.so"; fi
  echo "== $v" >> gpurun_out/clk36.log
  B200_W4A16_LIB=$L timeout 200 python scripts/clock_probe.py --m 16 --shape 16384x22016 --seconds 3 >> gpurun_out/clk36.log 2>&1
done
echo "== base 4096x22016" >> gpurun_out/clk36.log
timeout 200 python scripts/clock_probe.py --m 16 --shape 4096x22016 --seconds 2 >> gpurun_out/clk36.log 2>&1
cat gpurun_out/clk36.log | cut -c1-200
