set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
nvidia-smi --query-gpu=clocks.sm,clocks.mem,power.draw --format=csv
for v in base diag1 diag2 diag4 diag5 diag7 diag8 diag64 diag16 diag32; do
  if [ $v = base ]; then L=""; else L="libb200_w4a16_$v.so"; fi
  echo "== $v" >> gpurun_out/diag34.log
  B200_W4A16_LIB=$L timeout 200 python scripts/kernel_bench.py --m 16 --shapes 4096x22016,16384x22016 --iters 10 >> gpurun_out/diag34.log 2>&1
  B200_W4A16_LIB=$L timeout 200 python scripts/kernel_bench.py --m 1 --shapes 4096x22016 --iters 10 >> gpurun_out/diag34.log 2>&1
done
cat gpurun_out/diag34.log | cut -c1-150
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_all34.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all34.log
