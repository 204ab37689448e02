set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
python scripts/kernel_bench.py --m 16 > gpurun_out/kb_pdl.log 2>&1; tail -5 gpurun_out/kb_pdl.log
B200_W4A16_NO_PDL=1 python scripts/kernel_bench.py --m 16 > gpurun_out/kb_nopdl.log 2>&1; tail -5 gpurun_out/kb_nopdl.log
python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm -c 1 -o gpurun_out/prof_r1_gemm16 -f python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/ncu.log 2>&1
tail -3 gpurun_out/ncu.log
