cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python scripts/kernel_bench.py --m 16 --shapes 16384x22016 --once > gpurun_out/once38.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -c 1 -o gpurun_out/prof_r1d_decode16_16384x22016 -f python scripts/kernel_bench.py --m 16 --shapes 16384x22016 --once > gpurun_out/ncu38.log 2>&1
tail -2 gpurun_out/ncu38.log; ls -la gpurun_out
