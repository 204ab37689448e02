set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 4096 > gpurun_out/trace_4096.log 2>&1; head -60 gpurun_out/trace_4096.log
timeout 300 python scripts/trace_gemm.py --m 16 --k 4096 --n 22016 --ctas 0,73 > gpurun_out/trace_22016.log 2>&1; head -8 gpurun_out/trace_22016.log
timeout 300 python scripts/trace_gemm.py --m 64 --k 4096 --n 4096 --ctas 73 > gpurun_out/trace_4096_m64.log 2>&1; head -8 gpurun_out/trace_4096_m64.log
timeout 300 python bench.py --steps 3 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 3 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
timeout 300 python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_gemm -c 1 -o gpurun_out/prof_r1_gemm16 -f python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
