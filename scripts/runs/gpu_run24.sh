set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_r1b.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_r1b.log
timeout 300 python bench.py --steps 3 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_r1b.csv python bench.py --steps 3 --warmup 3 --layers 4 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
tail -2 gpurun_out/ncu_launches.log
timeout 300 python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/once.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -c 1 -o gpurun_out/prof_r1b_decode16 -f python scripts/kernel_bench.py --m 16 --shapes 4096x22016 --once > gpurun_out/ncu_full.log 2>&1
tail -3 gpurun_out/ncu_full.log
