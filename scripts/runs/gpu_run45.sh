set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 400 python -m pytest tests -x -q -m gpu > gpurun_out/t_all45.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all45.log
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r1d.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_r1d.log | cut -c1-1500
timeout 200 python scripts/kernel_bench.py --m 1,2,4,8,16 --shapes 4096x4096,4096x12288,4096x22016,11008x4096 --iters 10 > gpurun_out/kernel_bench_decode_r1d.jsonl 2>&1
timeout 200 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_short.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1d_launches_bench_short.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 200 python bench.py --steps 1 --warmup 3 --layers 1 --no-cpu-baseline > gpurun_out/bench_1layer.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -c 4 -o gpurun_out/prof_r1d_decode_layer -f python bench.py --steps 1 --warmup 3 --layers 1 --no-cpu-baseline > gpurun_out/ncu_layer.log 2>&1
tail -2 gpurun_out/ncu_layer.log
