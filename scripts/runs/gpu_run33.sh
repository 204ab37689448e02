set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_all.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_r1c.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_r1c.log | cut -c1-400
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench_ref.log | cut -c1-600
timeout 300 python bench.py --steps 1 --warmup 3 --layers 1 --no-cpu-baseline > gpurun_out/bench_1layer.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -c 4 -o gpurun_out/prof_r1c_decode_layer -f python bench.py --steps 1 --warmup 3 --layers 1 --no-cpu-baseline > gpurun_out/ncu_layer.log 2>&1
tail -2 gpurun_out/ncu_layer.log
