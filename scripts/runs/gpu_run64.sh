cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 200 python -m pytest tests -q -m gpu -x > gpurun_out/t_all64.log 2>&1; echo "all tests rc=$?"; tail -3 gpurun_out/t_all64.log | cut -c1-200
timeout 60 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench64.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench64.log | cut -c1-200
