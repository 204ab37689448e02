cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python scripts/moe_bench.py --tokens 1,4,16 --layers 3 --experts-per-rank 8 > gpurun_out/moe48.log 2>&1; echo "rc=$?"
timeout 300 python scripts/moe_bench.py --tokens 1,4,16,64 --layers 8 --experts-per-rank 1 >> gpurun_out/moe48.log 2>&1; echo "rc=$?"
tail -12 gpurun_out/moe48.log | cut -c1-300
