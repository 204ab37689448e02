set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024,4096 --layers 2 --experts-per-rank 8 --iters 5 > gpurun_out/moe54.log 2>&1; echo "rc=$?"
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024,4096 --layers 8 --experts-per-rank 1 --iters 5 >> gpurun_out/moe54.log 2>&1; echo "rc=$?"
cat gpurun_out/moe54.log | cut -c1-300
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/bench54.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench54.log | cut -c1-1500
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench54_ref.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/bench54_ref.log | cut -c1-600
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
