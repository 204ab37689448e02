set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python tests/golden/make_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; tail -3 gpurun_out/golden.log
timeout 600 python scripts/ref_kernel_bench.py > gpurun_out/ref_kernel_bench.jsonl 2>&1; cat gpurun_out/ref_kernel_bench.jsonl
