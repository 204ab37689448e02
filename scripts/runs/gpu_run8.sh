set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 120 scripts/probes/mma_probe > gpurun_out/mma_probe.log 2>&1; cat gpurun_out/mma_probe.log
timeout 600 python scripts/debug_parity.py --shapes 1024x4096x4096,1024x4096x4096,1024x4096x4096,512x5120x27648,2048x4096x4096,384x4096x12288 > gpurun_out/dbg_pdl.log 2>&1; tail -30 gpurun_out/dbg_pdl.log
timeout 900 python -m pytest tests -q -x -m gpu > gpurun_out/t_all.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/t_all.log
