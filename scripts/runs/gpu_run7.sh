set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python scripts/debug_parity.py > gpurun_out/dbg_pdl.log 2>&1; tail -60 gpurun_out/dbg_pdl.log
B200_W4A16_NO_PDL=1 timeout 600 python scripts/debug_parity.py --shapes 512x5120x27648,512x4096x4096 > gpurun_out/dbg_nopdl.log 2>&1; tail -30 gpurun_out/dbg_nopdl.log
