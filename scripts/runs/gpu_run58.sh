cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
for t in 32 100000; do
echo "cpt4_min_rows=$t"
B200_NORM_CPT4_MIN_ROWS=$t timeout 200 python scripts/norm_bench.py 32x8192 64x8192 128x8192 256x8192 512x8192 256x4096 | cut -c1-120
done
