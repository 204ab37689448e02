set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_moe.py -q -m gpu > gpurun_out/t_moe55.log 2>&1; echo "moe tests rc=$?"; tail -5 gpurun_out/t_moe55.log | cut -c1-250
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024,4096 --layers 2 --experts-per-rank 8 --iters 5 > gpurun_out/moe55.log 2>&1; echo "rc=$?"
timeout 300 python scripts/moe_bench.py --prefill --tokens 64,256,1024,4096 --layers 8 --experts-per-rank 1 --iters 5 >> gpurun_out/moe55.log 2>&1; echo "rc=$?"
cat gpurun_out/moe55.log | cut -c1-300
# launch list of the contract bench (short run), then full captures of the new kernels
timeout 300 python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench55_short.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r1e_launches_bench_short.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu55_list.log 2>&1; echo "ncu list rc=$?"
timeout 200 python scripts/mlp_bench.py --model llama2-7b --layers 2 --tokens 16 > gpurun_out/mlp55_once.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -s 8 -c 2 -o gpurun_out/prof_r1e_mlp_fused_decode16 -f python scripts/mlp_bench.py --model llama2-7b --layers 2 --tokens 16 > gpurun_out/ncu55_mlp.log 2>&1; echo "ncu mlp rc=$?"
timeout 200 python scripts/norm_bench.py > gpurun_out/norm55_once.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:add_rmsnorm -s 700 -c 1 -o gpurun_out/prof_r1e_add_rmsnorm_4096x8192 -f python scripts/norm_bench.py > gpurun_out/ncu55_norm.log 2>&1; echo "ncu norm rc=$?"
ls -la gpurun_out | tail -8
