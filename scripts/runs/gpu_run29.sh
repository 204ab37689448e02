set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 300 python -m pytest tests/test_gpu_allreduce.py -x -q -m gpu -k multi_gpu > gpurun_out/t_ar$N.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/t_ar$N.log
for mode in "" "--nccl-allreduce"; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 $mode > gpurun_out/bench_tp$N$mode.log 2>&1; echo "bench $mode rc=$?"; grep '^{' gpurun_out/bench_tp$N$mode.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['n_gpus'], d['ms_per_step'], d['value'], d['tokens_per_s'], d['allreduce'], d['gpu_launches'])"
done
