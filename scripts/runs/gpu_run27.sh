set -x
cd $GRAFT_REPO_ROOT; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce.py -x -q -m gpu > gpurun_out/t_ar.log 2>&1; echo "tests rc=$?"
tail -25 gpurun_out/t_ar.log
for mode in "" "--unfused-allreduce" "--nccl-allreduce"; do
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 $mode > gpurun_out/bench_tp2$mode.log 2>&1; echo "bench $mode rc=$?"; grep '^{' gpurun_out/bench_tp2$mode.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['value'], d['allreduce'], d['gpu_launches'])"
done
