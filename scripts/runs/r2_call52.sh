#!/bin/bash
# round 2, call 52: same-box head to head: the reference's own kernels (oracle/_ref, unmodified, rebuilt for
# sm_100a) against the final kernels on the benchmark shapes (7B layer, 70B layer, 13B prefill, Mixtral expert)
set -x
O=gpurun_out
S7=4096x4096,4096x12288,4096x22016,11008x4096
S70=8192x10240,8192x8192,8192x57344,28672x8192
S13=5120x15360,5120x5120,5120x27648,13824x5120
SMX=4096x28672,14336x4096
timeout 500 python scripts/ref_kernel_bench.py --m 1,16,64,256 --shapes $S7,$SMX > $O/r3z_ref_7b_mixtral.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 1,16,64,256 --shapes $S7,$SMX > $O/r3z_b200_7b_mixtral.jsonl 2>&1
timeout 500 python scripts/ref_kernel_bench.py --m 1,16 --shapes $S70 --iters 5 > $O/r3z_ref_70b.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 1,16 --shapes $S70 > $O/r3z_b200_70b.jsonl 2>&1
timeout 500 python scripts/ref_kernel_bench.py --m 512,2048 --shapes $S13 --iters 3 > $O/r3z_ref_13b.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --m 512,2048 --shapes $S13 --iters 5 > $O/r3z_b200_13b.jsonl 2>&1
tail -2 $O/r3z_ref_13b.jsonl | cut -c1-200; tail -2 $O/r3z_b200_13b.jsonl | cut -c1-200
