#!/bin/bash
# round 2, call 59: the same sweep, three interleaved repetitions of A (all SMs) / B (aligned >= 80 %, cap 4) /
# D (aligned >= 60 %, cap 4) on the shapes whose grid changes; decode kernel: aligned >= 80 % vs all SMs, twice
set -x
O=gpurun_out
SH=4096x4096,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,5120x5120,5120x15360,13824x5120,14336x4096,4096x12288
MS=32,64,128,256,512,1024
for i in 1 2 3; do
  B200_W4A16_ALIGNPCT=0 B200_W4A16_SPLITCAP=0 python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m $MS --shapes $SH --iters 10 > $O/r4g_kb_A_$i.jsonl 2>&1
  B200_W4A16_ALIGNPCT=80 B200_W4A16_SPLITCAP=4 python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m $MS --shapes $SH --iters 10 > $O/r4g_kb_B_$i.jsonl 2>&1
  B200_W4A16_ALIGNPCT=60 B200_W4A16_SPLITCAP=4 python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m $MS --shapes $SH --iters 10 > $O/r4g_kb_D_$i.jsonl 2>&1
done
DSH=4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,8192x8192,8192x10240,28672x8192,5120x5120,13824x5120,14336x4096,4096x28672
for i in 1 2; do
  for d in 0 80; do
    B200_W4A16_DALIGNPCT=$d python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m 1,16 --shapes $DSH > $O/r4g_kb_dec${d}_$i.jsonl 2>&1
  done
done
