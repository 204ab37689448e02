#!/bin/bash
# round 2, call 58: grid sizing when the output units are fewer than the SMs: unit-aligned ranges and a cap
# on the workers per unit (tuning build, env knobs), against the old all-SMs plan
set -x
O=gpurun_out
SH=4096x4096,4096x12288,4096x22016,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,5120x5120,5120x15360,13824x5120,5120x27648,4096x28672,14336x4096
MS=24,32,64,128,256,512,1024,2048
run() { # name, env...
  name=$1; shift
  env "$@" python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m $MS --shapes $SH --iters 10 > $O/r4f_kb_$name.jsonl 2>&1
}
run A B200_W4A16_ALIGNPCT=0 B200_W4A16_SPLITCAP=0
run B B200_W4A16_ALIGNPCT=80 B200_W4A16_SPLITCAP=4
run C B200_W4A16_ALIGNPCT=80 B200_W4A16_SPLITCAP=6
run D B200_W4A16_ALIGNPCT=60 B200_W4A16_SPLITCAP=4
run E B200_W4A16_ALIGNPCT=80 B200_W4A16_SPLITCAP=4 B200_W4A16_SPLITMINTOK=32
run F B200_W4A16_ALIGNPCT=90 B200_W4A16_SPLITCAP=4
DSH=4096x4096,4096x12288,11008x4096,8192x1280,1024x8192,8192x7168,3584x8192,8192x8192
for d in 0 60 80; do
  B200_W4A16_DALIGNPCT=$d python scripts/kernel_bench.py --lib libb200_w4a16_tune.so --m 1,8,16 --shapes $DSH > $O/r4f_kb_dec$d.jsonl 2>&1
done
