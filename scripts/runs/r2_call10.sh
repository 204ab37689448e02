#!/bin/bash
# round 2, call 10: decode vs tcdecode at M <= 16, tcdecode vs tcgen05 at 24..64, prefill sweep,
# decode / tcdecode timelines, launch list + one full ncu capture of the 70B decode layer
set -x
O=gpurun_out
SH7=4096x4096,4096x12288,4096x22016,11008x4096
SH70=8192x10240,8192x8192,8192x57344,28672x8192
SH70TP8=8192x1280,1024x8192,8192x7168,3584x8192
for k in decode tcdecode; do
timeout 300 python scripts/kernel_bench.py --kernel $k --m 1,8,16 --shapes $SH7,$SH70TP8 --iters 10 >> $O/r2j_kb_$k.jsonl 2>&1
timeout 300 python scripts/kernel_bench.py --kernel $k --m 8,16 --shapes $SH70 --iters 5 --min-mb 300 >> $O/r2j_kb_$k.jsonl 2>&1
done
for k in tcdecode tcgen05; do
timeout 300 python scripts/kernel_bench.py --kernel $k --m 24,32,48,64 --shapes 4096x4096,4096x22016,11008x4096,1024x8192,8192x7168,8192x57344 --iters 10 >> $O/r2j_kb_mid_$k.jsonl 2>&1
done
timeout 300 python scripts/kernel_bench.py --m 128,256,512,1024,2048,4096,8192 --shapes 5120x15360,5120x5120,5120x27648,13824x5120 --iters 5 --min-mb 100 >> $O/r2j_kb_prefill.jsonl 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > $O/r2j_trace_decode_4096.log 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 8192 --n 57344 > $O/r2j_trace_decode_57344.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 4096 --n 4096 --ctas 0,73 > $O/r2j_trace_tcdec_4096.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 8192 --n 57344 --ctas 0,73 > $O/r2j_trace_tcdec_57344.log 2>&1
timeout 120 python scripts/trace_gemm.py --m 2048 --k 5120 --n 15360 --ctas 0,73 > $O/r2j_trace_prefill.log 2>&1
# ncu: the bench command on 2 layers first plain, then the launch list, then one full capture
CMD="python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline --no-extras"
$CMD > $O/r2j_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/r2j_launches_bench_short.csv $CMD > $O/r2j_ncu_launch.log 2>&1
$CMD > $O/r2j_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -s 16 -c 4 -o $O/r2j_ncu_decode70b $CMD > $O/r2j_ncu_full.log 2>&1
ls -la $O
