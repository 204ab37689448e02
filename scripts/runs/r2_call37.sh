#!/bin/bash
# round 2, call 37: current captures for the record: launch list + full ncu of one 70B decode layer
# (final decode kernel), CPU-side suite on the box
set -x
O=gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --layers 2 --no-cpu-baseline --no-extras"
$CMD > $O/r3k_ncu_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r3k_launches_bench_short.csv $CMD > $O/r3k_ncu_launch.log 2>&1
$CMD > $O/r3k_ncu_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:w4a16_decode -s 16 -c 4 -o $O/r3k_ncu_decode70b $CMD > $O/r3k_ncu_full.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > $O/r3k_smoke.log 2>&1; tail -1 $O/r3k_smoke.log
