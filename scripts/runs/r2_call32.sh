#!/bin/bash
# round 2, call 32: compute-sanitizer memcheck over small launches of every kernel family
set -x
O=gpurun_out
timeout 300 python scripts/sanitize_smoke.py > $O/r3f_smoke_plain.log 2>&1 && tail -2 $O/r3f_smoke_plain.log
B200_W4A16_TWO=1 B200_W4A16_PAIR=2 timeout 300 python scripts/sanitize_smoke.py > $O/r3f_smoke_plain_two.log 2>&1 && tail -2 $O/r3f_smoke_plain_two.log
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python scripts/sanitize_smoke.py > $O/r3f_memcheck.log 2>&1; echo "memcheck rc=$?"
tail -5 $O/r3f_memcheck.log
B200_W4A16_TWO=1 B200_W4A16_PAIR=2 timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python scripts/sanitize_smoke.py > $O/r3f_memcheck_two.log 2>&1; echo "memcheck two rc=$?"
tail -5 $O/r3f_memcheck_two.log
