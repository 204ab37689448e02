#!/bin/bash
# round 2, call 49: b200_w4a16_gemm_noalloc (scratch built into the library): its test alone in a
# fresh process, then the whole GPU suite (it is collected last)
set -x
O=gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/r3w_noalloc_alone.log 2>&1; tail -4 $O/r3w_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/r3w_gpu_tests.log 2>&1; tail -4 $O/r3w_gpu_tests.log
