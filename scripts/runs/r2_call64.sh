#!/bin/bash
# round 2, call 64 (2 GPUs): the multi-rank all-reduce tests after the dispatch change at TP 2
set -x
timeout 300 python -m pytest tests/test_gpu_allreduce.py -m gpu -x -q -k test_multi_gpu_allreduce_matches_nccl > gpurun_out/r4l_mgpu_tests.log 2>&1; tail -3 gpurun_out/r4l_mgpu_tests.log
