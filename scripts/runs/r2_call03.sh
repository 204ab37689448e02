#!/bin/bash
set -x
timeout 120 python scripts/debug_determinism.py 16 4096 4096 > gpurun_out/r2c_det.log 2>&1
timeout 120 python scripts/debug_determinism.py 8 4096 4096 >> gpurun_out/r2c_det.log 2>&1
timeout 120 python scripts/trace_decode.py --m 16 --k 4096 --n 4096 > gpurun_out/r2c_trace_decode_4096.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 8192 --n 57344 --ctas 0,73 > gpurun_out/r2c_trace_tcdec_57344.log 2>&1
timeout 120 python scripts/trace_gemm.py --kernel tcdecode --m 16 --k 4096 --n 4096 --ctas 0,73 > gpurun_out/r2c_trace_tcdec_4096.log 2>&1
timeout 600 python -m pytest tests -m gpu -q --timeout 180 -k "not llama13b_prefill and not deterministic" > gpurun_out/r2c_gputests.log 2>&1
tail -5 gpurun_out/r2c_gputests.log
