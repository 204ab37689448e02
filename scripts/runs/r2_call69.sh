#!/bin/bash
# round 2, call 69: the reference's op signatures through the pybind shim of INTEGRATION.md section A, then the
# no-allocation test behind it (collection order), in one process; exit code of the process matters (statics)
timeout 200 python -m pytest tests/test_zx_integration_shim.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r4p_shim_tests.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r4p_shim_tests.log
