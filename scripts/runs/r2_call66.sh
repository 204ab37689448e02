#!/bin/bash
# round 2, call 66: the committed final state as the driver runs it: full GPU suite, smoke, default bench line
set -x
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r4n_gpu_tests.log 2>&1; tail -2 $O/r4n_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r4n_smoke.log 2>&1; tail -1 $O/r4n_smoke.log
python bench.py --no-extras > $O/r4n_bench_default.json 2>/dev/null; cut -c1-260 $O/r4n_bench_default.json
