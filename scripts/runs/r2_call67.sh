#!/bin/bash
# round 2, call 67: MoE tests after factoring the path policy into MixtralMoE.select_path
timeout 200 python -m pytest tests/test_moe.py -m gpu -x -q 2>&1 | tail -2
