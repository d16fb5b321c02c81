#!/bin/bash
set -x
O=gpurun_out/job30; mkdir -p $O
timeout 900 python -m pytest tests/test_grid_gpu.py -m gpu -q -x -k "projection or non_equi or deterministic_parts" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
