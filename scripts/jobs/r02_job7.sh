#!/bin/bash
set -x
O=gpurun_out/job7; mkdir -p $O
timeout 1200 python -m pytest tests/test_grid_gpu.py tests/test_harness_gpu.py -m gpu -q -k "toeplitz or harness or epoch or c4 or c5_sm" --durations=10 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
