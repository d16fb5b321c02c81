#!/bin/bash
set -x
O=gpurun_out/job20; mkdir -p $O
timeout 1200 python -m pytest tests/test_grid_gpu.py tests/test_cabi_cpu.py -q -x -k "not c5_shape" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
