#!/bin/bash
set -x
O=gpurun_out/job50; mkdir -p $O
( time timeout 300 python -m pytest tests/test_grid_gpu.py -m gpu -q -x -k "interpolation_degree or kmeans or non_equispaced or deterministic_parts or iterative_parts or learned_projection or toeplitz" ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
