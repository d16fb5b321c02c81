#!/bin/bash
set -x
O=gpurun_out/job39; mkdir -p $O
( time timeout 600 python -m pytest tests/test_grid_gpu.py -m gpu -q -x -k "kmeans or non_equispaced or bitexact or deterministic_parts" ) > $O/pytest_kmeans.log 2>&1; echo "pytest rc=$?" >> $O/pytest_kmeans.log
