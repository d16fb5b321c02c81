#!/bin/bash
set -x
O=gpurun_out/job10; mkdir -p $O
timeout 1500 python -m pytest tests/test_grid_gpu.py tests/test_harness_gpu.py tests/test_dist_gpu.py -m gpu -q --durations=8 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 > $O/time_grid_1m.log 2>&1
MAXIT=300 timeout 300 python scripts/time_grid_k1000.py 1000000 1000 > $O/k1000_maxit300.log 2>&1
MAXIT=300 KGP_KRON_GEMM_MIN=100000 timeout 300 python scripts/time_grid_k1000.py 1000000 1000 > $O/k1000_maxit300_nogemm.log 2>&1
MAXIT=60000 timeout 600 python scripts/time_grid_k1000.py 1000000 1000 > $O/k1000_full.log 2>&1
