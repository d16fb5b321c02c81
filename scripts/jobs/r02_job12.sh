#!/bin/bash
set -x
O=gpurun_out/job12; mkdir -p $O
timeout 900 python -m pytest tests/test_grid_gpu.py -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
KGP_DEBUG=1 MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 > $O/dbg.log 2>&1
MAXIT=30000 timeout 300 python scripts/time_grid.py 40000 100000 > $O/small.log 2>&1
