#!/bin/bash
set -x
O=gpurun_out/job29; mkdir -p $O
timeout 1500 python -m pytest tests/test_grid_gpu.py -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
MAXIT=20000 KGP_DEBUG=1 timeout 300 python scripts/time_grid.py 1000000 > $O/grid1m_fx.log 2>&1
