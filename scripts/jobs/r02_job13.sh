#!/bin/bash
set -x
O=gpurun_out/job13; mkdir -p $O
timeout 120 lab/gemm_lab quick > $O/lab_quick.log 2>&1; echo "rc=$?" >> $O/lab_quick.log
timeout 600 lab/gemm_lab > $O/lab_full.log 2>&1
timeout 900 python -m pytest tests/test_grid_gpu.py tests/test_exact_gpu.py -m gpu -q -x -k "not n16384 and not c5_shape" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_exact_classes.py 40000 > $O/classes.log 2>&1
timeout 300 python scripts/time_exact.py 40000 > $O/time_exact.log 2>&1
KGP_DEBUG=1 MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 > $O/grid_fx.log 2>&1
