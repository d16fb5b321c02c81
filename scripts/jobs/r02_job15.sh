#!/bin/bash
set -x
O=gpurun_out/job15; mkdir -p $O
timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_dist_gpu.py -m gpu -q -x -k "not n16384" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_small.py > $O/small.log 2>&1
timeout 300 python scripts/time_exact.py 40000 > $O/time_exact.log 2>&1
for t in 576 640; do
  KGP_B200_LIB=$PWD/build/libkgp_t$t.so MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 40000 > $O/grid_t$t.log 2>&1
done
MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 40000 > $O/grid_t512.log 2>&1
