#!/bin/bash
set -x
O=gpurun_out/job8; mkdir -p $O
KGP_DEBUG=1 MAXIT=300 timeout 300 python scripts/time_grid_k1000.py 1000000 1000 > $O/k1000_maxit300.log 2>&1
KGP_DEBUG=1 MAXIT=300 timeout 300 python scripts/time_grid_k1000.py 1000000 256 > $O/k256_maxit300.log 2>&1
