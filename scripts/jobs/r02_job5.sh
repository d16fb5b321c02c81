#!/bin/bash
set -x
O=gpurun_out/job5; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_exact_classes.py 40000 > $O/classes.log 2>&1
KGP_LOOKAHEAD=1 timeout 300 python scripts/time_exact.py 40000 > $O/time_la1.log 2>&1
KGP_LOOKAHEAD=1 KGP_TILE_REDUCE=1 timeout 300 python scripts/time_exact.py 40000 > $O/time_la1_tr1.log 2>&1
