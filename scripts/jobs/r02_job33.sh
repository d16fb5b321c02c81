#!/bin/bash
set -x
O=gpurun_out/job33; mkdir -p $O
timeout 1500 python -m pytest tests/test_grid_gpu.py -m gpu -q -x -k "cheby or slq or pred_var" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_probe_leg.py > $O/probe_sorted.log 2>&1
KGP_CHEBY_SORTED=0 timeout 300 python scripts/time_probe_leg.py > $O/probe_unsorted.log 2>&1
