#!/bin/bash
set -x
O=gpurun_out/job26; mkdir -p $O
timeout 1500 python -m pytest tests/test_grid_gpu.py -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/small_launches_c4.py > $O/c4.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c4_launches.csv python scripts/small_launches_c4.py > $O/ncu_c4.log 2>&1
