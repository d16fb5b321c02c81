#!/bin/bash
set -x
O=gpurun_out/job16; mkdir -p $O
python scripts/time_small.py > $O/small_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 200 --csv --log-file $O/launches_small.csv python scripts/time_small.py > $O/ncu_small.log 2>&1
timeout 600 python -m pytest tests/test_exact_gpu.py tests/test_dist_gpu.py -m gpu -q -x -k "not n16384" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
