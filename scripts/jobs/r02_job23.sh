#!/bin/bash
set -x
O=gpurun_out/job23; mkdir -p $O
timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_harness_gpu.py -m gpu -q -x > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_small.py > $O/small.log 2>&1
KGP_SMALL_NP=0 KGP_SMALL_TRSM=0 timeout 300 python scripts/time_small.py > $O/small_off.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/small512_launches.csv python scripts/small_launches.py 512 > $O/ncu512.log 2>&1
