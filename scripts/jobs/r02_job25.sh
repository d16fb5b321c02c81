#!/bin/bash
set -x
O=gpurun_out/job25; mkdir -p $O
timeout 300 python scripts/small_launches_c4.py > $O/c4.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/c4_launches.csv python scripts/small_launches_c4.py > $O/ncu_c4.log 2>&1
