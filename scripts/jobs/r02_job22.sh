#!/bin/bash
set -x
O=gpurun_out/job22; mkdir -p $O
timeout 300 python scripts/small_launches.py 512 > $O/small512.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/small512_launches.csv python scripts/small_launches.py 512 > $O/ncu512.log 2>&1
