#!/bin/bash
set -x
O=gpurun_out/job28; mkdir -p $O
KGP_SMALL_NP=8192 timeout 300 python scripts/time_small.py > $O/small_np8192.log 2>&1
timeout 300 python scripts/time_small.py > $O/small.log 2>&1
( time timeout 1200 python bench.py --steps 5 --warmup 3 ) > $O/bench_n1.log 2>&1; echo "rc=$?" >> $O/bench_n1.log
