#!/bin/bash
set -x
O=gpurun_out/job48; mkdir -p $O
( time timeout 1200 python -m pytest tests -m gpu -q -x --durations=8 ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
( time timeout 600 python bench.py ) > $O/bench_n1.log 2>&1; echo "rc=$?" >> $O/bench_n1.log
( time timeout 200 python __graft_entry__.py smoke ) > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
