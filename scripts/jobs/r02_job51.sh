#!/bin/bash
set -x
O=gpurun_out/job51; mkdir -p $O
( time timeout 400 python -m pytest tests -m gpu -q -x ) > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
( time timeout 100 python __graft_entry__.py smoke ) > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
