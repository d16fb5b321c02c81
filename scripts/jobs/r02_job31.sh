#!/bin/bash
set -x
O=gpurun_out/job31; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -x > $O/pytest_alone.log 2>&1; echo "pytest rc=$?" >> $O/pytest_alone.log
timeout 1800 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest rc=$?" >> $O/pytest_all.log
