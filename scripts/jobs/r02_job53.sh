#!/bin/bash
set -x
O=gpurun_out/job53; mkdir -p $O
( time timeout 40 python __graft_entry__.py smoke ) > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
( time timeout 60 python -m pytest tests/test_exact_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "parity_vs_oracle and not n16384 or no_alloc" ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
