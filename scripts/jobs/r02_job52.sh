#!/bin/bash
set -x
O=gpurun_out/job52; mkdir -p $O
( time timeout 100 python -m pytest tests/test_grid_gpu.py tests/test_zz_no_alloc.py -m gpu -q -x -k "interpolation_degree or deterministic_parts or kmeans or woodbury or no_alloc" ) > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
( time timeout 40 python __graft_entry__.py smoke ) > $O/smoke.log 2>&1; echo "rc=$?" >> $O/smoke.log
