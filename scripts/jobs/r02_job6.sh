#!/bin/bash
set -x
O=gpurun_out/job6; mkdir -p $O
timeout 1700 python -m pytest tests -m gpu -q --durations=15 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
