#!/bin/bash
set -x
O=gpurun_out/job17; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 300 python scripts/time_small.py > $O/small.log 2>&1
timeout 1200 python bench.py --steps 5 --warmup 3 > $O/bench_n1.log 2>&1; echo "rc=$?" >> $O/bench_n1.log
timeout 1500 python scripts/cpu_direct.py 32768 > $O/cpu_direct_32768.log 2>&1
