#!/bin/bash
set -x
O=gpurun_out/job35; mkdir -p $O
cat > /tmp/probe1.py <<'PY'
import sys, os, json
sys.path.insert(0, os.environ.get('GRAFT_REPO_ROOT', '/root/repo'))
import bench
r = bench.kissgp_probe_leg(1000000, 0, 1, 1, 1)
print(r['ms_per_eval'])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --launch-skip 1200 -c 2500 --csv --log-file $O/probe_launches.csv python /tmp/probe1.py > $O/ncu_probe.log 2>&1
