#!/bin/bash
# round-2 job 1: GPU tests, timings of the already-built alternatives, sanitizer evidence
set -x
O=gpurun_out/job1; mkdir -p $O
nvidia-smi -L > $O/smi.txt; nproc >> $O/smi.txt; lscpu | head -20 >> $O/smi.txt; free -g >> $O/smi.txt
python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
for la in 0 1; do for tr in 0 1; do
  KGP_LOOKAHEAD=$la KGP_TILE_REDUCE=$tr python scripts/time_exact.py 40000 > $O/time_exact_la${la}_tr${tr}.log 2>&1
done; done
KGP_LOOKAHEAD=1 python scripts/time_exact_classes.py 40000 > $O/classes_la1.log 2>&1
python scripts/time_exact_classes.py 40000 > $O/classes_la0.log 2>&1
python scripts/gemm_bench.py > $O/gemm_bench.log 2>&1
# sanitizers on the smoke-sized problems
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/san_small.py exact > $O/san_memcheck_exact.log 2>&1; echo "rc=$?" >> $O/san_memcheck_exact.log
for m in 3 2 1; do
  KGP_CG_MODE=$m timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python scripts/san_small.py grid > $O/san_memcheck_grid_m$m.log 2>&1; echo "rc=$?" >> $O/san_memcheck_grid_m$m.log
done
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/san_small.py exact > $O/san_racecheck_exact.log 2>&1; echo "rc=$?" >> $O/san_racecheck_exact.log
for m in 3 2 1; do
  KGP_CG_MODE=$m timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python scripts/san_small.py grid > $O/san_racecheck_grid_m$m.log 2>&1; echo "rc=$?" >> $O/san_racecheck_grid_m$m.log
done
timeout 600 compute-sanitizer --tool synccheck --error-exitcode 9 python scripts/san_small.py all > $O/san_synccheck.log 2>&1; echo "rc=$?" >> $O/san_synccheck.log
ls -la $O
