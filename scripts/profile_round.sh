#!/bin/bash
# One-GPU evidence run for profiles/: bench line, ncu launch lists (exact step, KISS-GP eval), ncu --set full of the two dominant kernels.
set -x
python bench.py --steps 3 --warmup 3 > gpurun_out/bench_n1.log 2>&1 && tail -1 gpurun_out/bench_n1.log > gpurun_out/bench_n1.json
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_exact.csv python bench.py --steps 1 --warmup 1 --no-cpu --no-kissgp > gpurun_out/ncu_exact.log 2>&1
MAXIT=30000 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_grid.csv python scripts/time_grid.py 1000000 > gpurun_out/ncu_gridl.log 2>&1
scripts/ncu_grid.sh 1000000
ncu --set full --clock-control none --import-source on -k regex:dgemm_nt -s 3 -c 5 -o gpurun_out/dgemm_full -f python bench.py --steps 1 --warmup 1 --no-cpu --no-kissgp > gpurun_out/ncu_dgemm.log 2>&1
ncu -i gpurun_out/dgemm_full.ncu-rep --page raw --csv > gpurun_out/dgemm_full_raw.csv 2>/dev/null
ls -la gpurun_out
