#!/bin/bash
set -x
O=gpurun_out/job9; mkdir -p $O
MAXIT=30000 timeout 300 python scripts/time_grid.py 1000000 > $O/time_grid_plain.log 2>&1 && \
MAXIT=30000 ncu --set full --clock-control none --import-source on -k regex:cg_persistent2 -c 1 -o $O/cgp -f python scripts/time_grid.py 1000000 > $O/ncu.log 2>&1
ncu -i $O/cgp.ncu-rep --page raw --csv > $O/cgp_raw.csv 2>/dev/null
ncu -i $O/cgp.ncu-rep --page source --csv --print-source sass > $O/cgp_sass.csv 2>/dev/null
ncu -i $O/cgp.ncu-rep --page source --csv --print-source cuda > $O/cgp_cuda.csv 2>/dev/null
rm -f $O/cgp.ncu-rep
ls -la $O
