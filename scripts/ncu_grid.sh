#!/bin/bash
# ncu --set full capture of the persistent CG kernel at n=$1 (one launch = one whole solve)
n=${1:-1000000}
MAXIT=30000 ncu --set full --clock-control none --import-source on -k regex:cg_persistent2 -c 1 -o gpurun_out/cgp_$n -f python scripts/time_grid.py $n > gpurun_out/ncu_grid_$n.log 2>&1
ncu -i gpurun_out/cgp_$n.ncu-rep --page raw --csv > gpurun_out/cgp_${n}_raw.csv 2>/dev/null
ncu -i gpurun_out/cgp_$n.ncu-rep --page source --csv --print-source cuda > gpurun_out/cgp_${n}_src.csv 2>/dev/null
