#!/bin/bash
# Round-2 one-GPU evidence run for profiles/: bench line, reference-arm line, ncu launch list of one exact step, ncu --set full of the GEMM
set -x
O=gpurun_out/r02; mkdir -p $O
python bench.py --steps 5 --warmup 3 > $O/bench_n1.log 2>&1; tail -1 $O/bench_n1.log > $O/bench_n1.json
python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_ref.log 2>&1; tail -1 $O/bench_ref.log > $O/bench_ref.json
python scripts/time_exact.py 40000 > $O/time_exact_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/launches_exact.csv python scripts/time_exact.py 40000 > $O/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:dgemm_nt_tma -s 3 -c 6 -o $O/dgemm_full -f python scripts/time_exact.py 40000 > $O/ncu_dgemm.log 2>&1
ncu -i $O/dgemm_full.ncu-rep --page raw --csv > $O/dgemm_full_raw.csv 2>/dev/null
rm -f $O/dgemm_full.ncu-rep
ls -la $O
