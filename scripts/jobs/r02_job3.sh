#!/bin/bash
# round-2 job 3: TMA GEMM variants (BM=128 one CTA/SM, BM=64 two CTAs/SM), no-load ceiling, ncu source-level profile
set -x
O=gpurun_out/job3; mkdir -p $O
timeout 120 lab/gemm_lab quick > $O/lab_quick.log 2>&1; echo "rc=$?" >> $O/lab_quick.log
timeout 600 lab/gemm_lab > $O/lab_full.log 2>&1; echo "rc=$?" >> $O/lab_full.log
for bm in 128 64; do
  KGP_TG_BM=$bm timeout 300 python scripts/time_exact_classes.py 40000 > $O/classes_bm$bm.log 2>&1
done
timeout 120 lab/gemm_lab prof 128 > $O/prof_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:dgemm_nt_tma -s 1 -c 1 -o $O/tma128 -f lab/gemm_lab prof 128 > $O/ncu_tma128.log 2>&1
ncu -i $O/tma128.ncu-rep --page raw --csv > $O/tma128_raw.csv 2>/dev/null
ncu -i $O/tma128.ncu-rep --page source --csv --print-source sass > $O/tma128_sass.csv 2>/dev/null
ls -la $O
