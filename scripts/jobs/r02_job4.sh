#!/bin/bash
set -x
O=gpurun_out/job4; mkdir -p $O
timeout 120 lab/gemm_lab quick > $O/lab_quick.log 2>&1; echo "rc=$?" >> $O/lab_quick.log
timeout 600 lab/gemm_lab > $O/lab_full.log 2>&1; echo "rc=$?" >> $O/lab_full.log
