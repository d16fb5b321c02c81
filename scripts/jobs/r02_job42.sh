#!/bin/bash
# 8 GPUs: all-gather of Z by grouped broadcasts (KGP_DIST_ZPUSH=0) against pipelined peer copies (default for P > 2), n = 40000
set -x
O=gpurun_out/job42; mkdir -p $O
export KGP_DIST_NOTPD=0
for ZP in 0 1; do
  ( KGP_DEBUG=1 KGP_DIST_ZPUSH=$ZP timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29720+ZP)) scripts/dist_exact.py 40000 ) > $O/zpush$ZP.log 2>&1; echo "rc=$?" >> $O/zpush$ZP.log
done
