#!/bin/bash
# 4 GPUs: Z all-gather by grouped broadcasts (KGP_DIST_ZPUSH=0) against pipelined peer copies (default for P > 2), n = 40000 and a ragged size
set -x
O=gpurun_out/job43; mkdir -p $O
export KGP_DIST_NOTPD=0
( KGP_DIST_ZPUSH=0 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29730 scripts/dist_exact.py 40000 ) > $O/zpush0.log 2>&1; echo "rc=$?" >> $O/zpush0.log
( timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29731 scripts/dist_exact.py 5555 40000 ) > $O/zpush1.log 2>&1; echo "rc=$?" >> $O/zpush1.log
( KGP_DIST_ZPUSH=1 timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29732 scripts/dist_exact.py 40000 ) > $O/n2_zpush1.log 2>&1; echo "rc=$?" >> $O/n2_zpush1.log
