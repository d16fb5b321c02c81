#!/bin/bash
# 2 GPUs: the peer-copy all-gather of Z (KGP_DIST_ZPUSH=1) -- correctness against the unsharded evaluation at three sizes, timing
# against the default of P = 2 (pipelined broadcasts under a staged Z Z^T)
set -x
O=gpurun_out/job41; mkdir -p $O
( KGP_DEBUG=1 KGP_DIST_ZPUSH=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 scripts/dist_exact.py 3000 7777 40000 ) > $O/zpush1.log 2>&1; echo "rc=$?" >> $O/zpush1.log
( KGP_DIST_NOTPD=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 scripts/dist_exact.py 40000 ) > $O/default.log 2>&1; echo "rc=$?" >> $O/default.log
