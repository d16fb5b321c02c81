#!/bin/bash
# 8 GPUs: the sharded exact evaluation at n=40000 with look-ahead 1 (default) and look-ahead 2 (KGP_DIST_LA2=1)
set -x
O=gpurun_out/job40; mkdir -p $O
export KGP_DIST_NOTPD=0
for LA in 0 1; do
  ( KGP_DIST_LA2=$LA timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port $((29700+LA)) scripts/dist_exact.py 40000 ) > $O/la$LA.log 2>&1; echo "rc=$?" >> $O/la$LA.log
done
