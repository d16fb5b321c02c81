#!/bin/bash
set -x
O=gpurun_out/job45; mkdir -p $O
for ZP in 0 d; do
  if [ $ZP = d ]; then unset KGP_DIST_ZPUSH; else export KGP_DIST_ZPUSH=$ZP; fi
  ( timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2974$( [ $ZP = d ] && echo 1 || echo 0 ) scripts/dist_exact_steps.py 40000 6 ) > $O/steps_$ZP.log 2>&1; echo "rc=$?" >> $O/steps_$ZP.log
done
