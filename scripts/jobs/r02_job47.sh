#!/bin/bash
# 8 GPUs, eight back-to-back evaluations per setting: steady-state time per step with the broadcast gather and the peer-copy gather
set -x
O=gpurun_out/job47; mkdir -p $O
( KGP_DIST_ZPUSH=0 timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29760 scripts/dist_exact_steps.py 40000 8 ) > $O/steps_0.log 2>&1; echo "rc=$?" >> $O/steps_0.log
( timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29761 scripts/dist_exact_steps.py 40000 8 ) > $O/steps_d.log 2>&1; echo "rc=$?" >> $O/steps_d.log
