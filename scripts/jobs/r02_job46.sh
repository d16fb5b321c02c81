#!/bin/bash
set -x
O=gpurun_out/job46; mkdir -p $O
( timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29751 scripts/dist_exact_steps.py 40000 10 ) > $O/steps_d.log 2>&1; echo "rc=$?" >> $O/steps_d.log
