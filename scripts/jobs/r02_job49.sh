#!/bin/bash
set -x
O=gpurun_out/job49; mkdir -p $O
( time timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29770 bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu --no-kissgp ) > $O/bench_n4.log 2>&1; echo "rc=$?" >> $O/bench_n4.log
