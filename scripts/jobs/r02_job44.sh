#!/bin/bash
# 2 GPUs: the world-2 tests (both gather modes of Z, sharded probes, row-sharded CG) and the full N=2 bench line
set -x
O=gpurun_out/job44; mkdir -p $O
( time timeout 500 python -m pytest tests/test_dist_gpu.py -m gpu -q -x ) > $O/pytest_dist.log 2>&1; echo "pytest rc=$?" >> $O/pytest_dist.log
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 ) > $O/bench_n2.log 2>&1; echo "rc=$?" >> $O/bench_n2.log
