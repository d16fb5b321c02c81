#!/bin/bash
# 2-GPU validation: world-2 tests (sharded exact, NOT_PD status on every rank, sharded probes, row-sharded CG) + the N=2 bench line
set -x
O=gpurun_out/job14; mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 1200 python -m pytest tests/test_dist_gpu.py -m gpu -q -x > $O/pytest_dist.log 2>&1; echo "pytest rc=$?" >> $O/pytest_dist.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 3 --warmup 3 > $O/bench_n2.log 2>&1; echo "rc=$?" >> $O/bench_n2.log
tail -2 $O/bench_n2.log | head -1 > $O/bench_n2.json
