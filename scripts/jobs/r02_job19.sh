#!/bin/bash
# 8 GPUs: exact-path strong scaling line (no KISS-GP legs), with and without the pipelined Z all-gather
set -x
O=gpurun_out/job19; mkdir -p $O
nvidia-smi -L > $O/smi.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 4 --warmup 3 --no-kissgp --no-cpu > $O/bench_n8.log 2>&1; echo "rc=$?" >> $O/bench_n8.log
KGP_DIST_ZPIPE=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29622 bench.py --gpus 8 --steps 4 --warmup 3 --no-kissgp --no-cpu > $O/bench_n8_nozpipe.log 2>&1; echo "rc=$?" >> $O/bench_n8_nozpipe.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29623 bench.py --gpus 4 --steps 4 --warmup 3 --no-kissgp --no-cpu > $O/bench_n4.log 2>&1; echo "rc=$?" >> $O/bench_n4.log
