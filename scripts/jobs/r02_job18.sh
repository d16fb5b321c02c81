#!/bin/bash
# 2 GPUs: new potrf_diag128 (panel warp), pipelined Z all-gather, unpack off the panel chain
set -x
O=gpurun_out/job18; mkdir -p $O
timeout 900 python -m pytest tests/test_exact_gpu.py tests/test_dist_gpu.py -m gpu -q -x -k "not n16384" > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python scripts/time_small.py > $O/small.log 2>&1
timeout 300 python scripts/time_exact.py 40000 > $O/time_exact.log 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --steps 3 --warmup 3 --no-kissgp > $O/bench_n2.log 2>&1; echo "rc=$?" >> $O/bench_n2.log
KGP_DIST_ZPIPE=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --steps 3 --warmup 3 --no-kissgp > $O/bench_n2_nozpipe.log 2>&1; echo "rc=$?" >> $O/bench_n2_nozpipe.log
