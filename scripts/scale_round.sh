#!/bin/bash
# bench.py at N = 1, 2, 4, 8 on one box (the driver's SCALE run), lines collected in gpurun_out/scale.jsonl
: > gpurun_out/scale.jsonl
python bench.py --gpus 1 --steps 3 --warmup 3 --no-cpu --no-kissgp 2>/dev/null | tail -1 >> gpurun_out/scale.jsonl
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600+N)) bench.py --gpus $N --steps 3 --warmup 3 --no-cpu 2>/dev/null | tail -1 >> gpurun_out/scale.jsonl
done
python - <<'PY'
import json
for l in open('gpurun_out/scale.jsonl'):
    d = json.loads(l); print(d['n_gpus'], d['scaling'], 'value %.4f e2e %.4f ms %.1f gemm frac %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac']))
PY
