"""torchrun: per-step device time (every rank) and wall time of back-to-back sharded exact evaluations at n (default 40000)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch, torch.distributed as td
from kgp_b200 import B200GP, dist

rank, world, local = dist.env_rank()
torch.cuda.set_device(local)
td.init_process_group('nccl', device_id=torch.device('cuda', local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40000; K = int(sys.argv[2]) if len(sys.argv) > 2 else 6
rng = np.random.RandomState(43); D = 8
X = rng.randn(n, D); y = np.sin(X[:, 0] * X[:, 1]) + X[:, 2] ** 2 - X[:, 3]; y = (y - y.mean()) / y.std() + 0.1 * rng.randn(n)
hyp = {'lik': np.log(0.1), 'mean': [], 'cov': np.concatenate([np.log(2.0) * np.ones(D), [0.0]])[:, None]}
gp = B200GP(engine_kwargs={'device': local})
gp.configure(D, dict(hyp), {}, 'infExact', 'meanZero', 'covSEard', 'likGauss', 'dlikExact', verbose=0)
dist.attach(gp); gp.update_data('tr', X, y[:, None])
for _ in range(3):
    gp.eng.eval(0, 2, gp._n_hyp)
td.barrier(); torch.cuda.synchronize()
rows = []; t0 = time.perf_counter()
for _ in range(K):
    ta = time.perf_counter(); gp.eng.eval(0, 2, gp._n_hyp); tb = time.perf_counter()
    st = gp.stats(); rows.append((st['ms_total'], st['ms_kbuild'], st['ms_potrf'], st['ms_inverse'], st['ms_solve'], st['ms_reduce'], (tb - ta) * 1e3))
td.barrier(); torch.cuda.synchronize(); wall = (time.perf_counter() - t0) / K * 1e3
for r in range(world):
    td.barrier()
    if r == rank:
        print('rank %d wall/step %.1f ms | per step (dev total, kbuild, potrf, inverse, solve, reduce, host wall): %s' % (
            rank, wall, ' '.join('(%.1f %.1f %.1f %.1f %.1f %.1f %.1f)' % t for t in rows)), flush=True)
td.destroy_process_group()
