"""torchrun check + timing of the ROW-SHARDED KISS-GP solve (SURVEY 8(e) KISS-GP (ii)): every rank owns n/P of the
points; the two reductions of each CG iteration cross the ranks over peer memory (CUDA IPC / NVLink) inside the
persistent kernel.  Result must equal the single-GPU evaluation to the CG tolerance."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch, torch.distributed as td
from kgp_b200 import B200GP, dist

rank, world, local = dist.env_rank()
torch.cuda.set_device(local)
td.init_process_group('nccl', device_id=torch.device('cuda', local))
def rel(a, b): return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
k = float(sys.argv[2]) if len(sys.argv) > 2 else 40.
tol = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-10
rng = np.random.RandomState(44)
uniform = len(sys.argv) > 4 and sys.argv[4] == 'uniform'
H = rng.rand(n, 2) if uniform else np.maximum(0, rng.normal(0.5, 1, (n, 2)))      # uniform (C5-like) or ReLU-like with exact zeros (C3)
y = np.sin(3 * H[:, 0]) * np.cos(2 * H[:, 1]) + 0.3 * rng.randn(n)
opt = {'cg_maxit': 40000, 'cg_tol': tol}
c5 = len(sys.argv) > 4 and sys.argv[4] == 'c5'           # the bench's KISS-GP leg: covSM(Q=6) x 2 + meanSum, thousands of iterations
if c5:
    import bench
    H, y, hc5 = bench.synth_c5(n)
def mk(attach):
    gp = B200GP(engine_kwargs={'device': local})
    if c5:
        gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': [list(a) for a in hc5]}, opt, 'infGrid', 'meanSum', 'covSM', 'likGauss',
                     'dlikGrid', grid_kwargs={'eq': 1, 'k': k}, cov_args=[6], mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    else:
        gp.configure(2, {'lik': np.log(0.3), 'mean': [], 'cov': np.array([[0.5], [1.0]])}, opt, 'infGrid', 'meanZero', 'covSEiso', 'likGauss',
                     'dlikGrid', grid_kwargs={'eq': 1, 'k': k}, verbose=0)
    if attach: dist.attach(gp)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    return gp
gpd = mk(True)
for rep in range(3):
    td.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    nlZ, g = gpd.eng.eval(0, 2, gpd._n_hyp)
    wall = time.perf_counter() - t0
    st = gpd.stats()
dX = gpd.get_dlik_dx('tr'); al = gpd.alpha()
res = torch.tensor([nlZ], dtype=torch.float64, device='cuda'); lo = res.clone(); hi = res.clone()
td.all_reduce(lo, op=td.ReduceOp.MIN); td.all_reduce(hi, op=td.ReduceOp.MAX)
if rank == 0:
    gps = mk(False)
    for rep in range(2):
        nlZ1, g1 = gps.eng.eval(0, 2, gps._n_hyp)
    st1 = gps.stats(); dX1 = gps.get_dlik_dx('tr'); al1 = gps.alpha()
    print('n=%d world=%d row-sharded CG: rel nlZ %.2e dnlZ %.2e dX %.2e alpha %.2e ranks_agree %.1e | iters %d vs %d | solve %.2f ms (%.2f us/it) vs %.2f ms (%.2f us/it) | eval %.2f vs %.2f ms'
          % (n, world, rel(nlZ, nlZ1), rel(g, g1), rel(dX, dX1), rel(al, al1), float(hi - lo) / abs(nlZ), st['cg_iters'], st1['cg_iters'],
             st['ms_solve'], 1e3 * st['ms_solve'] / max(st['cg_iters'] - 1, 1), st1['ms_solve'], 1e3 * st1['ms_solve'] / max(st1['cg_iters'] - 1, 1),
             st['ms_total'], st1['ms_total']), flush=True)
td.destroy_process_group()
