"""torchrun check of KISS-GP probe sharding: opt.ldB2_cheby with the Hutchinson probes split over the ranks
(each rank runs its share of the Krylov recurrences, one all-reduce of the grid-sized accumulators) must equal
the single-GPU evaluation with the same probes."""
import os, sys
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
rng = np.random.RandomState(44)
H = np.maximum(0, rng.normal(0.5, 1, (n, 2))); y = np.sin(3 * H[:, 0]) * np.cos(2 * H[:, 1]) + 0.3 * rng.randn(n)
V = np.sign(rng.randn(n, 10)); V[V == 0] = 1.0
opt = {'cg_maxit': 5000, 'cg_tol': 1e-10, 'ldB2_cheby': 1, 'ldB2_cheby_degree': 15, 'ldB2_cheby_hutch': 10}
def mk(attach):
    gp = B200GP(engine_kwargs={'device': local})
    gp.configure(2, {'lik': np.log(0.3), 'mean': [], 'cov': np.array([[0.5], [1.0]])}, opt, 'infGrid', 'meanZero', 'covSEiso', 'likGauss',
                 'dlikGrid', grid_kwargs={'eq': 1, 'k': 40.}, verbose=0)
    if attach: dist.attach(gp)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    gp.eng.set_probes(V, 1.0 + n * np.exp(2.0) / 0.09)
    return gp
gpd = mk(True)
nlZ, g = gpd.eng.eval(0, 2, gpd._n_hyp)
if rank == 0:
    gps = mk(False)
    nlZ1, g1 = gps.eng.eval(0, 2, gps._n_hyp)
    print('n=%d world=%d cheby probes sharded: nlZ %.12g vs %.12g rel %.2e dnlZ %.2e' % (n, world, nlZ, nlZ1, rel(nlZ, nlZ1), rel(g, g1)), flush=True)
td.destroy_process_group()
