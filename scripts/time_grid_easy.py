"""A well-conditioned KISS-GP problem (uniform points, covSEiso, few dozen CG iterations): fixed costs of a solve."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
rng = np.random.RandomState(44)
H = rng.rand(n, 2); y = np.sin(3 * H[:, 0]) * np.cos(2 * H[:, 1]) + 0.3 * rng.randn(n)
gp = B200GP()
gp.configure(2, {'lik': np.log(0.3), 'mean': [], 'cov': np.array([[0.5], [1.0]])}, {'cg_maxit': 40000, 'cg_tol': 1e-6}, 'infGrid', 'meanZero', 'covSEiso',
             'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, verbose=0)
gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
for rep in range(4):
    t0 = time.perf_counter(); nlZ, g = gp.eng.eval(0, 2, gp._n_hyp); w = time.perf_counter() - t0
    st = gp.stats()
    print('rep %d wall %.2f ms dev %.2f (setup %.2f solve %.2f rest %.2f) iters %d' % (rep, 1e3 * w, st['ms_total'], st['ms_kbuild'], st['ms_solve'], st['ms_reduce'], st['cg_iters']), flush=True)
