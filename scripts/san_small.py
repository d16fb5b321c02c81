"""Smoke-sized exact + KISS-GP evaluations for compute-sanitizer (memcheck / racecheck); no oracle, no timing."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
which = sys.argv[1] if len(sys.argv) > 1 else 'all'
rng = np.random.RandomState(42)
if which in ('all', 'exact'):
    n, D = 300, 3
    X = rng.randn(n, D); y = np.sin(X.sum(1)) + 0.1 * rng.randn(n)
    gp = B200GP()
    gp.configure(D, {'lik': -2.0, 'mean': [], 'cov': np.array([[-0.7], [0.0]])}, {}, 'infExact', 'meanZero', 'covSEiso', 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    nlZ, g = gp.eng.eval(0, 2, 3); dX = gp.get_dlik_dx('tr'); ymu, ys2 = gp.predict(X[:50], return_var=True)
    print('exact', nlZ, g, float(np.abs(dX).sum()), float(ymu.sum()), flush=True)
    del gp
if which in ('all', 'grid'):
    n = 800; H = rng.rand(n, 2); y2 = np.sin(4 * H[:, 0]) + np.cos(3 * H[:, 1]) + 0.1 * rng.randn(n)
    gk = B200GP()
    gk.configure(2, {'lik': np.log(0.2), 'mean': [], 'cov': [[np.log(0.3)], [0.0]]}, {'cg_maxit': 400, 'cg_tol': 1e-8}, 'infGrid', 'meanZero', 'covSEiso',
                 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 12.}, verbose=0)
    gk.update_data('tr', H, y2[:, None]); gk.update_grid('tr')
    nlZ, g = gk.eng.eval(0, 2, gk._n_hyp); dX = gk.get_dlik_dx('tr'); ymu = gk.predict(H[:40])
    print('grid mode', os.environ.get('KGP_CG_MODE', '3'), nlZ, g, float(np.abs(dX).sum()), float(ymu.sum()), gk.stats()['cg_iters'], flush=True)
