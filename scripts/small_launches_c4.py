"""Evaluations of C4 (KISS-GP n=460, 1-D grid k=100, cg 500/1e-4) for an ncu launch list: the latency-bound small grid case."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
rng = np.random.RandomState(45)
H = np.tanh(rng.randn(460, 1)); y = np.sin(4 * H[:, 0]) + np.exp(-2) * rng.randn(460)
gp = B200GP()
gp.configure(1, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, {'cg_maxit': 500, 'cg_tol': 1e-4}, 'infGrid', 'meanZero', 'covSEiso', 'likGauss',
             'dlikGrid', grid_kwargs={'eq': 1, 'k': 100.}, verbose=0)
gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
v0 = gp._hyp_vec().copy()
for i in range(3):
    gp._set_hyp_vec(v0 + 2e-3 * (1 if i % 2 else -1)); gp._push_model()
    nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
print(nlZ, gp.stats())
