"""Two evaluations of C1 (exact n=512 d=32) for an ncu launch list: which kernels make up the latency-bound small case."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
n, D = (int(sys.argv[1]) if len(sys.argv) > 1 else 512), 32
rng = np.random.RandomState(42)
X = np.tanh(rng.randn(n, D)); y = np.sin(X.sum(1) / 4) + 0.1 * rng.randn(n)
gp = B200GP()
gp.configure(D, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, {}, 'infExact', 'meanZero', 'covSEiso', 'likGauss', 'dlikExact', verbose=0)
gp.update_data('tr', X, y[:, None])
for _ in range(3):
    nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
print(nlZ, gp.stats())
