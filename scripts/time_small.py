"""Latency of the reference-sized configurations (C1 exact n=512 d=32; n=2048): device ms per evaluation and phases."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
for n, D in ((512, 32), (1024, 8), (2048, 8), (4096, 8)):
    rng = np.random.RandomState(42)
    X = np.tanh(rng.randn(n, D)); y = np.sin(X.sum(1) / 4) + 0.1 * rng.randn(n)
    gp = B200GP()
    gp.configure(D, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, {}, 'infExact', 'meanZero', 'covSEiso', 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    for _ in range(5):
        nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
    t = []; t0 = time.perf_counter()
    for _ in range(30):
        gp.eng.eval(0, 2, gp._n_hyp); t.append(gp.stats())
    wall = (time.perf_counter() - t0) / 30
    med = lambda k: float(np.median([s[k] for s in t]))
    print('n=%d D=%d nlZ %.10g: dev %.3f ms (kb %.3f potrf %.3f inv %.3f solve %.3f red %.3f) wall %.3f ms, %d launches' % (
        n, D, nlZ, med('ms_total'), med('ms_kbuild'), med('ms_potrf'), med('ms_inverse'), med('ms_solve'), med('ms_reduce'), 1e3 * wall, t[-1]['kernel_launches']), flush=True)
