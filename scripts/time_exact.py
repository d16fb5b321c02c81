"""Timing of the exact path at large n (no oracle)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as g
g.build()
from kgp_b200 import B200GP
for n in [int(a) for a in sys.argv[1:]]:
    D = 8
    rng = np.random.RandomState(43)
    X = rng.randn(n, D); y = np.sin(X[:, 0] * X[:, 1]) + X[:, 2] ** 2 - X[:, 3] + 0.5 * np.cos(X.sum(1))
    y = (y - y.mean()) / y.std() + 0.1 * rng.randn(n)
    hc = np.concatenate([np.log(2.0) * np.ones(D), [0.0]])
    gp = B200GP()
    gp.configure(D, {'lik': np.log(0.1), 'mean': [], 'cov': hc[:, None].copy()}, {}, 'infExact', 'meanZero', 'covSEard', 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    for rep in range(2):
        t0 = time.time()
        nlZ, gr = gp.eng.eval(0, 2, hc.size + 1)
        t1 = time.time()
        st = gp.stats()
        print('n=%d rep %d: nlZ %.10g wall %.3fs dev %.1f ms (kb %.1f potrf %.1f inv %.1f solve %.1f red %.1f) gemm %.4g flop -> %.2f TF/s overall, %d launches' % (
            n, rep, nlZ, t1 - t0, st['ms_total'], st['ms_kbuild'], st['ms_potrf'], st['ms_inverse'], st['ms_solve'], st['ms_reduce'],
            st['gemm_flops'], st['gemm_flops'] / st['ms_total'] / 1e9, st['kernel_launches']), flush=True)
    print('   dnlZ', gr, flush=True)
    t0 = time.time(); ymu, ys2 = gp.predict(X[:1000] + 0.01, return_var=True); print('   predict 1000: %.3fs' % (time.time() - t0), ymu[:2, 0], ys2[:2, 0], flush=True)
    del gp
