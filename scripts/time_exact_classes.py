"""Per-call-site GEMM throughput of one exact evaluation (profiling events around every dgemm_nt launch)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as g
g.build()
from kgp_b200 import B200GP
NAMES = ['potrf panel update', 'potrf trailing SYRK', 'Z-sweep GEMM', 'trsm block update', 'Z Z^T', 'predict']
for n in [int(a) for a in sys.argv[1:]]:
    D = 8
    rng = np.random.RandomState(43)
    X = rng.randn(n, D); y = np.sin(X[:, 0] * X[:, 1]) + X[:, 2] ** 2 - X[:, 3] + 0.5 * np.cos(X.sum(1))
    y = (y - y.mean()) / y.std() + 0.1 * rng.randn(n)
    hc = np.concatenate([np.log(2.0) * np.ones(D), [0.0]])
    gp = B200GP()
    gp.configure(D, {'lik': np.log(0.1), 'mean': [], 'cov': hc[:, None].copy()}, {}, 'infExact', 'meanZero', 'covSEard', 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    gp.eng.eval(0, 2, hc.size + 1)
    gp.eng.set_profiling(True)
    gp.eng.eval(0, 2, hc.size + 1)
    st = gp.stats()
    print('n=%d dev %.1f ms (kb %.1f potrf %.1f inv %.1f solve %.1f red %.1f); gemm %.1f ms' % (
        n, st['ms_total'], st['ms_kbuild'], st['ms_potrf'], st['ms_inverse'], st['ms_solve'], st['ms_reduce'], st['ms_gemm']))
    for i, nm in enumerate(NAMES):
        ms, fl = st['gemm_class_ms'][i], st['gemm_class_flops'][i]
        if ms > 0:
            print('   %-22s %9.2f ms  %8.3e flop  %6.2f TF/s' % (nm, ms, fl, fl / ms / 1e9))
