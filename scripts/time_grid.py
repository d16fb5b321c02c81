"""Timing of the KISS-GP path at large n (config C5: covSM Q=6 x2, meanSum, k=70)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
def synth_c5(n, seed=46, Q=6):
    rng = np.random.RandomState(seed)
    H = rng.rand(n, 2); y = np.sin(2 * np.pi * 3 * H).sum(1) + 0.5 * H[:, 0] + 0.3 * rng.randn(n)
    hyp_cov = [[np.log(np.ones((Q, 1)))], [np.log(np.maximum(0.05 * rng.rand(Q, 1), 1e-8))], [np.log(np.abs(rng.randn(Q, 1) + 1))]]
    return H, y, hyp_cov
for n in ([int(a) for a in sys.argv[1:]] if __name__ == "__main__" else []):
    H, y, hc = synth_c5(n)
    gp = B200GP()
    gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc}, {'cg_maxit': int(os.environ.get('MAXIT', 2000)), 'cg_tol': 1e-6}, 'infGrid', 'meanSum', 'covSM',
                 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6], mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    for rep in range(3):
        t0 = time.time(); nlZ, g = gp.eng.eval(0, 2, gp._n_hyp); t1 = time.time(); st = gp.stats()
        it = st['cg_iters'] - 1
        bytes_iter = n * (16 * 2 + 88) + 32 * 4900
        print('n=%d rep %d nlZ %.8g wall %.4fs dev %.2f ms cg %.2f ms iters %d -> %.1f us/iter, %.0f GB/s algorithmic (%.1f MB/iter)' % (
            n, rep, nlZ, t1 - t0, st['ms_total'], st['ms_solve'], st['cg_iters'], 1e3 * st['ms_solve'] / max(it, 1),
            bytes_iter * it / (st['ms_solve'] * 1e-3) / 1e9, bytes_iter / 1e6), flush=True)
