"""SURVEY 8(d) C5 variant with the k = 1000 grid (N = 10^6 nodes, Toeplitz factors via nested xg): timing, no oracle."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
from bench import synth_c5
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
k = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
toep = (sys.argv[3] != 'dense') if len(sys.argv) > 3 else True
H, y, hc = synth_c5(n)
ax = [np.linspace(H[:, d].min() - 5.0 / k, H[:, d].max() + 5.0 / k, k) for d in range(2)]
xg = [[a] for a in ax] if toep else ax
gp = B200GP()
gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc}, {'cg_maxit': int(os.environ.get('MAXIT', 30000)), 'cg_tol': 1e-6}, 'infGrid', 'meanSum', 'covSM',
             'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': float(k), 'xg': xg}, cov_args=[6], mean_args=['{@meanLinear, @meanConst}'], verbose=0)
gp.update_data('tr', H, y[:, None])
for rep in range(2):
    t0 = time.time()
    try:
        nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
    except Exception as e:
        print('eval failed:', e); nlZ = float('nan')
    t1 = time.time(); st = gp.stats(); it = max(st['cg_iters'] - 1, 1)
    print('n=%d k=%d toep=%s rep %d nlZ %.8g wall %.3fs dev %.1f ms (setup %.1f cg %.1f rest %.1f) iters %d -> %.1f us/iter, %d launches' % (
        n, k, toep, rep, nlZ, t1 - t0, st['ms_total'], st['ms_kbuild'], st['ms_solve'], st['ms_reduce'], st['cg_iters'], 1e3 * st['ms_solve'] / it, st['kernel_launches']), flush=True)
