"""KISS-GP at large n with the Woodbury-preconditioned CG (opt.cg_precond) next to the reference CG."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from kgp_b200 import B200GP
sys.path.insert(0, os.path.join(ROOT, 'scripts'))
from time_grid import synth_c5
for n in [int(a) for a in sys.argv[1:]]:
    H, y, hc = synth_c5(n)
    res = {}
    for pre in (1, 0):
        gp = B200GP()
        gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': [[a.copy() for a in b] for b in hc]}, {'cg_maxit': 30000, 'cg_tol': 1e-6, 'cg_precond': pre}, 'infGrid', 'meanSum', 'covSM',
                     'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6], mean_args=['{@meanLinear, @meanConst}'], verbose=0)
        t0 = time.time(); gp.update_data('tr', H, y[:, None]); gp.update_grid('tr'); t_up = time.time() - t0
        for rep in range(3):
            t0 = time.time(); nlZ, g = gp.eng.eval(0, 2, gp._n_hyp); t1 = time.time(); st = gp.stats()
            print('n=%d precond=%d rep %d nlZ %.10g wall %.4fs dev %.2f ms (setup %.2f cg %.2f rest %.2f) iters %d relres %.2e' % (
                n, pre, rep, nlZ, t1 - t0, st['ms_total'], st['ms_kbuild'], st['ms_solve'], st['ms_reduce'], st['cg_iters'], st['cg_relres']), flush=True)
        res[pre] = (nlZ, g, gp.alpha())
        del gp
    rel = lambda a, b: float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(b)))
    print('   precond vs plain: nlZ %.2e dnlZ %.2e alpha %.2e' % (rel(res[1][0], res[0][0]), rel(res[1][1], res[0][1]), rel(res[1][2], res[0][2])), flush=True)
