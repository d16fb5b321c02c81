"""Development check of the KISS-GP path on a GPU box: parity vs oracle."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import numpy as np
import __graft_entry__ as g
g.build()
from kgp_b200 import B200GP
import gpml_grid as gg

def rel(a, b):
    a = np.asarray(a, dtype=float).ravel(); b = np.asarray(b, dtype=float).ravel()
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))

def case(name, X, y, cov, hyp_cov, lik, k, opt, mean='meanZero', hyp_mean=(), mean_args=None, Q=0, Xs=None):
    n, D = X.shape
    gp = B200GP()
    hyp = {'lik': lik, 'mean': [[v] for v in hyp_mean], 'cov': hyp_cov}
    gp.configure(D, hyp, opt, 'infGrid', mean, cov, 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': k},
                 cov_args=[Q] if Q else None, mean_args=mean_args, verbose=0)
    gp.update_data('tr', X, y[:, None])
    gp.update_grid('tr')
    xg = gp.grid(); xg0 = gg.creategrid(X, 1, k)
    print(name, 'grid', [a.size for a in xg], 'axes err', max(np.max(np.abs(a - b)) for a, b in zip(xg, xg0)))
    t0 = time.time(); nlZ, gr = gp.eng.eval(0, 2, gp._n_hyp); t1 = time.time()
    st = gp.stats()
    dX = gp.get_dlik_dx('tr')
    ymu, ys2 = gp.predict(Xs, return_var=True)
    hv = gp._hyp_vec(); nc = gp.hyp['cov'].size
    oh = {'cov': hv[:nc], 'lik': lik, 'mean': hv[nc + 1:]}
    om = {'meanZero': ['meanZero'], 'meanSum': ['meanLinear', 'meanConst']}[mean]
    t2 = time.time(); post, nlZ0, d0 = gg.inf_grid(oh, om, cov, xg0, X, y, opt, Q=Q); t3 = time.time()
    ot = dict(opt); ot['cg_tol'] = 1e-13; ot['cg_maxit'] = 20000
    postT, nlZT, dT = gg.inf_grid(oh, om, cov, xg0, X, y, ot, Q=Q)
    dX0 = gg.dlik_grid(oh, om, cov, xg0, X, y, opt, Q=Q)
    p0 = gg.predict_grid(oh, om, cov, xg0, post, Xs, Q=Q)
    g0 = np.concatenate([d0['cov'], [d0['lik']], d0['mean']]); gT = np.concatenate([dT['cov'], [dT['lik']], dT['mean']])
    print('  nlZ %.12g vs %.12g | rel nlZ %.2e dnlZ %.2e dX %.2e alpha %.2e ymu %.2e ys2 %.2e | cg it %d vs %d relres %.2e | budget(nlZ %.1e dnlZ %.1e) | gpu %.4fs dev %.2f ms (setup %.2f cg %.2f rest %.2f) %d launches; cpu %.3fs' % (
        nlZ, nlZ0, rel(nlZ, nlZ0), rel(gr, g0), rel(dX, dX0), rel(gp.alpha(), post['alpha']), rel(ymu, p0[2]), rel(ys2, p0[3]),
        st['cg_iters'], post['cg_iter'], st['cg_relres'], rel(nlZ0, nlZT), rel(g0, gT), t1 - t0, st['ms_total'], st['ms_kbuild'], st['ms_solve'], st['ms_reduce'],
        st['kernel_launches'], t3 - t2), flush=True)
    print('  dnlZ gpu', gr[:8], '\n  oracle  ', g0[:8], flush=True)

rng = np.random.RandomState(45)
h = np.tanh(rng.randn(460, 1)); y = np.sin(4 * h[:, 0]) + np.exp(-2) * rng.randn(460)
case('C4', h, y, 'covSEiso', [[-0.7], [0.0]], -2.0, 100., {'cg_maxit': 500, 'cg_tol': 1e-4}, Xs=np.tanh(rng.randn(77, 1)) * 1.3)
rng = np.random.RandomState(42)
X = rng.rand(100, 1); y = rng.rand(100)
case('gradcheck(n<N)', X, y, 'covSEiso', [[np.log(0.5)], [0.0]], np.log(0.1), 128., {'cg_maxit': 500, 'cg_tol': 1e-8}, Xs=rng.rand(20, 1))
rng = np.random.RandomState(44)
for n, k in ((2000, 30.), (40000, 70.)):
    H = np.maximum(0, rng.normal(0.5, 1, (n, 2))); y = np.sin(3 * H[:, 0]) * np.cos(2 * H[:, 1]) + 0.3 * rng.randn(n)
    case('C3 n=%d' % n, H, y, 'covSEiso', [[0.5], [1.0]], np.log(0.3), k, {'cg_maxit': 2000, 'cg_tol': 1e-6}, Xs=np.maximum(0, rng.normal(0.5, 1.2, (500, 2))))
Q = 3
def initCovSM(Q, D, rng):
    w0 = np.log(np.ones((Q, 1))); mu = np.log(np.maximum(0.05 * rng.rand(Q * D, 1), 1e-8)); v = np.log(np.abs(rng.randn(Q * D, 1) + 1))
    return [[w0], [mu], [v]]
rng = np.random.RandomState(46)
n = 5000
H = rng.rand(n, 2); y = np.sin(2 * np.pi * 3 * H).sum(1) + 0.5 * H[:, 0] + 0.3 * rng.randn(n)
case('C5-like SM', H, y, 'covSM', initCovSM(Q, 1, rng), np.log(0.3), 40., {'cg_maxit': 2000, 'cg_tol': 1e-6}, mean='meanSum',
     hyp_mean=[0.1, -0.1, 0.2], mean_args=['{@meanLinear, @meanConst}'], Q=Q, Xs=rng.rand(300, 2))
rng = np.random.RandomState(3)
X = rng.randn(500, 1); X[::3] = X[1::3][:X[::3].shape[0]]
gp = B200GP(); gp.configure(1, {'lik': -1., 'mean': [], 'cov': [[0.], [0.]]}, {}, 'infGrid', 'meanZero', 'covSEiso', 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 0.5}, verbose=0)
gp.update_data('tr', X); gp.update_grid('tr')
print('fraction k=0.5: grid', gp.grid()[0].size, 'oracle', gg.creategrid(X, 1, 0.5)[0].size, 'err', np.max(np.abs(gp.grid()[0] - gg.creategrid(X, 1, 0.5)[0])))
