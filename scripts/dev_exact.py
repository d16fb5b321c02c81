"""Development check of the exact path on a GPU box: parity vs oracle at several sizes + peak probes."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import numpy as np
import __graft_entry__ as g
g.build()
from kgp_b200 import B200GP
import gpml_exact as orc

def rel(a, b):
    a = np.asarray(a, dtype=float); b = np.asarray(b, dtype=float)
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))

def case(n, D, cov, seed, ns=200, lik=np.log(0.1)):
    rng = np.random.RandomState(seed)
    X = rng.randn(n, D); y = np.sin(X[:, 0] * X[:, -1]) + 0.5 * np.cos(X.sum(1)) + 0.1 * rng.randn(n)
    y = (y - y.mean()) / y.std()
    Xs = rng.randn(ns, D)
    if cov == 'covSEard':
        hc = np.concatenate([np.log(2.0) * np.ones(D) + 0.1 * rng.randn(D), [0.0]])
        mode = 'ard'
    else:
        hc = np.array([np.log(1.5), 0.1]); mode = 'iso'
    gp = B200GP()
    gp.configure(D, {'lik': lik, 'mean': [], 'cov': hc[:, None].copy()}, {}, 'infExact', 'meanZero', cov, 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    t0 = time.time()
    nlZ, gr = gp.eng.eval(0, 2, hc.size + 1)
    t1 = time.time()
    st = gp.stats()
    dX = gp.get_dlik_dx('tr')
    ymu, ys2 = gp.predict(Xs, return_var=True)
    oh = {'cov': hc, 'lik': lik, 'mean': []}
    t2 = time.time()
    post, nlZ0, d0 = orc.inf_exact(oh, ['meanZero'], mode, X, y)
    t3 = time.time()
    dX0 = orc.dlik_exact(oh, ['meanZero'], mode, X, y)
    p0 = orc.gp_predict(oh, ['meanZero'], mode, X, y, Xs, post)
    g0 = np.concatenate([d0['cov'], [d0['lik']]])
    print('n=%d D=%d %s: nlZ %.12g vs %.12g | rel nlZ %.2e dnlZ %.2e dX %.2e ymu %.2e ys2 %.2e alpha %.2e | gpu %.3fs (dev %.1f ms: kb %.1f potrf %.1f inv %.1f solve %.1f red %.1f; gemm %.3g flop, %d launches) cpu %.2fs' % (
        n, D, cov, nlZ, nlZ0, rel(nlZ, nlZ0), rel(gr, g0), rel(dX, dX0), rel(ymu[:, 0], p0[0]), rel(ys2[:, 0], p0[1]),
        rel(gp.alpha(), post['alpha']), t1 - t0, st['ms_total'], st['ms_kbuild'], st['ms_potrf'], st['ms_inverse'], st['ms_solve'], st['ms_reduce'],
        st['gemm_flops'], st['kernel_launches'], t3 - t2), flush=True)
    print('   dnlZ gpu', gr, 'oracle', g0, flush=True)
    return gp

if __name__ == '__main__':
    sizes = [int(a) for a in sys.argv[1:]] or [100, 512, 1000, 2048, 4000]
    g.smoke()
    gp = case(100, 1, 'covSEiso', 42, lik=-2.0)
    print('DMMA peak TF/s:', gp.eng.bench_dmma_peak(20000), flush=True)
    for nn, kk in [(4096, 4096), (8192, 8192), (16384, 512)]:
        print('dgemm_nt n=%d k=%d: TF/s, ms =' % (nn, kk), gp.eng.bench_dgemm(nn, kk, 3), flush=True)
    for n in sizes:
        case(n, 8, 'covSEard', 43)
    case(512, 32, 'covSEiso', 42, lik=-2.0)
