"""CPU: self-consistency of the KISS-GP oracle (the reference pins no infGrid numbers):
interpolation invariants, Kronecker MVM vs the explicit product, CG vs a dense solve, analytic
hyper- and input-gradients vs central differences (mirrors tests/backend/matlab/grad_check_likGrid.m)."""
import numpy as np

import gpml_grid as gg


def c3_like(n=150, k=12., seed=0):
    rng = np.random.RandomState(seed)
    X = np.maximum(0, rng.normal(0.5, 1, (n, 2)))            # ReLU-like features with exact zeros
    y = np.sin(3 * X[:, 0]) * np.cos(2 * X[:, 1]) + 0.3 * rng.randn(n)
    return X, y, gg.creategrid(X, 1, k)


def test_creategrid_matches_definition():
    X, _, xg = c3_like()
    for j, ax in enumerate(xg):
        u = np.unique(X[:, j]); du = (u[-1] - u[0]) / 12
        assert ax.size == 12 and np.isclose(ax[0], u[0] - 5 * du) and np.isclose(ax[-1], u[-1] + 5 * du)
        assert np.max(np.abs(np.diff(ax) - (ax[1] - ax[0]))) < 1e-12
    assert gg.creategrid(np.arange(10.)[:, None], 1, 0.5)[0].size == 5      # 0<=k<=1: fraction of unique points
    assert gg.creategrid(np.arange(10.)[:, None], 1, 3.)[0].size == 5       # at least 5 grid points


def test_interp_rows_sum_to_one_and_snap_at_boundary():
    X, _, xg = c3_like()
    Mx, dM = gg.interpgrid(xg, X, want_deriv=True)
    assert Mx.shape == (150, 144) and np.allclose(Mx.sum(1), 1.0, atol=1e-14)
    assert (Mx.getnnz(axis=1) <= 16).all()
    for d in dM:
        assert np.allclose(d.sum(1), 0.0, atol=1e-9)
    ax = np.linspace(0., 1., 11)
    J, C, dC = gg.eqinterp(ax, np.array([-0.3, 0.04, 0.5, 0.97, 1.4]))
    assert (C[[0, 1, 3, 4]].max(1) == 1.0).all() and (dC[[0, 1, 3, 4]] == 0).all()   # snapped to a single node
    assert J[0, 0] == 0 and J[-1, 0] == 7 and np.argmax(C[0]) == 0 and np.argmax(C[-1]) == 3
    assert np.isclose(C[2] @ ax[J[2]], 0.5)                                       # cubic reproduces linear functions


def test_kronmvm_equals_explicit_kronecker():
    X, _, xg = c3_like()
    hs = gg.split_hyp('covSEiso', [-0.3, 0.2, 0.1, -0.1], 2)
    Ks = [gg.factor_cov('covSEiso', hs[i], xg[i])[0] for i in range(2)]
    v = np.random.RandomState(1).randn(144, 3)
    assert np.allclose(np.kron(Ks[1], Ks[0]) @ v, gg.kronmvm(Ks, v), atol=1e-12)


def test_cg_matches_dense_solve_and_reports_iterations():
    rng = np.random.RandomState(2)
    A = rng.randn(40, 40); A = A @ A.T + 40 * np.eye(40); b = rng.randn(40, 2)
    x, flag, relres, it = gg.conjgrad(lambda v: A @ v, b, 1e-12, 200)
    assert flag == 1 and np.allclose(x, np.linalg.solve(A, b), rtol=1e-9) and 2 <= it <= 41
    x, flag, relres, it = gg.conjgrad(lambda v: A @ v, b, 1e-12, 3)
    assert flag == 0 and it == 3                                   # maxit allows maxit-1 MVMs (apx.m:289)


def _fd(f, v, h=1e-5):
    return np.array([(f(v + h * e) - f(v - h * e)) / (2 * h) for e in np.eye(v.size)])


def test_hyper_gradients_seiso_and_sm():
    X, y, xg = c3_like()
    opt = {'cg_tol': 1e-12, 'cg_maxit': 2000}
    v0 = np.array([-0.3, 0.2, 0.1, -0.1, np.log(0.3)])
    f = lambda v: gg.inf_grid({'cov': v[:4], 'lik': v[4], 'mean': []}, ['meanZero'], 'covSEiso', xg, X, y, opt)
    _, _, d = f(v0)
    g = np.concatenate([d['cov'], [d['lik']]])
    assert np.max(np.abs(g - _fd(lambda v: f(v)[1], v0))) / np.max(np.abs(g)) < 1e-6
    Q = 2; rng = np.random.RandomState(3)
    hsm = np.concatenate([np.zeros(Q), np.log(0.05 * rng.rand(Q) + 0.01), np.log(np.abs(rng.randn(Q) + 1))])
    v0 = np.concatenate([hsm, hsm + 0.1, [np.log(0.3)], [0.1, -0.2, 0.3]])
    f = lambda v: gg.inf_grid({'cov': v[:6 * Q], 'lik': v[6 * Q], 'mean': v[6 * Q + 1:]}, ['meanLinear', 'meanConst'],
                              'covSM', xg, X, y, opt, Q=Q)
    _, _, d = f(v0)
    g = np.concatenate([d['cov'], [d['lik']], d['mean']])
    assert np.max(np.abs(g - _fd(lambda v: f(v)[1], v0))) / np.max(np.abs(g)) < 1e-6


def test_dlik_grid_gradcheck_mirror():
    """grad_check_likGrid.m:13-34: N=100, D=1, covGrid('create',X,1,128), default CG -> ratio < 1e-2."""
    rng = np.random.RandomState(42)
    X = rng.rand(100, 1); y = rng.rand(100)
    xg = gg.creategrid(X, 1, 128.)
    hyp = {'cov': np.log([0.5, 1.0]), 'lik': np.log(0.1), 'mean': []}
    opt = {'cg_tol': 1e-6, 'cg_maxit': 2000}
    an = gg.dlik_grid(hyp, ['meanZero'], 'covSEiso', xg, X, y, opt)
    f = lambda xv: gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, xv[:, None], y, opt, want_grad=False)[1]
    fd = _fd(f, X[:, 0].copy())
    assert np.linalg.norm(an[:, 0] - fd) / np.linalg.norm(an[:, 0]) < 1e-2


def test_cg_non_convergence_raises():
    X, y, xg = c3_like()
    import pytest
    with pytest.raises(gg.CGNotConverged):
        gg.inf_grid({'cov': [-0.3, 0.2, 0.1, -0.1], 'lik': np.log(0.01), 'mean': []}, ['meanZero'], 'covSEiso', xg, X, y,
                    {'cg_tol': 1e-12, 'cg_maxit': 3})


def test_chebyshev_estimator_with_injected_probes():
    """logdet_sample (apx.m:336-391) with explicit probes / emax: derivative consistent with differences,
    and the value tracks the dense log-det when B is well conditioned."""
    X, y, xg = c3_like(n=120, k=8.)
    rng = np.random.RandomState(5); V = np.sign(rng.randn(120, 30))
    opt = {'cg_tol': 1e-12, 'cg_maxit': 2000}
    v0 = np.array([-0.3, -1.0, 0.1, -1.0, np.log(0.8)])
    hs = gg.split_hyp('covSEiso', v0[:4], 2); Ks = [gg.factor_cov('covSEiso', hs[i], xg[i])[0] for i in range(2)]
    Mx = gg.interpgrid(xg, X).toarray(); Kd = Mx @ np.kron(Ks[1], Ks[0]) @ Mx.T
    B = np.eye(120) + Kd / np.exp(2 * v0[4]); emax = np.linalg.eigvalsh(B).max() * 1.01
    f = lambda v: gg.inf_grid({'cov': v[:4], 'lik': v[4], 'mean': []}, ['meanZero'], 'covSEiso', xg, X, y, opt,
                              cheby={'V': V, 'd': 15, 'emax': emax})
    post, nlZ, d = f(v0)
    assert abs(post['ldB2'] - np.linalg.slogdet(B)[1] / 2) < 0.15 * abs(np.linalg.slogdet(B)[1] / 2) + 0.5
    fd = _fd(lambda v: f(np.concatenate([v, v0[4:]]))[1], v0[:4])           # cov derivatives (emax fixed)
    assert np.max(np.abs(d['cov'] - fd)) / np.max(np.abs(fd)) < 1e-5


def test_pred_var_matches_closed_form_statistically():
    """pred_var_grid (infGrid.m:91-101,147-152): E[vg] = diag(Ku M'(M Ku M' + sn2 I)^-1 M Ku); the reference states a
    relative accuracy of sqrt(2/ns) (infGrid.m:99).  Pins the restatement to the identity it estimates."""
    rng = np.random.RandomState(3)
    n = 300; x = rng.rand(n, 2); y = np.sin(4 * x[:, 0]) + np.cos(3 * x[:, 1]) + 0.1 * rng.randn(n)
    xg = gg.creategrid(x, 1, 12)
    hyp = {'cov': np.array([np.log(0.3), 0.0, np.log(0.3), 0.0]), 'lik': np.log(0.1), 'mean': []}
    post, _, _ = gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, x, y, {'cg_maxit': 2000, 'cg_tol': 1e-10}, want_grad=False)
    N = int(np.prod([len(a) for a in xg])); ns = 600
    vg = gg.pred_var_grid(post, rng.randn(N, ns), rng.randn(n, ns))
    Ks = post['_Ks']; Ku = np.kron(Ks[1], Ks[0]); M = post['_Mx'].toarray()
    ex = np.diag(Ku @ M.T @ np.linalg.solve(M @ Ku @ M.T + 0.01 * np.eye(n), M @ Ku))
    big = ex > 0.2 * ex.max()
    err = np.abs(vg[big] / ex[big] - 1.0)
    assert err.mean() < 2.0 * np.sqrt(2.0 / ns) and err.max() < 6.0 * np.sqrt(2.0 / ns)
    xs = rng.rand(40, 2)
    f0 = gg.predict_grid(hyp, ['meanZero'], 'covSEiso', xg, post, xs)
    f1 = gg.predict_grid(hyp, ['meanZero'], 'covSEiso', xg, post, xs, vg=vg)
    assert np.allclose(f0[0], f1[0]) and np.all(f1[1] <= f0[1] + 1e-12) and np.all(f1[1] >= 0) and f1[1].min() < 0.5 * f0[1].max()


def test_slq_matches_the_dense_quantities_it_estimates():
    """logdet_slq (extension): per probe, the Lanczos Gauss quadrature equals v' log(B) v of the dense
    matrix, and the Hutchinson derivative equals (1/2m) sum v' B^-1 dB v formed densely (dB by central differences)."""
    rng = np.random.RandomState(3)
    n = 200; x = rng.rand(n, 2); y = np.sin(4 * x[:, 0]) + np.cos(3 * x[:, 1]) + 0.1 * rng.randn(n)
    xg = gg.creategrid(x, 1, 10)
    hyp = {'cov': np.array([np.log(0.3), 0.0, np.log(0.3), 0.0]), 'lik': np.log(0.1), 'mean': []}
    opt = {'cg_maxit': 5000, 'cg_tol': 1e-12}
    m = 8; V = np.sign(rng.randn(n, m))
    post, nlZ, d = gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, x, y, opt, slq={'V': V, 'steps': 60})
    M = post['_Mx'].toarray()

    def Bof(hc):
        hs = gg.split_hyp('covSEiso', hc, 2, 0)
        Ks = [gg.factor_cov('covSEiso', hs[i], xg[i], 0)[0] for i in range(2)]
        return np.eye(n) + M @ np.kron(Ks[1], Ks[0]) @ M.T / 0.01
    B = Bof(hyp['cov']); th, S = np.linalg.eigh(B); logB = (S * np.log(th)) @ S.T
    ex = sum(V[:, i] @ logB @ V[:, i] for i in range(m)) / (2 * m)
    assert abs(post['ldB2'] - ex) < 1e-9 * abs(ex)
    Bi = np.linalg.inv(B); ld = []
    for i in range(4):
        e = np.zeros(4); e[i] = 1e-6
        dB = (Bof(hyp['cov'] + e) - Bof(hyp['cov'] - e)) / 2e-6
        ld.append(sum(V[:, k] @ Bi @ dB @ V[:, k] for k in range(m)) / (2 * m))
    hs = gg.split_hyp('covSEiso', hyp['cov'], 2, 0); fac = [gg.factor_cov('covSEiso', hs[i], xg[i], 0) for i in range(2)]
    fit = -gg.dirder([f[0] for f in fac], [f[1] for f in fac], post['_Mx'], post['alpha'], post['alpha']) / 2
    assert np.allclose(d['cov'] - fit, ld, rtol=1e-6)
    # lik derivative: B = I + K/sn2 -> d ldB2/d log sn = -(1/m) sum v' B^-1 (B - I) v / 1 ... checked by differences of the value
    f = lambda lik: gg.inf_grid({'cov': hyp['cov'], 'lik': lik, 'mean': []}, ['meanZero'], 'covSEiso', xg, x, y, opt,
                                slq={'V': V, 'steps': 60}, want_grad=False)[1]
    fd = (f(hyp['lik'] + 1e-5) - f(hyp['lik'] - 1e-5)) / 2e-5
    assert abs(d['lik'] - fd) < 1e-5 * abs(fd)               # dB commutes with B for the noise scale: exact derivative


def test_bench_kissgp_cpu_sample_runs():
    """bench.py's CPU arm of the KISS-GP leg (oracle pieces on a bounded sample) stays runnable."""
    import os, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    r = bench.kissgp_cpu_sample(20000, 100, sample_iters=3)
    assert r['kind'] == 'port' and r['unit'] == 'evals/s' and r['value'] > 0 and r['cores'] >= 1
