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


def demo_grid_data(n=30):
    """The set-up of GPML/doc/demoGrid1d.m:6-12 (f = a x + b + sin x, x = 1 + 4u + sign(u), covSEiso ell 1 sf 2, sn 0.5,
    meanSum) with the draws taken from GPML's portable generator gpml_randn (doc/gpml_randn.m) instead of Octave's
    rand/randn('seed',3), which cannot be reproduced outside Octave."""
    import gpml_exact as orc
    from scipy.stats import norm
    a, b, sn = 0.3, 1.2, 0.5
    u = 2 * norm.cdf(orc.gpml_randn(0.3, n)) - 1                         # uniform(-1,1) from the portable normal stream
    x = 1 + 4 * u + np.sign(u)
    y = a * x + b + np.sin(x) + sn * orc.gpml_randn(0.15, n)
    hyp = {'cov': np.log([1.0, 2.0]), 'lik': np.log(sn), 'mean': np.array([a, b])}
    return x[:, None], y, hyp


def test_grid_oracle_agrees_with_kat_pinned_exact_oracle():
    """KISS-GP parity is unpinned by the reference (no infGrid number is recorded anywhere), so the grid oracle is tied
    to the exact oracle -- itself pinned on GPML's manual values -- the way demoGrid1d.m:22-33 compares the two paths:
    on a fine grid the SKI covariance approaches the exact one, hence alpha, the data-fit term and the predictive mean
    must agree to the interpolation error; CG alpha must equal a dense solve with the SKI matrix; the Fiedler bound
    must lie above the dense 1/2 log det of the SKI matrix (it is an upper bound, apx.m:312-318)."""
    import gpml_exact as orc
    X, y, hyp = demo_grid_data()
    mean = ['meanLinear', 'meanConst']
    xs = np.linspace(-5.5, 7.5, 300)[:, None]
    post_e, nlZ_e, _ = orc.inf_exact(hyp, mean, 'iso', X, y)
    ymu_e = orc.gp_predict(hyp, mean, 'iso', X, y, xs, post_e)[0]
    m, _ = orc.mean_eval(mean, hyp['mean'], X)
    fit_e = (y - m) @ post_e['alpha'] / 2
    errs = []
    for ng in (40, 160, 640):
        xg = [np.linspace(-6, 8, ng)]
        opt = {'cg_maxit': 5000, 'cg_tol': 1e-12}
        post, nlZ, _ = gg.inf_grid(hyp, mean, 'covSEiso', xg, X, y, opt)
        ymu = gg.predict_grid(hyp, mean, 'covSEiso', xg, post, xs)[2]
        fit = (y - m) @ post['alpha'] / 2
        errs.append((np.max(np.abs(post['alpha'] - post_e['alpha'])) / np.max(np.abs(post_e['alpha'])),
                     abs(fit - fit_e) / abs(fit_e), np.max(np.abs(ymu - ymu_e)) / np.max(np.abs(ymu_e))))
        # CG alpha = dense solve with the explicit SKI matrix; Fiedler bound >= dense 1/2 log det
        Mx = gg.interpgrid(xg, X).toarray(); K = gg.factor_cov('covSEiso', hyp['cov'], xg[0])[0]
        sn2 = np.exp(2 * hyp['lik']); Kski = Mx @ K @ Mx.T
        assert np.allclose(post['alpha'], np.linalg.solve(Kski + sn2 * np.eye(len(y)), y - m), rtol=1e-8, atol=1e-10)
        ld_dense = 0.5 * np.linalg.slogdet(np.eye(len(y)) + Kski / sn2)[1]
        ld_fiedler = nlZ - fit - len(y) * np.log(2 * np.pi * sn2) / 2
        assert ld_fiedler >= ld_dense - 1e-9
    # cubic interpolation: the error falls with the grid spacing (>= 30x from 40 to 640 nodes) and is small on the fine grid
    for k in range(3):
        assert errs[2][k] < 2e-5 and errs[2][k] < errs[0][k] / 30, errs


# ---- Toeplitz factors (nested-cell xg; apxGrid.m:258-281,335-346,571-588,617-627,642-650; apx.m:256-266) -------------
def _toep_case(kind, seed=3, n=400):
    rng = np.random.RandomState(seed)
    X = rng.rand(n, 2) * np.array([2.0, 1.0]); y = np.sin(3 * X[:, 0]) + np.cos(5 * X[:, 1]) + 0.2 * rng.randn(n)
    a1 = np.linspace(-0.3, 2.3, 64); a2 = np.linspace(-0.2, 1.2, 20)
    if kind == 'covSEiso':
        hc, Q = np.array([np.log(0.4), 0.1, np.log(0.3), -0.2]), 0
    else:
        Q = 2
        blk = lambda: np.concatenate([np.log(rng.rand(Q) + 0.5), np.log(0.5 * rng.rand(Q) + 0.1), np.log(0.4 * rng.rand(Q) + 0.2)])
        hc = np.concatenate([blk(), blk()])
    return X, y, a1, a2, hc, Q


def test_toeplitz_mvm_equals_dense_product():
    X, y, a1, a2, hc, Q = _toep_case('covSM')
    f = gg.ToepFactor('covSM', hc[:6], a1, Q)
    Kd = f.dense()
    assert np.allclose(Kd, gg.factor_cov('covSM', hc[:6], a1, Q)[0], atol=1e-13)      # same matrix as the dense factor
    A = np.random.RandomState(0).randn(64, 7)
    assert np.allclose(f.mvm(A), Kd @ A, atol=1e-12)
    # mixed Kronecker product: Toeplitz first factor, dense second factor
    K2 = gg.factor_cov('covSM', hc[6:], a2, Q)[0]
    v = np.random.RandomState(1).randn(64 * 20, 3)
    assert np.allclose(gg.kronmvm([f, K2], v), np.kron(K2, Kd) @ v, atol=1e-11)


def test_whittle_eigenvalues_follow_the_closed_form():
    X, y, a1, a2, hc, Q = _toep_case('covSEiso')
    f = gg.ToepFactor('covSEiso', hc[:2], a1)
    c, xw = gg.circ_whittle(f, 3)
    n, h = f.n, f.h
    ell, sf2 = np.exp(hc[0]), np.exp(2 * hc[1])
    k = lambda t: sf2 * np.exp(-0.5 * (t / ell) ** 2)
    c0 = np.array([sum(k((j + r * n) * h) for r in range(-3, 3)) for j in range(n)])
    assert np.allclose(c, c0, rtol=1e-13) and xw.size == 6 * n and xw[0] == 1 - 3 * n
    _, ee, _ = gg.eigkron([f])
    assert np.allclose(ee[0], np.maximum(np.real(np.fft.fft(c0)), 0), atol=1e-12)
    ed = np.linalg.eigvalsh(f.dense())
    assert abs(ee[0].sum() - ed.sum()) / ed.sum() < 0.2 and not np.allclose(np.sort(ee[0]), ed, rtol=1e-3)   # a different spectrum


import pytest


@pytest.mark.parametrize('kind', ['covSEiso', 'covSM'])
def test_toeplitz_mode_inference_and_derivatives(kind):
    """Toeplitz mode changes only the log-det term (Whittle eigenvalues); alpha, the data-fit term and dirder equal the
    dense-factor values; the analytic hyper-gradient matches central differences of the Toeplitz-mode nlZ."""
    X, y, a1, a2, hc, Q = _toep_case(kind)
    opt = {'cg_maxit': 4000, 'cg_tol': 1e-13}
    hyp = {'cov': hc, 'lik': np.log(0.2), 'mean': []}
    xg_t, xg_d = [[a1], a2], [a1, a2]
    pt, nt, dt = gg.inf_grid(hyp, ['meanZero'], kind, xg_t, X, y, opt, Q=Q)
    pd, nd, dd = gg.inf_grid(hyp, ['meanZero'], kind, xg_d, X, y, opt, Q=Q)
    assert np.allclose(pt['alpha'], pd['alpha'], rtol=1e-8, atol=1e-10)
    assert abs(pt['ldB2'] - pd['ldB2']) > 1e-6 * abs(pd['ldB2'])                     # a different numerical definition
    assert abs((nt - pt['ldB2']) - (nd - pd['ldB2'])) < 1e-8 * abs(nd)
    assert np.allclose(gg.predict_grid(hyp, ['meanZero'], kind, xg_t, pt, X[:50], Q=Q)[0],
                       gg.predict_grid(hyp, ['meanZero'], kind, xg_d, pd, X[:50], Q=Q)[0], rtol=1e-8, atol=1e-10)

    def f(v):
        return gg.inf_grid({'cov': v[:-1], 'lik': v[-1], 'mean': []}, ['meanZero'], kind, xg_t, X, y, opt, Q=Q, want_grad=False)[1]
    v0 = np.concatenate([hc, [hyp['lik']]]); g = np.concatenate([dt['cov'], [dt['lik']]])
    fd = _fd(f, v0, 1e-5)
    assert np.max(np.abs(fd - g)) < 2e-6 * np.max(np.abs(g)), (fd, g)
    # both Toeplitz: only ldB2 and its derivative change again
    pt2, nt2, dt2 = gg.inf_grid(hyp, ['meanZero'], kind, [[a1], [a2]], X, y, opt, Q=Q)
    assert np.allclose(pt2['alpha'], pd['alpha'], rtol=1e-8, atol=1e-10)
    assert np.allclose(gg.dlik_grid(hyp, ['meanZero'], kind, [[a1], [a2]], X, y, opt, Q=Q),
                       gg.dlik_grid(hyp, ['meanZero'], kind, xg_d, X, y, opt, Q=Q), rtol=1e-7, atol=1e-9)


def test_neqinterp_matches_definition_and_gradcheck():
    """Non-equispaced axes (apxGrid.m:453-465 dispatch on equi(), neqinterp :547-557): linear weights on the bracketing nodes, weight 1
    on the end node outside the axis, rows sum to 1; dlik_grid (deriv_x.m:11-15: w_i = 1) against central differences of nlZ."""
    ax = np.array([0., 0.3, 0.5, 1.1, 1.2, 2.0])
    J, C, dC = gg.neqinterp(ax, np.array([-0.2, 0.29, 0.8, 2.5]))
    assert J.tolist() == [[0, 1], [0, 1], [2, 3], [4, 5]]
    assert np.allclose(C, [[1, 0], [1 / 30, 29 / 30], [0.5, 0.5], [0, 1]]) and np.allclose(C.sum(1), 1)
    assert np.allclose(dC[[0, 3]], 0) and np.allclose(dC[2], [-1 / 0.6, 1 / 0.6])
    assert gg.equi(np.linspace(0, 1, 9)) and not gg.equi(ax)
    rng = np.random.RandomState(5)
    n = 120; X = rng.rand(n, 2); y = np.sin(4 * X[:, 0]) + X[:, 1] + 0.1 * rng.randn(n)
    a1 = np.sort(np.concatenate([[-0.3, 1.3], rng.rand(22) * 1.4 - 0.2])); a2 = np.linspace(-0.4, 1.4, 19)       # neq x eq
    hyp = {'cov': np.array([np.log(0.3), 0.0, np.log(0.4), 0.1]), 'lik': np.log(0.1), 'mean': []}
    opt = {'cg_maxit': 3000, 'cg_tol': 1e-13}
    Mx = gg.interpgrid([a1, a2], X)
    assert np.allclose(Mx.sum(1), 1.0) and (Mx.getnnz(axis=1) <= 8).all()
    dX = gg.dlik_grid(hyp, ['meanZero'], 'covSEiso', [a1, a2], X, y, opt)
    f = lambda xx: gg.inf_grid(hyp, ['meanZero'], 'covSEiso', [a1, a2], xx, y, opt, want_grad=False)[1]
    for (i, d) in ((3, 0), (40, 1), (77, 0)):
        e = np.zeros_like(X); e[i, d] = 1e-6
        # the Fiedler bound does not depend on x, the data-fit term does; linear interpolation is differentiable inside a cell
        assert abs((f(X + e) - f(X - e)) / 2e-6 - dX[i, d]) < 1e-5 * max(1.0, abs(dX[i, d]))


def proj_problem(seed=5, n=400, D=5, d=2):
    rng = np.random.RandomState(seed)
    X = rng.randn(n, D)
    P = 0.4 * rng.randn(d, D)
    y = np.sin((X @ P.T).sum(1)) + 0.1 * rng.randn(n)
    xg = [np.linspace(-6.0, 6.0, 40), np.linspace(-6.0, 6.0, 36)]
    return X, y, P, xg


@pytest.mark.parametrize('mode', ['none', 'orth', 'norm'])
def test_proj_gradient_matches_finite_differences(mode):
    """hyp.proj (infGrid.m:56-65,82-87,155-186): dnlZ.proj from deriv_proj + the orth / norm chain rule against central
    differences of nlZ in the entries of hyp.proj; dlik_grid_proj = dxP hP against differences in x."""
    X, y, P, xg = proj_problem()
    opt = {'cg_tol': 1e-13, 'cg_maxit': 5000}
    hyp = {'cov': np.array([0.1, 0.2, -0.1, 0.1]), 'lik': np.log(0.2), 'mean': np.array([0.3, -0.2, 0.1])}
    mean = ['meanLinear', 'meanConst']
    f = lambda pv: gg.inf_grid_proj(dict(hyp, proj=pv.reshape(P.shape)), mean, 'covSEiso', xg, X, y, opt, mode=mode)
    post, nlZ, d, hP = f(P.ravel())
    fd = _fd(lambda pv: f(pv)[1], P.ravel().copy()).reshape(P.shape)
    assert np.max(np.abs(d['proj'] - fd)) / np.max(np.abs(fd)) < 2e-5
    if mode == 'orth':
        assert np.allclose(hP @ hP.T, np.eye(2), atol=1e-12)
    if mode == 'norm':
        assert np.allclose(np.sum(hP * hP, 1), 1.0, atol=1e-12)
    dX = gg.dlik_grid_proj(dict(hyp, proj=P), mean, 'covSEiso', xg, X, y, opt, mode=mode)
    assert dX.shape == X.shape
    for (i, j) in ((0, 0), (7, 3), (123, 4)):
        def fx(t):
            Xt = X.copy(); Xt[i, j] += t
            return gg.inf_grid_proj(dict(hyp, proj=P), mean, 'covSEiso', xg, Xt, y, opt, mode=mode)[1]
        num = (fx(1e-5) - fx(-1e-5)) / 2e-5
        assert abs(num - dX[i, j]) < 2e-5 * max(1.0, np.max(np.abs(dX)))


def test_kmeans_grid_eq0_structure_and_host_routine():
    """covGrid('create', x, 0, k) (apxGrid.m:393-415): k-means centres of the unique values plus boundary nodes.  Octave's
    kmeans is randomly initialised (not reproducible), so the deterministic Lloyd stand-in is checked by what defines a k-means
    solution -- every value sits in the cluster of its nearest centre, every centre is its cluster's mean -- by the structure of
    the axis the reference builds around it, and bit for bit against the engine's host routine (kgp_kmeans1d_host, no GPU)."""
    import ctypes as C
    from kgp_b200 import _cabi
    rng = np.random.RandomState(11)
    x = np.concatenate([rng.randn(300) * 0.2 - 2.0, rng.randn(500) * 0.5 + 1.0, rng.rand(100) * 6 - 3])[:, None]
    x = np.round(x, 3)                                                  # duplicates: unique() matters
    u = np.unique(x[:, 0])
    for k in (12, 40, 0.05):
        xg = gg.creategrid(x, 0, k)[0]
        ngj = int(np.ceil(k * u.size)) if k <= 1 else int(k)
        assert xg.size == ngj and np.all(np.diff(xg) > 0)
        du = (u[-1] - u[0]) / ngj
        assert xg[0] == u[0] - 5 * du and xg[-1] == u[-1] + 5 * du       # bu (apxGrid.m:407)
        cent = xg[1:-1]                                                 # nb = 2: one boundary node per side
        lab = np.argmin(np.abs(u[:, None] - cent[None, :]), axis=1)
        assert np.all(np.diff(lab) >= 0)
        for c in range(cent.size):
            assert abs(u[lab == c].mean() - cent[c]) < 1e-12            # a fixed point of Lloyd's iteration
        assert not gg.equi(xg)
        out = np.zeros(cent.size)
        rc = _cabi.lib().kgp_kmeans1d_host(u.ctypes.data_as(C.c_void_p), u.size, cent.size, out.ctypes.data_as(C.c_void_p))
        assert rc == 0 and np.array_equal(out, cent)
    # more clusters requested than unique values: every value is its own centre, the rest are boundary nodes (apxGrid.m:412-415)
    xs = np.array([0.0, 1.0, 1.5, 4.0])[:, None]
    xg = gg.creategrid(xs, 0, 9)[0]
    assert xg.size == 9 and np.array_equal(xg[2:6], xs[:, 0]) and np.all(np.diff(xg) > 0)
    assert _cabi.lib().kgp_kmeans1d_host(u.ctypes.data_as(C.c_void_p), 3, 4, out.ctypes.data_as(C.c_void_p)) == -1
    # the oracle runs infGrid on such a grid (neqinterp path) and its gradient passes the finite-difference check
    X = rng.rand(150, 2); y = np.sin(5 * X[:, 0]) + X[:, 1] + 0.1 * rng.randn(150)
    xg2 = gg.creategrid(X, 0, 16)
    hyp = {'cov': np.array([np.log(0.3), 0.0, np.log(0.4), 0.1]), 'lik': np.log(0.1), 'mean': []}
    opt = {'cg_maxit': 3000, 'cg_tol': 1e-13}
    _, nl, d = gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg2, X, y, opt)
    for i in range(4):
        hp = dict(hyp); hm = dict(hyp); e = np.zeros(4); e[i] = 1e-6
        hp['cov'] = hyp['cov'] + e; hm['cov'] = hyp['cov'] - e
        fd = (gg.inf_grid(hp, ['meanZero'], 'covSEiso', xg2, X, y, opt, want_grad=False)[1] -
              gg.inf_grid(hm, ['meanZero'], 'covSEiso', xg2, X, y, opt, want_grad=False)[1]) / 2e-6
        assert abs(fd - d['cov'][i]) < 1e-5 * max(1.0, abs(fd))


@pytest.mark.parametrize('deg', [1, 0])
def test_eqinterp_low_degrees(deg):
    """opt.deg (apx.m:27-32,122; apxGrid.m:495-523): linear (1) and nearest-node (0) interpolation on equispaced axes -- two
    nodes per axis, rows sum to 1, boundary rule on the two end nodes; deg = 1 reproduces linear functions exactly inside the grid and
    its dlik_grid passes the central-difference check (deg = 0 has zero input gradient: dk = 0, apxGrid.m:498)."""
    s = np.linspace(-1.0, 2.0, 13); h = s[1] - s[0]
    t = np.array([-1.5, -1.0, -0.9, 0.123, 0.5, s[5], s[5] + 0.5 * h, 1.99, 2.0, 2.7])
    J, C, dC = gg.eqinterp(s, t, deg)
    assert J.shape == (t.size, 2) and np.allclose(C.sum(1), 1.0) and J.min() >= 0 and J.max() <= s.size - 1
    inside = (t >= s[0]) & (t < s[-1])
    if deg == 1:
        assert np.allclose((C * s[J]).sum(1)[inside], t[inside])                       # exact on linear functions
        off = inside & (np.abs((t - s[0]) / h - np.round((t - s[0]) / h)) > 1e-9)      # at a node dk(0) = -sign(0) = 0: one-sided weights
        assert np.allclose((dC * s[J]).sum(1)[off] / (s[-1] - s[0]), 1.0)              # d/dt of the same, deriv_x's w_i scaling
    else:
        assert set(np.unique(C)) <= {0.0, 1.0} and np.all(dC == 0)
        near = s[J[np.arange(t.size), np.argmax(C, axis=1)]]
        assert np.all(np.abs(near - t)[inside] <= 0.5 * h + 1e-12)
        assert near[6] == s[5]                                                       # a tie at w = 0.5 goes to the lower node
    assert np.array_equal(C[0], [1.0, 0.0]) and J[0, 0] == 0                         # below the axis: first node
    assert np.array_equal(C[-1], [0.0, 1.0]) and J[-1, 1] == s.size - 1              # above: last node
    rng = np.random.RandomState(9)
    n = 140; X = rng.rand(n, 2); y = np.sin(4 * X[:, 0]) + X[:, 1] + 0.1 * rng.randn(n)
    xg = [np.linspace(-0.2, 1.2, 17), np.linspace(-0.3, 1.3, 15)]
    hyp = {'cov': np.array([np.log(0.3), 0.0, np.log(0.4), 0.1]), 'lik': np.log(0.1), 'mean': []}
    opt = {'cg_maxit': 3000, 'cg_tol': 1e-13, 'deg': deg}
    Mx = gg.interpgrid(xg, X, deg=deg)
    assert np.allclose(Mx.sum(1), 1.0) and (Mx.getnnz(axis=1) <= 4).all()
    _, nl3, _ = gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, X, y, dict(opt, deg=3))
    _, nl, d = gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, X, y, opt)
    assert abs(nl - nl3) > 1e-6                                                      # the option is honoured
    for i in range(4):                                                               # hyper-gradient vs central differences
        e = np.zeros(4); e[i] = 1e-6
        hp = dict(hyp, cov=hyp['cov'] + e); hm = dict(hyp, cov=hyp['cov'] - e)
        fd = (gg.inf_grid(hp, ['meanZero'], 'covSEiso', xg, X, y, opt, want_grad=False)[1] -
              gg.inf_grid(hm, ['meanZero'], 'covSEiso', xg, X, y, opt, want_grad=False)[1]) / 2e-6
        assert abs(fd - d['cov'][i]) < 1e-5 * max(1.0, abs(fd))
    # one degree per grid dimension (apxGrid.m:448): cubic x this degree
    Mm = gg.interpgrid(xg, X, deg=[3, deg])
    assert np.allclose(Mm.sum(1), 1.0) and (Mm.getnnz(axis=1) <= 8).all() and abs(Mm - Mx).max() > 1e-3
    dX = gg.dlik_grid(hyp, ['meanZero'], 'covSEiso', xg, X, y, opt)
    if deg == 0:
        assert np.all(dX == 0)
    else:
        f = lambda xx: gg.inf_grid(hyp, ['meanZero'], 'covSEiso', xg, xx, y, opt, want_grad=False)[1]
        for (i, dd) in ((3, 0), (40, 1), (77, 0)):
            e = np.zeros_like(X); e[i, dd] = 1e-6
            assert abs((f(X + e) - f(X - e)) / 2e-6 - dX[i, dd]) < 1e-5 * max(1.0, abs(dX[i, dd]))
