"""CPU: pins the oracle on the only numeric anchors the reference tree holds
(GPML/doc/index.html:971,1019-1021; data from GPML/doc/demoRegression.m:10-26,67-76)
and mirrors the reference's MATLAB gradient checks (tests/backend/matlab/grad_check_*.m)."""
import numpy as np
import scipy.linalg as sla

import gpml_exact as orc
from kgp_b200.minimize import minimize


def demo_data():
    n = 20
    x = orc.gpml_randn(0.3, n)[:, None]
    hc = np.log([0.25, 1.0])
    K = orc.cov_matern_iso(3, hc, x)
    mu = 0.5 * x[:, 0] + 1.0
    y = sla.cholesky(K).T @ orc.gpml_randn(0.15, n) + mu + 0.1 * orc.gpml_randn(0.2, n)
    return x, y, K


def test_gpml_randn_is_deterministic():
    a = orc.gpml_randn(0.3, 20)
    assert a.shape == (20,) and np.all(np.isfinite(a))
    assert np.array_equal(a, orc.gpml_randn(0.3, 20))


def test_kat1_nlml_11_97():
    x, y, K = demo_data()
    hyp = {'cov': np.log([0.25, 1.0]), 'lik': np.log(0.1), 'mean': np.array([0.5, 1.0])}
    _, nlZ, _ = orc.inf_exact(hyp, ['meanLinear', 'meanConst'], None, x, y, want_grad=False, K=K.copy())
    assert '%.2f' % nlZ == '11.97'                      # GPML/doc/index.html:971
    assert abs(nlZ - 11.9706317) < 1e-6


def test_kat2_minimize_14_13():
    x, y, _ = demo_data()

    def f(v):
        hyp = {'cov': v[:2], 'lik': v[2], 'mean': []}
        _, nl, d = orc.inf_exact(hyp, ['meanZero'], 'iso', x, y)
        return nl, np.concatenate([d['cov'], [d['lik']]])

    X, fX, nev = minimize(np.array([0.0, 0.0, np.log(0.1)]), f, -100)
    nl = f(X)[0]
    assert '%.2f' % nl == '14.13' and '%.2f' % np.exp(X[2]) == '0.15'      # index.html:1019-1021
    assert nev <= 101 and all(b <= a + 1e-12 for a, b in zip(fX, fX[1:]))


def _fd(fun, v, h=1e-5):
    g = np.zeros_like(v)
    for i in range(v.size):
        e = np.zeros_like(v); e[i] = h
        g[i] = (fun(v + e) - fun(v - e)) / (2 * h)
    return g


def test_hyper_gradients_match_finite_differences():
    rng = np.random.RandomState(0)
    X = rng.randn(60, 4); y = np.sin(X[:, 0]) + 0.1 * rng.randn(60)
    for mode, nl in (('iso', 1), ('ard', 4)):
        v0 = np.concatenate([0.3 * rng.randn(nl), [0.2, -1.0, 0.3, -0.2, 0.1, 0.05, 0.7]])

        def split(v):
            return {'cov': v[:nl + 1], 'lik': v[nl + 1], 'mean': v[nl + 2:]}
        mean = ['meanLinear', 'meanConst']
        _, _, d = orc.inf_exact(split(v0), mean, mode, X, y)
        g = np.concatenate([d['cov'], [d['lik']], d['mean']])
        fd = _fd(lambda v: orc.inf_exact(split(v), mean, mode, X, y, want_grad=False)[1], v0)
        assert np.max(np.abs(g - fd)) / np.max(np.abs(fd)) < 1e-6


def test_dlik_exact_gradcheck_mirror():
    """tests/backend/matlab/grad_check_likExact.m:13-33: N=100, D=10, hyp log[0.5;1], lik log 0.1, < 1e-6."""
    rng = np.random.RandomState(42)
    X = rng.rand(100, 10); y = rng.rand(100)
    hyp = {'cov': np.log([0.5, 1.0]), 'lik': np.log(0.1), 'mean': []}
    an = orc.dlik_exact(hyp, ['meanZero'], 'iso', X, y)
    f = lambda xv: orc.inf_exact(hyp, ['meanZero'], 'iso', xv.reshape(100, 10), y, want_grad=False)[1]
    idx = rng.choice(1000, 40, replace=False)
    fd = np.zeros(40)
    for k, q in enumerate(idx):
        e = np.zeros(1000); e[q] = 1e-5
        fd[k] = (f(X.ravel() + e) - f(X.ravel() - e)) / 2e-5
    assert np.linalg.norm(an.ravel()[idx] - fd) / np.linalg.norm(an.ravel()[idx]) < 1e-6


def test_ard_quirk_forms_agree_for_equal_lengthscales():
    """dcovSEard.m:7,16 divides by ell_1^2 for every dimension."""
    rng = np.random.RandomState(1)
    X = rng.randn(50, 3); y = rng.randn(50)
    hyp = {'cov': np.array([0.3, 0.3, 0.3, 0.1]), 'lik': -1.0, 'mean': []}
    a = orc.dlik_exact(hyp, ['meanZero'], 'ard', X, y, ard_reference_form=True)
    b = orc.dlik_exact(hyp, ['meanZero'], 'ard', X, y, ard_reference_form=False)
    assert np.allclose(a, b, rtol=1e-12)
    hyp['cov'] = np.array([0.3, -0.2, 0.5, 0.1])
    a = orc.dlik_exact(hyp, ['meanZero'], 'ard', X, y, ard_reference_form=True)
    b = orc.dlik_exact(hyp, ['meanZero'], 'ard', X, y, ard_reference_form=False)
    assert not np.allclose(a, b, rtol=1e-3)


def test_golden_fixture_matches_oracle():
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), 'golden', 'exact_mean_777.npz'))
    rng = np.random.RandomState(7)
    X = rng.randn(777, 3); y = 0.5 * X[:, 0] - 0.2 * X[:, 2] + 1.0 + np.sin(2 * X[:, 1]) + 0.2 * rng.randn(777)
    hyp = {'cov': g['hyp_cov'], 'lik': float(g['hyp_lik']), 'mean': g['hyp_mean']}
    _, nlZ, d = orc.inf_exact(hyp, ['meanLinear', 'meanConst'], 'ard', X, y)
    assert abs(nlZ - float(g['nlZ'])) < 1e-9 * max(1, abs(nlZ))
    assert np.allclose(np.concatenate([d['cov'], [d['lik']], d['mean']]), g['dnlZ'], rtol=1e-9)
