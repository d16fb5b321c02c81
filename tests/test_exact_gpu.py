"""GPU parity tests of the exact (infExact) path: CUDA engine through the C ABI vs the CPU oracle
and the committed golden vectors.  Tolerance: 1e-8 relative (north_star) on nlZ, dnlZ, dnlZ/dX,
ymu, ys2 -- vectors are compared in the max norm (|a-b|_inf / |b|_inf)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-8
GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def rel(a, b):
    a = np.asarray(a, dtype=float).ravel(); b = np.asarray(b, dtype=float).ravel()
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def make_gp(D, hyp_cov, lik, cov, mean='meanZero', hyp_mean=(), mean_args=None, opt=None):
    from kgp_b200 import B200GP
    gp = B200GP()
    hyp = {'lik': lik, 'mean': list(hyp_mean), 'cov': np.asarray(hyp_cov, dtype=float)[:, None].copy()}
    gp.configure(D, hyp, opt or {}, 'infExact', mean, cov, 'likGauss', 'dlikExact', mean_args=mean_args, verbose=0)
    return gp


def full_eval(gp, X, y):
    gp.update_data('tr', X, y[:, None])
    nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
    return nlZ, g, gp.get_dlik_dx('tr')


@pytest.mark.parametrize('n,D,cov,seed', [(100, 1, 'covSEiso', 42), (128, 2, 'covSEard', 1), (129, 5, 'covSEiso', 2),
                                           (640, 8, 'covSEard', 3), (1500, 16, 'covSEard', 4), (2000, 3, 'covSEiso', 5)])
def test_parity_vs_oracle(built_lib, n, D, cov, seed):
    import gpml_exact as orc
    rng = np.random.RandomState(seed)
    X = rng.randn(n, D); y = np.sin(X[:, 0]) + 0.3 * np.cos(X.sum(1)) + 0.1 * rng.randn(n)
    Xs = rng.randn(137, D)
    mode = 'ard' if cov == 'covSEard' else 'iso'
    nl = D if mode == 'ard' else 1
    hc = np.concatenate([np.log(1.5) + 0.2 * rng.randn(nl), [0.1]]); lik = np.log(0.1)
    gp = make_gp(D, hc, lik, cov)
    nlZ, g, dX = full_eval(gp, X, y)
    ymu, ys2 = gp.predict(Xs, return_var=True)
    oh = {'cov': hc, 'lik': lik, 'mean': []}
    post, nlZ0, d0 = orc.inf_exact(oh, ['meanZero'], mode, X, y)
    dX0 = orc.dlik_exact(oh, ['meanZero'], mode, X, y)
    p0 = orc.gp_predict(oh, ['meanZero'], mode, X, y, Xs, post)
    assert rel(nlZ, nlZ0) < TOL
    assert rel(g, np.concatenate([d0['cov'], [d0['lik']]])) < TOL
    assert rel(dX, dX0) < TOL
    assert rel(gp.alpha(), post['alpha']) < TOL
    assert rel(ymu, p0[0]) < TOL and rel(ys2, p0[1]) < TOL
    assert ymu.shape == (137, 1) and ys2.shape == (137, 1) and dX.shape == (n, D)


@pytest.mark.parametrize('name,cov,mean,margs', [('exact_c1', 'covSEiso', 'meanZero', None),
                                                 ('exact_c2_8192', 'covSEard', 'meanZero', None),
                                                 ('exact_mean_777', 'covSEard', 'meanSum', ['{@meanLinear, @meanConst}'])])
def test_golden_vectors(built_lib, name, cov, mean, margs):
    from bench import synth_c2
    g = np.load(os.path.join(GOLD, name + '.npz'))
    if name == 'exact_c1':
        rng = np.random.RandomState(42)
        X = np.tanh(rng.randn(512, 32)); y = np.sin(X.sum(1) / 4) + 0.1 * rng.randn(512)
    elif name == 'exact_c2_8192':
        X, y, _ = synth_c2(8192)
    else:
        rng = np.random.RandomState(7)
        X = rng.randn(777, 3); y = 0.5 * X[:, 0] - 0.2 * X[:, 2] + 1.0 + np.sin(2 * X[:, 1]) + 0.2 * rng.randn(777)
    gp = make_gp(X.shape[1], g['hyp_cov'], float(g['hyp_lik']), cov, mean, g['hyp_mean'], margs)
    nlZ, gr, dX = full_eval(gp, X, y)
    ymu, ys2 = gp.predict(g['Xs'], return_var=True)
    assert rel(nlZ, g['nlZ']) < TOL
    assert rel(gr, g['dnlZ']) < TOL
    assert rel(dX[g['dX_rows']], g['dX']) < TOL
    assert rel(gp.alpha()[g['alpha_rows']], g['alpha']) < TOL
    assert rel(ymu, g['ymu']) < TOL and rel(ys2, g['ys2']) < TOL


def test_ard_reference_form_quirk(built_lib):
    """opt.ard_reference_form reproduces dcovSEard.m:7,16 (every dim divided by ell_1^2)."""
    import gpml_exact as orc
    rng = np.random.RandomState(11)
    X = rng.randn(200, 3); y = rng.randn(200)
    hc = np.array([0.3, -0.2, 0.5, 0.1])
    gp = make_gp(3, hc, -1.0, 'covSEard', opt={'ard_reference_form': 1})
    _, _, dX = full_eval(gp, X, y)
    oh = {'cov': hc, 'lik': -1.0, 'mean': []}
    assert rel(dX, orc.dlik_exact(oh, ['meanZero'], 'ard', X, y, ard_reference_form=True)) < TOL


def test_backend_interface_mirror(built_lib):
    """tests/backend/test_gpml.py scenarios (N=100, D=1, seed 42): same calls, same types and shapes --
    plus the numbers, which the reference's tests never check."""
    import gpml_exact as orc
    from kgp_b200.minimize import minimize
    from kgp_b200 import GP_BACKEND
    rng = np.random.RandomState(42)
    N, D = 100, 1
    X_tr = rng.normal(size=(N, D)); y_tr = rng.normal(size=(N, 1))
    X_tst = rng.normal(size=(N, D)); y_tst = rng.normal(size=(N, 1))
    gp = GP_BACKEND()
    cfg = {'input_dim': 1, 'hyp': {'lik': -2.0, 'mean': [], 'cov': np.array([[-0.7], [0.0]])}, 'opt': {},
           'inf': 'infExact', 'mean': 'meanZero', 'cov': 'covSEiso', 'lik': 'likGauss', 'dlik': 'dlikExact', 'verbose': 0}
    gp.configure(**cfg)
    gp.update_data('tr', X_tr, y_tr)
    gp.update_data('tst', X_tst)
    nlZ = gp.evaluate('tr')
    assert isinstance(nlZ, float)
    nlZ2 = gp.evaluate('tmp', X_tst, y_tst)
    ymu, ys2 = gp.predict(X_tst, X_tr, y_tr, return_var=True)
    assert type(ymu) is np.ndarray and ymu.shape == (N, 1) and ys2.shape == (N, 1)
    dl = gp.get_dlik_dx('tr')
    assert type(dl) is np.ndarray and dl.shape == (N, D)
    nl3, ymu3 = gp.eval_predict(X_tst, X_tr, y_tr)
    assert ymu3.shape == (N, 1) and abs(nl3 - nlZ) < 1e-9 * abs(nlZ)
    oh = {'cov': np.array([-0.7, 0.0]), 'lik': -2.0, 'mean': []}
    assert rel(nlZ, orc.inf_exact(oh, ['meanZero'], 'iso', X_tr, y_tr, want_grad=False)[1]) < TOL
    assert rel(nlZ2, orc.inf_exact(oh, ['meanZero'], 'iso', X_tst, y_tst, want_grad=False)[1]) < TOL
    # train: same trajectory as the oracle's minimize on the oracle's gp
    hyp = gp.train(5, X_tr, y_tr)
    assert set(hyp.keys()) == {'cov', 'lik', 'mean'}

    def f(v):
        _, nl, d = orc.inf_exact({'cov': v[:2], 'lik': v[2], 'mean': []}, ['meanZero'], 'iso', X_tr, y_tr)
        return nl, np.concatenate([d['cov'], [d['lik']]])
    v, fX, nev = minimize(np.array([-0.7, 0.0, -2.0]), f, -5)
    assert nev == gp.last_train['n_evals']
    assert rel(np.concatenate([hyp['cov'].ravel(), [hyp['lik']]]), v) < 1e-7
    assert rel(gp.last_train['fX'], fX) < TOL


def test_not_positive_definite_is_reported(built_lib):
    """Octave's chol error -> KGP_NOT_PD; evaluate() turns it into NaN + warning like gp.m:117-123."""
    from kgp_b200 import _cabi
    rng = np.random.RandomState(3)
    X = rng.randn(300, 2); X[1] = X[0]; y = rng.randn(300)
    gp = make_gp(2, [0.0, 20.0], -40.0, 'covSEiso')       # sf^2/sn^2 = e^120: duplicates make B singular in FP64
    gp.update_data('tr', X, y[:, None])
    with pytest.raises(_cabi.KgpError) as ei:
        gp.eng.eval(0, 1, 3)
    assert ei.value.code == _cabi.NOT_PD
    with pytest.warns(UserWarning):
        assert np.isnan(gp.evaluate('tr'))
    with pytest.raises(_cabi.KgpError):
        gp.get_dlik_dx('tr')


def test_full_size_properties_n40000(built_lib):
    """BASELINE configs[1] at full size, through size-independent properties:
    (1) K alpha + sn2 alpha = y  (prediction at the training inputs closes the linear system),
    (2) dnlZ . dir matches a central difference of nlZ along a hyper-parameter direction,
    (3) translation invariance: sum_i dnlZ/dx_i = 0."""
    from bench import synth_c2
    n, D = 40000, 8
    X, y, hyp = synth_c2(n)
    gp = make_gp(D, hyp['cov'], hyp['lik'], 'covSEard')
    nlZ, g, dX = full_eval(gp, X, y)
    alpha = gp.alpha(); sn2 = np.exp(2 * hyp['lik'])
    sub = np.arange(0, n, 40)
    fmu, _ = gp.eng.predict(X[sub], want_var=False)
    assert rel(fmu + sn2 * alpha[sub], y[sub]) < 1e-8
    assert np.max(np.abs(dX.sum(0))) < 1e-8 * np.abs(dX).sum(0).max()
    rng = np.random.RandomState(0)
    dirv = rng.randn(D + 2); dirv /= np.linalg.norm(dirv); h = 1e-4
    vals = []
    for s in (+1, -1):
        v = np.concatenate([hyp['cov'], [hyp['lik']]]) + s * h * dirv
        gp.eng.set_model('infExact', 'covSEard', v[:D + 1], v[D + 1])
        vals.append(gp.eng.eval(0, 0, D + 2)[0])
    fd = (vals[0] - vals[1]) / (2 * h)
    assert abs(fd - g @ dirv) < 1e-5 * max(abs(fd), np.linalg.norm(g) * 1e-2)


def test_parity_vs_oracle_n16384_c2(built_lib):
    """BASELINE configs[1] shape (covSEard, d=8, the bench's synthetic inputs and hyper-parameters) at n = 16 384 -- the
    largest size the CPU oracle finishes in a couple of minutes on the box's 16 cores: every output against the oracle
    at the 1e-8 relative bar north_star states (cond(B) ~ 1e6 here, so cond * eps ~ 1e-10 leaves room)."""
    import gpml_exact as orc
    from bench import synth_c2
    n, D = 16384, 8
    X, y, hyp = synth_c2(n)
    Xs = np.random.RandomState(9).randn(500, D)
    gp = make_gp(D, hyp['cov'], hyp['lik'], 'covSEard')
    nlZ, g, dX = full_eval(gp, X, y)
    alpha = gp.alpha()
    ymu, ys2 = gp.predict(Xs, return_var=True)
    oh = {'cov': hyp['cov'], 'lik': hyp['lik'], 'mean': []}
    post, nlZ0, d0 = orc.inf_exact(oh, ['meanZero'], 'ard', X, y)
    g0 = np.concatenate([d0['cov'], [d0['lik']]])
    assert rel(nlZ, nlZ0) < TOL and rel(g, g0) < TOL and rel(alpha, post['alpha']) < TOL
    p0 = orc.gp_predict(oh, ['meanZero'], 'ard', X, y, Xs, post)
    assert rel(ymu[:, 0], p0[0]) < TOL and rel(ys2[:, 0], p0[1]) < TOL
    del post['Q']
    dX0 = orc.dlik_exact(oh, ['meanZero'], 'ard', X, y)
    assert rel(dX, dX0) < TOL


@pytest.mark.parametrize('n,D,Q,mean', [(300, 1, 2, 'meanZero'), (700, 3, 4, 'meanSum'), (1500, 8, 6, 'meanZero')])
def test_exact_covsm_vs_oracle(built_lib, n, D, Q, mean):
    """infExact with the multivariate spectral-mixture kernel (covSM.m:124-156, Q > 0): nlZ, dnlZ, alpha, predictions
    at 1e-8 against the oracle restatement of covSMfast.  dnlZ/dX is an extension (the reference ships no dcovSM.m, so
    its dlikExact cannot run with covSM): checked against the oracle's contraction of covSMfast's own x-derivative."""
    import gpml_exact as orc
    from kgp_b200 import B200GP
    rng = np.random.RandomState(100 + n)
    X = rng.randn(n, D) * 0.7; y = np.sin(2 * X.sum(1)) + 0.2 * rng.randn(n); Xs = rng.randn(90, D) * 0.7
    blocks = [np.log(rng.rand(Q, 1) + 0.5), np.log(0.3 * rng.rand(D * Q, 1) + 0.02), np.log(0.3 * rng.rand(D * Q, 1) + 0.05)]
    hc = np.concatenate([b.ravel() for b in blocks])
    hm = rng.randn(D + 1) * 0.1 if mean == 'meanSum' else np.zeros(0)
    gp = B200GP()
    gp.configure(D, {'lik': np.log(0.25), 'mean': [[v] for v in hm], 'cov': blocks}, {}, 'infExact', mean, 'covSM', 'likGauss', 'dlikExact',
                 cov_args=[Q], mean_args=['{@meanLinear, @meanConst}'] if mean == 'meanSum' else None, verbose=0)
    nlZ, g, dX = full_eval(gp, X, y)
    ymu, ys2 = gp.predict(Xs, return_var=True)
    oh = {'cov': hc, 'lik': np.log(0.25), 'mean': hm}
    om = ['meanLinear', 'meanConst'] if mean == 'meanSum' else ['meanZero']
    post, nlZ0, d0 = orc.inf_exact(oh, om, ('sm', Q), X, y)
    g0 = np.concatenate([d0['cov'], [d0['lik']], d0['mean']])
    p0 = orc.gp_predict(oh, om, ('sm', Q), X, y, Xs, post)
    assert rel(nlZ, nlZ0) < TOL and rel(g, g0) < TOL and rel(gp.alpha(), post['alpha']) < TOL
    assert rel(ymu[:, 0], p0[0]) < TOL and rel(ys2[:, 0], p0[1]) < TOL
    assert rel(dX, orc.dlik_exact_sm(oh, om, Q, X, y)) < TOL
