"""CPU FP64 oracle for the exact (infExact) GP path -- TEST INFRASTRUCTURE ONLY.

This file is a NumPy/SciPy restatement of the reference's GPML-on-Octave
arithmetic for the path behind ``kgp.backend`` (exact inference).  It is the
parity checker for the CUDA engine and the timed CPU baseline of ``bench.py``;
nothing under ``kgp_b200/`` may import it.

Reference restated (``GPML/`` = ``kgp/backend/gpml/`` of alshedivat/keras-gp):
  covariance   GPML/cov/covMaha.m:83-98,125-191, covSE.m:23-24, covScale.m:49-93
  inference    GPML/inf/infGaussLik.m:19-33, GPML/cov/apx.m:92-98,141-173
  solve        GPML/util/solve_chol.m:25
  prediction   GPML/gp.m:128-181, GPML/lik/likGauss.m:25-40
  means        GPML/mean/{meanZero,meanConst,meanLinear,meanSum}.m
  input grad   kgp/backend/utils/{dlikExact,dcovSEiso,dcovSEard}.m
  KAT helpers  GPML/doc/gpml_randn.m:18-37, GPML/cov/covMatern.m (oracle only)

Parity status: pinned (loosely) on GPML's two documented known-answer values
(GPML/doc/index.html:971,1019-1021) -- see tests/test_oracle_kat.py.  The KGP
layer's own tests assert no numbers, so beyond the KATs this restatement *is*
the oracle ("parity unpinned" for everything the KATs do not reach).
"""
import numpy as np
import scipy.linalg as sla


# ----------------------------------------------------------------------------
# means  (GPML/mean/*.m).  A mean spec is a list of names, e.g. ['meanZero'] or
# ['meanLinear', 'meanConst'] (= {@meanSum,{@meanLinear,@meanConst}}).
# ----------------------------------------------------------------------------
def mean_nhyp(mean, D):
    n = 0
    for name in mean:
        n += {'meanZero': 0, 'meanConst': 1, 'meanLinear': D}[name]
    return n


def mean_eval(mean, hyp_mean, x):
    """returns m (n,), dm(q)->(n_mean,)   [meanSum.m:13-37]"""
    n, D = x.shape
    hyp_mean = np.asarray(hyp_mean, dtype=np.float64).reshape(-1)
    m = np.zeros(n)
    parts = []
    o = 0
    for name in mean:
        if name == 'meanZero':
            parts.append(lambda q: np.zeros(0))
        elif name == 'meanConst':
            c = hyp_mean[o]; o += 1
            m = m + c
            parts.append(lambda q: np.array([q.sum()]))
        elif name == 'meanLinear':
            w = hyp_mean[o:o + D]; o += D
            m = m + x @ w
            parts.append(lambda q: x.T @ q)
        else:
            raise ValueError(name)
    assert o == hyp_mean.size, 'Number of mean hyperparameters disagree'
    dm = lambda q: np.concatenate([p(q) for p in parts]) if parts else np.zeros(0)
    return m, dm


# ----------------------------------------------------------------------------
# Mahalanobis distance  (covMaha.m:125-154)
# ----------------------------------------------------------------------------
def _metric(mode, hyp_len, D):
    if mode == 'iso':
        return np.full(D, np.exp(-2.0 * hyp_len[0]))
    return np.exp(-2.0 * np.asarray(hyp_len))


def maha(x, a, z=None):
    """D2 = max(s_i + s_j - 2 x A z', 0) on mean-centred inputs; a = diag of A."""
    n = x.shape[0]
    if z is None:
        mu = x.mean(axis=0)
        xc = x - mu
        ax = xc * a
        sax = (xc * ax).sum(1)
        D2 = sax[:, None] + (sax[None, :] - 2.0 * (xc @ ax.T))
        return np.maximum(D2, 0.0), xc, xc
    m = z.shape[0]
    mu = (m / (n + m)) * z.mean(axis=0) + (n / (n + m)) * x.mean(axis=0)
    zc = z - mu
    xc = x - mu
    ax = xc * a
    az = zc * a
    sax = (xc * ax).sum(1)
    saz = (zc * az).sum(1)
    D2 = sax[:, None] + (saz[None, :] - 2.0 * (xc @ az.T))
    return np.maximum(D2, 0.0), xc, zc


def cov_se(mode, hyp_cov, x, z=None):
    """covSEiso / covSEard: K = sf^2 exp(-D2/2)  (covSE.m:23, covScale.m:74).

    hyp_cov = [log ell (1 or D entries), log sf].  z may be None (K(x,x)),
    'diag' (vector sf^2) or an array (cross covariance)."""
    hyp_cov = np.asarray(hyp_cov, dtype=np.float64).reshape(-1)
    n, D = x.shape
    nl = 1 if mode == 'iso' else D
    assert hyp_cov.size == nl + 1, 'Wrong number of hyperparameters'
    sf2 = np.exp(2.0 * hyp_cov[nl])
    if isinstance(z, str) and z == 'diag':
        return sf2 * np.ones(n)
    a = _metric(mode, hyp_cov[:nl], D)
    D2, _, _ = maha(x, a, z)
    return sf2 * np.exp(-D2 / 2.0)


def cov_se_dirder(mode, hyp_cov, x, Q):
    """dK(Q): d tr(Q' K)/d hyp_cov for K = covSE{iso,ard}(x,x).

    covScale.m:82-93 + covMaha.m:100-119,156-191 restated.  Q is n x n (need
    not be symmetric)."""
    hyp_cov = np.asarray(hyp_cov, dtype=np.float64).reshape(-1)
    n, D = x.shape
    nl = 1 if mode == 'iso' else D
    sf2 = np.exp(2.0 * hyp_cov[nl])
    a = _metric(mode, hyp_cov[:nl], D)
    D2, xc, _ = maha(x, a, None)
    K0 = np.exp(-D2 / 2.0)
    R = (-0.5 * K0) * (Q * sf2)                      # covMaha.m:101 with Q.*S
    if mode == 'iso':
        dlen = np.array([-2.0 * np.dot(R.ravel(), D2.ravel())])   # covMaha.m:104
    else:
        q2 = R.sum(1); q1 = R.sum(0)
        y = (q1 + q2)[:, None] * xc - (R + R.T) @ xc              # covMaha.m:169
        dAdiag = (xc * y).sum(0)                                   # :172/:174
        dlen = -2.0 * a * dAdiag                                   # covMaha.m:75
    dsf = 2.0 * sf2 * (Q * K0).sum()                 # covScale.m:87-93 (scalar sf)
    return np.concatenate([dlen, [dsf]])


# ----------------------------------------------------------------------------
# spectral mixture, multivariate form (Q > 0): covSM.m:124-156 covSMfast
#   hyp = [log w (Q); log m (D x Q, column-major); log sqrt(v) (D x Q)]
# covmode ('sm', Q) selects it in inf_exact / gp_predict.
# ----------------------------------------------------------------------------
def _sm_unpack(Q, hyp, D):
    hyp = np.asarray(hyp, dtype=np.float64).reshape(-1)
    assert hyp.size == (1 + 2 * D) * Q, 'Wrong number of hyperparameters'
    w = np.exp(hyp[:Q])
    m = np.exp(hyp[Q:Q + D * Q].reshape(Q, D).T)                 # reshape(.,D,Q), column-major
    v = np.exp(2.0 * hyp[Q + D * Q:Q + 2 * D * Q].reshape(Q, D).T)
    return w, m, v


def cov_sm(Q, hyp, x, z=None):
    """K = sum_q w_q exp(-0.5 sum_d T_d^2 v_dq) cos(sum_d T_d m_dq), T = 2 pi (x - z)   (covSM.m:127-149)"""
    n, D = x.shape
    w, m, v = _sm_unpack(Q, hyp, D)
    if isinstance(z, str) and z == 'diag':
        return np.full(n, w.sum())
    zz = x if z is None else z
    K = np.zeros((n, zz.shape[0]))
    for q in range(Q):                                # q by q: the reference's (n^2 x Q) intermediates do not fit at large n
        T2 = np.zeros_like(K); Tm = np.zeros_like(K)
        for d in range(D):
            Td = 2 * np.pi * (x[:, d][:, None] - zz[:, d][None, :])
            T2 += Td * Td * v[d, q]; Tm += Td * m[d, q]
        K += w[q] * np.exp(-0.5 * T2) * np.cos(Tm)
    return K


def cov_sm_dirder(Q, hyp, x, R, want_dx=False):
    """dKdhyp(R) of covSMfast (covSM.m:150-153) and, as in dirder_fast (covSM.m:158-168), dx (n x D)."""
    n, D = x.shape
    w, m, v = _sm_unpack(Q, hyp, D)
    dw = np.zeros(Q); dm = np.zeros((D, Q)); dv = np.zeros((D, Q)); dx = np.zeros((n, D))
    T = [2 * np.pi * (x[:, d][:, None] - x[:, d][None, :]) for d in range(D)]
    for q in range(Q):
        T2 = sum(T[d] * T[d] * v[d, q] for d in range(D)); Tm = sum(T[d] * m[d, q] for d in range(D))
        E = np.exp(-0.5 * T2); H = E * np.cos(Tm); S = E * np.sin(Tm)
        dw[q] = (H * R).sum() * w[q]
        for d in range(D):
            dm[d, q] = -((S * R * w[q]) * T[d]).sum() * m[d, q]
            dv[d, q] = -((H * R * w[q]) * T[d] * T[d]).sum() * v[d, q]
            if want_dx:
                A = (R * w[q]) * S * m[d, q] + (R * w[q]) * H * v[d, q] * T[d]
                dx[:, d] += -2 * np.pi * A.sum(1) + 2 * np.pi * A.sum(0)
    dhyp = np.concatenate([dw, dm.T.reshape(-1), dv.T.reshape(-1)])
    return (dhyp, dx) if want_dx else dhyp


# Matern (oracle only; used by KAT-1)   covMatern.m / covMaterniso.m
def cov_matern_iso(d, hyp_cov, x, z=None):
    ell, sf = np.exp(hyp_cov[0]), np.exp(hyp_cov[1])
    D = x.shape[1]
    D2, _, _ = maha(x, np.full(D, ell ** -2), z)
    t = np.sqrt(d * D2)
    f = {1: lambda t: 1.0, 3: lambda t: 1 + t, 5: lambda t: 1 + t * (1 + t / 3)}[d]
    return sf ** 2 * f(t) * np.exp(-t)


# ----------------------------------------------------------------------------
# exact inference  (infGaussLik.m + apx.m ldB2_exact)
# ----------------------------------------------------------------------------
class NotPD(Exception):
    pass


def _chol_upper(B):
    try:
        return sla.cholesky(B, lower=False, overwrite_a=True, check_finite=False)
    except sla.LinAlgError as e:       # Octave: "chol: input matrix must be positive definite"
        raise NotPD(str(e))


def solve_chol(L, B):
    """X = L\\(L'\\B), L upper (solve_chol.m:25)."""
    t = sla.solve_triangular(L, B, trans='T', lower=False, check_finite=False)
    return sla.solve_triangular(L, t, trans='N', lower=False, check_finite=False)


def inf_exact(hyp, mean, covmode, x, y, want_grad=True, K=None, dK=None):
    """[post nlZ dnlZ] = infGaussLik(hyp, mean, cov, 'likGauss', x, y).

    hyp = {'cov': array, 'lik': float, 'mean': array}; covmode in {'iso','ard'}
    (or None with explicit K/dK callables for the KATs).  Like the reference
    this always forms Q = B^-1 (apx.m:163 executes because infGaussLik.m:23
    requests five outputs).  Returns post dict, nlZ, dnlZ dict."""
    x = np.asarray(x, dtype=np.float64)
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    n, D = x.shape
    m, dm = mean_eval(mean, hyp.get('mean', []), x)
    sn2 = np.exp(2.0 * float(hyp['lik']))
    if K is None and isinstance(covmode, tuple):            # ('sm', Q): covSM.m multivariate spectral mixture
        K = cov_sm(covmode[1], hyp['cov'], x)
        dK = lambda R: cov_sm_dirder(covmode[1], hyp['cov'], x, R)
    elif K is None:
        K = cov_se(covmode, hyp['cov'], x)
        dK = lambda Q: cov_se_dirder(covmode, hyp['cov'], x, Q)
    sW = 1.0 / np.sqrt(sn2)
    B = K / sn2
    B[np.diag_indices(n)] += 1.0                      # apx.m:160
    L = _chol_upper(B)
    ldB2 = np.log(np.diag(L)).sum()                   # apx.m:161
    alpha = solve_chol(L, (y - m) * sW) * sW          # apx.m:162, infGaussLik.m:25
    post = {'alpha': alpha, 'L': L, 'sW': np.full(n, sW)}
    nlZ = (y - m) @ alpha / 2.0 + ldB2 + n * np.log(2 * np.pi * sn2) / 2.0
    if not want_grad:
        return post, nlZ, None
    Q = solve_chol(L, np.eye(n))                      # apx.m:163 (sW cancels: scalar W)
    dW = (Q * K).sum(1) / 2.0                         # apx.m:166
    R = Q / sn2 - np.outer(alpha, alpha)              # apx.m:171-173
    dnlZ = {'cov': dK(R) / 2.0,
            'mean': -dm(alpha),
            'lik': -sn2 * (alpha @ alpha) - 2.0 * dW.sum() / sn2 + n}
    post['Q'] = Q
    return post, nlZ, dnlZ


def dlik_exact(hyp, mean, covmode, x, y, ard_reference_form=False):
    """dnlZ/dX (n x D)  -- kgp/backend/utils/dlikExact.m:9-25.

    The reference forms the n x n x D tensor of dcovSE{iso,ard}.m; here the
    contraction over j is done dimension by dimension (identical arithmetic,
    no tensor).  ard_reference_form=True reproduces dcovSEard.m:7,16, which
    divides every dimension by ell_1^2; False uses ell_d."""
    x = np.asarray(x, dtype=np.float64)
    n, D = x.shape
    sn2 = np.exp(2.0 * float(hyp['lik']))
    sl = 1.0 if sn2 < 1e-6 else sn2                   # dlikExact.m:11
    post, _, _ = inf_exact(hyp, mean, covmode, x, y, want_grad=False)
    alpha, L = post['alpha'], post['L']
    dlik_dcov = solve_chol(L, np.eye(n)) / sl - np.outer(alpha, alpha)   # :19
    K = cov_se(covmode, hyp['cov'], x)
    hc = np.asarray(hyp['cov'], dtype=np.float64).reshape(-1)
    if covmode == 'iso' or ard_reference_form:
        ell2 = np.full(D, np.exp(2.0 * hc[0]))
    else:
        ell2 = np.exp(2.0 * hc[:D])
    G = dlik_dcov * K
    r = G.sum(1)
    # sum_j G_ij * (-(x_id - x_jd)/ell_d^2)
    return -(r[:, None] * x - G @ x) / ell2[None, :]


def dlik_exact_sm(hyp, mean, Q, x, y):
    """dnlZ/dX for exact covSM -- an EXTENSION: the reference ships no dcovSM.m, so dlikExact cannot run with covSM
    there.  Same contraction as dlikExact.m:19-25 with the x-derivative of covSMfast (covSM.m:158-168):
    dnlZ/dx = 1/2 dx(R), R = K^-1 - alpha alpha' (the factor 1/2 because dirder_fast's dx counts both arguments)."""
    x = np.asarray(x, dtype=np.float64)
    n = x.shape[0]
    sn2 = np.exp(2.0 * float(hyp['lik']))
    post, _, _ = inf_exact(hyp, mean, ('sm', Q), x, y, want_grad=False)
    R = solve_chol(post['L'], np.eye(n)) / sn2 - np.outer(post['alpha'], post['alpha'])
    return cov_sm_dirder(Q, hyp['cov'], x, R, want_dx=True)[1] / 2.0


def gp_predict(hyp, mean, covmode, x, y, xs, post=None):
    """[ymu ys2 fmu fs2] = gp(hyp, inf, mean, cov, lik, x, y, xs)  (gp.m:128-181)."""
    x = np.asarray(x, dtype=np.float64)
    xs = np.asarray(xs, dtype=np.float64)
    if post is None:
        post, _, _ = inf_exact(hyp, mean, covmode, x, y, want_grad=False)
    alpha, L, sW = post['alpha'], post['L'], post['sW']
    sn2 = np.exp(2.0 * float(hyp['lik']))
    ns = xs.shape[0]
    fmu = np.zeros(ns); fs2 = np.zeros(ns)
    for s in range(0, ns, 1000):                      # gp.m:142-146
        idx = slice(s, min(s + 1000, ns))
        if isinstance(covmode, tuple):
            kss = cov_sm(covmode[1], hyp['cov'], xs[idx], 'diag')
            Ks = cov_sm(covmode[1], hyp['cov'], x, xs[idx])
        else:
            kss = cov_se(covmode, hyp['cov'], xs[idx], 'diag')
            Ks = cov_se(covmode, hyp['cov'], x, xs[idx])
        ms, _ = mean_eval(mean, hyp.get('mean', []), xs[idx])
        fmu[idx] = ms + Ks.T @ alpha
        V = sla.solve_triangular(L, sW[:, None] * Ks, trans='T', lower=False,
                                 check_finite=False)
        fs2[idx] = np.maximum(kss - (V * V).sum(0), 0.0)
    return fmu.copy(), fs2 + sn2, fmu, fs2            # likGauss.m:34-39


# ----------------------------------------------------------------------------
# GPML's portable PRNG (doc/gpml_randn.m:18-37) -- data for the KATs
# ----------------------------------------------------------------------------
def gpml_randn(seed, n):
    N = int(np.ceil(n / 2) * 2)
    a = 7 ** 5; m = 2 ** 31 - 1
    q = m // a; r = m % a
    u = np.zeros(N + 1)
    u[0] = np.fix(seed * 2 ** 31)
    for i in range(1, N + 1):
        u[i] = a * (u[i - 1] % q) - r * np.fix(u[i - 1] / q)
        if u[i] < 0:
            u[i] += m
    u = u[1:] / 2 ** 31
    w = np.sqrt(-2 * np.log(u[:N // 2]))
    xx = np.concatenate([w * np.cos(2 * np.pi * u[N // 2:]), w * np.sin(2 * np.pi * u[N // 2:])])
    return xx[:n]
