"""CPU FP64 oracle for the KISS-GP / MSGP (infGrid) path -- TEST INFRASTRUCTURE ONLY.

NumPy/SciPy restatement of the reference's grid inference; parity checker for the
CUDA engine and timed CPU baseline.  Nothing under kgp_b200/ may import it.

Reference restated (GPML/ = kgp/backend/gpml/ of alshedivat/keras-gp):
  grid creation      GPML/cov/apxGrid.m:389-416 (creategrid; eq=0: kmeans -> deterministic Lloyd stand-in, see kmeans1d)
  interpolation      GPML/cov/apxGrid.m:440-490 (interpgrid), 495-523 (eqinterp), 527-539 (kcub/dkcub)
  Kronecker MVM      GPML/cov/apxGrid.m:302-329 (kronmvm), dense factors (plain numeric axes)
  CG                 GPML/cov/apx.m:275-302 (conjgrad/linsolve), 218-226 (ldB2_grid)
  Fiedler log-det    GPML/cov/apxGrid.m:571-588 (eigkron), GPML/cov/apx.m:312-318, 246-269
  Chebyshev log-det  GPML/cov/apx.m:336-391 (probes and emax are explicit inputs)
  hyper-derivatives  GPML/cov/apxGrid.m:631-652 (dirder), GPML/cov/apx.m:271-273
  inference          GPML/inf/infGrid.m:53-89,131-145, GPML/inf/infGaussLik.m:19-33
  factor kernels     covSEiso on a 1-D axis (covMaha/covSE/covScale), GPML/cov/covSM.m:124-156
  input gradient     kgp/backend/utils/dlikGrid.m:8-40, deriv_x.m:4-26

Parity status: "parity unpinned" -- the reference holds no numeric value for infGrid
anywhere; this restatement is checked by self-consistency (finite differences,
explicit Kronecker products, dense solves) in tests/test_oracle_grid.py.
"""
import numpy as np
import scipy.sparse as sp

from gpml_exact import mean_eval


# ----------------------------------------------------------------------------
# grid creation (apxGrid.m:389-416)
# ----------------------------------------------------------------------------
def kmeans1d(u, K):
    """Stand-in for `[idx,xgj] = kmeans(u, K)` (apxGrid.m:411) on the sorted unique values u.  Octave's kmeans (statistics
    package) starts from random centres, so no implementation can reproduce its output; this is Lloyd's algorithm from an
    equal-count initial partition (deterministic).  In 1-D the nearest-centre rule is a set of thresholds at the midpoints of
    neighbouring centres (a tie goes to the lower cluster, as Octave's min does); segment means come from sequential prefix
    sums, so the engine's host routine (kgp_kmeans1d_host) agrees to the last bit."""
    u = np.asarray(u, dtype=np.float64)
    nu = u.size
    assert 1 <= K <= nu
    S = np.concatenate([[0.0], np.cumsum(u)])          # sequential sums
    bnd = [(c * nu) // K for c in range(K + 1)]
    cent = [(S[bnd[c + 1]] - S[bnd[c]]) / float(bnd[c + 1] - bnd[c]) for c in range(K)]
    for _ in range(200):
        nb = [0] * (K + 1)
        nb[K] = nu
        for c in range(K - 1):
            nb[c + 1] = max(int(np.searchsorted(u, 0.5 * (cent[c] + cent[c + 1]), side='right')), nb[c])
        if nb == bnd:
            break
        bnd = nb
        for c in range(K):
            if bnd[c + 1] > bnd[c]:                    # an emptied cluster keeps its centre
                cent[c] = (S[bnd[c + 1]] - S[bnd[c]]) / float(bnd[c + 1] - bnd[c])
    return np.array(cent)


def _linspace(lo, hi, n):
    """Octave linspace: base + i*delta, the last point exact."""
    v = lo + np.arange(n, dtype=np.float64) * ((hi - lo) / float(n - 1) if n > 1 else 0.0)
    if n > 0:
        v[-1] = hi
    return v


def creategrid(x, eq=1, k=1.0):
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    xg = []
    for j in range(x.shape[1]):
        u = np.unique(x[:, j])
        if u.size < 2:
            raise ValueError('Two few unique points.')
        if 0 <= k <= 1:
            ngj = float(np.ceil(k * u.size))
        else:
            ngj = float(k)                           # stays fractional in du (apxGrid.m:404-406)
        du = (u[-1] - u[0]) / ngj
        bu = (u[0] - 5 * du, u[-1] + 5 * du)
        if eq:
            xg.append(np.linspace(bu[0], bu[1], int(np.floor(max(ngj, 5)))))   # Octave linspace floors n
        else:                                        # apxGrid.m:410-415: cluster centres + boundary nodes
            ngi = int(np.floor(ngj))
            xgj = kmeans1d(u, min(u.size, ngi - 2))
            nb = ngi - xgj.size; nb1 = nb // 2; nb2 = nb - nb1
            xg1 = _linspace(bu[0], xgj[0], nb1 + 1); xg2 = _linspace(xgj[-1], bu[1], nb2 + 1)
            xg.append(np.concatenate([xg1[:nb1], xgj, xg2[1:]]))
    return xg


# ----------------------------------------------------------------------------
# cubic convolution interpolation (apxGrid.m:495-539)
# ----------------------------------------------------------------------------
def kcub(x):
    x = np.abs(x)
    y = np.zeros_like(x)
    q = x <= 1
    y[q] = ((1.5 * x[q] - 2.5) * x[q]) * x[q] + 1
    q = (1 < x) & (x <= 2)
    y[q] = ((-0.5 * x[q] + 2.5) * x[q] - 4) * x[q] + 2
    return y


def dkcub(x):
    y = np.sign(x); x = np.abs(x)
    q = x <= 1
    y[q] = y[q] * (4.5 * x[q] - 5.0) * x[q]
    q = (1 < x) & (x <= 2)
    y[q] = y[q] * ((-1.5 * x[q] + 5.0) * x[q] - 4.0)
    y[x > 2] = 0
    return y


def eqinterp(s, t, deg=3):
    """J (nt,nc) 0-based node indices, C (nt,nc) weights, dC (nt,nc) derivative weights * (ns-1); apxGrid.m:495-523.
    deg = 3: Keys cubic on 4 nodes (the default, apx.m:122); deg = 1: linear, deg = 0: nearest node, both on 2 nodes
    (apxGrid.m:498-500).  The GP-interpolation branch (deg > 3, apxGrid.m:501-504) is not restated."""
    s = np.asarray(s, dtype=np.float64); t = np.asarray(t, dtype=np.float64)
    if deg == 3:
        it = np.array([-2, -1, 0, 1]); k, dk = kcub, dkcub
    elif deg == 1:
        it = np.array([-1, 0]); k = lambda x: np.maximum(1.0 - np.abs(x), 0.0); dk = lambda x: -(np.abs(x) <= 1).astype(np.float64) * np.sign(x)
    elif deg == 0:
        it = np.array([-1, 0]); k = lambda x: ((-0.5 < x) & (x <= 0.5)).astype(np.float64); dk = lambda x: 0.0 * x
    else:
        raise ValueError('interpolation degree %r is not restated (0, 1, 3)' % (deg,))
    nc = it.size
    ds = s[1] - s[0]; ns = s.size
    assert ns >= nc
    j = np.floor((t - s[0]) / ds) + 1
    w = (t - s[0]) / ds - j + 1
    j = j - it[nc - 1]
    C = np.zeros((t.size, nc)); dC = np.zeros((t.size, nc))
    for i in range(nc):
        C[:, i] = k(w + it[nc - 1 - i])
        dC[:, i] = dk(w + it[nc - 1 - i]) * (ns - 1)
    idx = np.where(j < nc + it[0])[0]                     # lower boundary (:515-518)
    C[idx] = 0; dC[idx] = 0
    if idx.size:
        jid = np.argmin(np.abs(s[None, :nc] - t[idx, None]), axis=1)
        C[idx, jid] = 1
    j[idx] = 1
    v = ns - nc + 1
    idx = np.where(j > v)[0]                              # upper boundary (:519-522)
    C[idx] = 0; dC[idx] = 0
    if idx.size:
        jid = np.argmin(np.abs(s[None, ns - nc:] - t[idx, None]), axis=1)
        C[idx, jid] = 1
    j[idx] = v
    J = (j[:, None] + np.arange(nc)[None, :] - 1).astype(np.int64)   # 0-based
    return J, C, dC


def equi(ax):
    """apxGrid.m:684-695: the axis is equispaced when max |diff(ax) - (ax[1] - ax[0])| < 1e-9."""
    ax = np.asarray(ax, dtype=np.float64)
    return ax.size < 2 or float(np.max(np.abs(np.diff(ax) - (ax[1] - ax[0])))) < 1e-9


def neqinterp(s, t):
    """apxGrid.m:547-557: piecewise linear interpolation on a non-equispaced (increasing) axis; J (nt,2), C, dC.
    The reference sorts s and maps back through `ord`; only increasing axes are restated (ord = identity)."""
    s = np.asarray(s, dtype=np.float64); t = np.asarray(t, dtype=np.float64)
    ns = s.size
    assert ns >= 2 and np.all(np.diff(s) >= 1e-10), 'Some source points are equal.'
    ii = np.searchsorted(s[1:ns - 1], t, side='right')            # histc(t, [-inf; s(2:end-1); inf]) - 1
    d0 = t - s[ii]; d1 = s[ii + 1] - t
    d0n = d0 < 0; d1n = d1 < 0
    d0 = np.where(d0n, 0.0, d0); d1 = np.where(d1n, 0.0, d1)
    J = np.stack([ii, ii + 1], axis=1).astype(np.int64)
    C = np.stack([d1 / (d1 + d0), d0 / (d1 + d0)], axis=1)
    nz = 1.0 - (d1n | d0n)
    dC = np.stack([-nz / (d1 + d0), nz / (d1 + d0)], axis=1)
    return J, C, dC


def interpgrid(xg, x, want_deriv=False, deg=3):
    """Sparse Mx (n x N) and optionally dMx[i]; grid linear index has the first axis fastest.  deg: interpolation degree of the
    equispaced axes (opt.deg, apxGrid.m:448); non-equispaced axes ignore it (apx.m:28-32)."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    n, p = x.shape
    ng = [a.size for a in xg]; N = int(np.prod(ng))
    J = np.zeros((n, 1), dtype=np.int64); C = np.ones((n, 1))
    dC = [np.ones((n, 1)) for _ in range(p)]
    s = 1
    for i in range(p):
        di = int(np.atleast_1d(deg)[i % np.size(deg)])                                  # deg = deg(:).*ones(p,1) (apxGrid.m:448)
        Ji, Ci, dCi = eqinterp(xg[i], x[:, i], di) if equi(xg[i]) else neqinterp(xg[i], x[:, i])    # apxGrid.m:453-465
        J = (J[:, :, None] + s * Ji[:, None, :]).reshape(n, -1)
        C = (C[:, :, None] * Ci[:, None, :]).reshape(n, -1)
        for jd in range(p):
            dC[jd] = (dC[jd][:, :, None] * (dCi if jd == i else Ci)[:, None, :]).reshape(n, -1)
        s *= ng[i]
    I = np.repeat(np.arange(n)[:, None], C.shape[1], axis=1)
    Mx = sp.csr_matrix((C.ravel(), (I.ravel(), J.ravel())), shape=(n, N))
    if not want_deriv:
        return Mx
    dMx = [sp.csr_matrix((d.ravel(), (I.ravel(), J.ravel())), shape=(n, N)) for d in dC]
    return Mx, dMx


# ----------------------------------------------------------------------------
# 1-D factor kernels on a grid axis and their directional derivatives
# ----------------------------------------------------------------------------
def factor_nhyp(kind, Q=0):
    return 2 if kind == 'covSEiso' else 3 * Q


def factor_cov(kind, hyp, ax, Q=0):
    """returns K (ng x ng) and dK(R) -> d tr(R'K)/dhyp   (R need not be symmetric)."""
    hyp = np.asarray(hyp, dtype=np.float64).reshape(-1)
    ax = np.asarray(ax, dtype=np.float64)
    if kind == 'covSEiso':
        ell, sf2 = np.exp(hyp[0]), np.exp(2 * hyp[1])
        xc = (ax - ax.mean()) / ell                       # covMaha.m:138-143 centring
        sq = xc * xc
        D2 = np.maximum(sq[:, None] + (sq[None, :] - 2.0 * np.outer(xc, xc)), 0.0)
        K0 = np.exp(-D2 / 2.0)
        K = sf2 * K0

        def dK(R):
            return np.array([np.sum(R * K * D2), 2.0 * np.sum(R * K)])   # covMaha.m:104, covScale.m:87-93
        return K, dK
    if kind == 'covSM':                                   # covSM.m:124-156, D=1
        w = np.exp(hyp[:Q]); m = np.exp(hyp[Q:2 * Q]); v = np.exp(2 * hyp[2 * Q:3 * Q])
        T = 2 * np.pi * (ax[:, None] - ax[None, :])
        E = np.exp(-0.5 * (T * T)[:, :, None] * v[None, None, :])
        Tm = T[:, :, None] * m[None, None, :]
        H = E * np.cos(Tm)
        K = H @ w

        def dK(R):
            dw = np.einsum('ij,ijq->q', R, H) * w
            dm = -np.einsum('ij,ijq,ij->q', R, E * np.sin(Tm), T) * w * m
            dv = -np.einsum('ij,ijq,ij->q', R, H, T * T) * w * v
            return np.concatenate([dw, dm, dv])
        return K, dK
    raise ValueError(kind)


def factor_diag(kind, hyp, Q=0):
    """k(x,x) of one factor (covGrid 'diag' mode, apxGrid.m:239-249)."""
    hyp = np.asarray(hyp, dtype=np.float64).reshape(-1)
    return np.exp(2 * hyp[1]) if kind == 'covSEiso' else np.exp(hyp[:Q]).sum()


# ----------------------------------------------------------------------------
# Toeplitz factors: xg{i} given as a nested cell {axis} with an equispaced axis (apxGrid.m:258-281).  GPML then keeps
# only the covariance as a function of the lag (kii), multiplies by FFT through a Strang circulant embedding of length
# 2 ng - 1 (bttbmvmsymfft, apxGrid.m:335-346), takes the Fiedler eigenvalues from a WHITTLE circulant embedding
# (circ with s = 3, apxGrid.m:576-580,617-627; apx.m:236) -- which are not the eigenvalues of the Toeplitz matrix, so
# ldB2 differs from the dense-factor value on the same axis -- and evaluates all hyper-derivatives as sums over lags
# (apx.m:256-266, apxGrid.m:642-650).  Only 1-D sub-grids (Toeplitz, not BTTB) are restated.
# In this module such a grid entry is a list / tuple holding the one axis array:  xg = [[ax1], ax2]  <->  {{ax1}, ax2}.
# ----------------------------------------------------------------------------
def is_toep(xgi):
    return isinstance(xgi, (list, tuple))


def axis_of(xgi):
    if is_toep(xgi):
        assert len(xgi) == 1, 'only 1-D Toeplitz sub-grids are restated (no BTTB)'
        return np.asarray(xgi[0], dtype=np.float64).reshape(-1)
    return np.asarray(xgi, dtype=np.float64).reshape(-1)


def flat_grid(xg):
    return [axis_of(a) for a in xg]


class ToepFactor(object):
    """K.kron(i).factor of a Toeplitz grid dimension: struct('descr','toep','mvm',..,'kii',..,'size',..)."""

    def __init__(self, kind, hyp, ax, Q=0):
        self.kind, self.Q = kind, Q
        self.hyp = np.asarray(hyp, dtype=np.float64).reshape(-1)
        ax = np.asarray(ax, dtype=np.float64)
        dev = np.abs(np.diff(ax) - (ax[1] - ax[0]))
        if dev.max() >= 1e-9:
            raise ValueError('Subgrid not equispaced.')                        # apxGrid.m:262-263
        self.n = ax.size
        self.h = (ax[-1] - ax[0]) / (self.n - 1)                               # wi./(ni-1), apxGrid.m:265
        xc = np.concatenate([np.arange(1, self.n + 1), np.arange(2 - self.n, 1)])   # Strang embedding grid, apxGrid.m:267
        self.xc = xc
        self.fi = np.real(np.fft.fft(self.kii(xc)))                            # apxGrid.m:268-269
        self.shape = (self.n, self.n)

    def _parts(self, idx):
        tau = (np.asarray(idx, dtype=np.float64) - 1.0) * self.h              # kii(n) = k((n-1) h, 0)
        if self.kind == 'covSEiso':
            ell, sf2 = np.exp(self.hyp[0]), np.exp(2 * self.hyp[1])
            d2 = (tau / ell) ** 2
            return tau, d2, sf2 * np.exp(-d2 / 2.0)
        Q = self.Q
        w = np.exp(self.hyp[:Q]); m = np.exp(self.hyp[Q:2 * Q]); v = np.exp(2 * self.hyp[2 * Q:3 * Q])
        T = 2 * np.pi * tau
        E = np.exp(-0.5 * (T * T)[:, None] * v[None, :]); Tm = T[:, None] * m[None, :]
        return T, (E, Tm, w, m, v), (E * np.cos(Tm)) @ w

    def kii(self, idx):
        return self._parts(idx)[2]

    def dkii(self, idx, R):
        """second output of kii: d sum(R .* k(lags)) / d hyp  (the dK functional of the cross-covariance call)."""
        R = np.asarray(R, dtype=np.float64).reshape(-1)
        if self.kind == 'covSEiso':
            _, d2, k = self._parts(idx)
            return np.array([np.sum(R * k * d2), 2.0 * np.sum(R * k)])
        T, (E, Tm, w, m, v), _ = self._parts(idx)
        H = E * np.cos(Tm)
        dw = (R @ H) * w
        dm = -((R * T) @ (E * np.sin(Tm))) * w * m
        dv = -((R * T * T) @ H) * w * v
        return np.concatenate([dw, dm, dv])

    def mvm(self, A):
        """bttbmvmsymfft for p = 1 (apxGrid.m:335-346): A is (n, m)."""
        L = 2 * self.n - 1
        B = np.fft.ifft(self.fi[:, None] * np.fft.fft(A, L, axis=0), axis=0)
        return np.real(B[:self.n])

    def dense(self):
        i = np.arange(self.n)
        return self.kii((np.abs(i[:, None] - i[None, :]) + 1).reshape(-1)).reshape(self.n, self.n)


def circ_whittle(f, s=3):
    """circ(kii, n, s) for p = 1 (apxGrid.m:617-627): c[j] = sum_{r=-s}^{s-1} k((j + r n) h); returns c and the lag grid."""
    n = f.n
    xw = np.arange(1 - s * n, s * n + 1)
    c = f.kii(xw).reshape(2 * s, n).sum(0)                 # reshape(.,[n, 2s]) column-major, summed over the 2nd dimension
    return c, xw


def _fsize(K):
    return K.n if isinstance(K, ToepFactor) else K.shape[0]


def split_hyp(kind, hyp_cov, p, Q=0):
    nh = factor_nhyp(kind, Q)
    hyp_cov = np.asarray(hyp_cov, dtype=np.float64).reshape(-1)
    assert hyp_cov.size == nh * p, 'Number of cov function hyperparameters disagree with cov function'
    return [hyp_cov[i * nh:(i + 1) * nh] for i in range(p)]


def kronmvm(Ks, v):
    """(K_p kron ... kron K_1) v for v (N,) or (N,m); first axis fastest (apxGrid.m:302-329)."""
    v = np.asarray(v, dtype=np.float64)
    one = v.ndim == 1
    if one:
        v = v[:, None]
    m = v.shape[1]
    ng = [_fsize(K) for K in Ks]
    t = v.T.reshape([m] + ng[::-1])                       # axes: rhs, dim p, ..., dim 1
    for i, K in enumerate(Ks):
        axn = t.ndim - 1 - i
        if isinstance(K, ToepFactor):                     # FFT product along this mode (apxGrid.m:326)
            tt = np.moveaxis(t, axn, 0); shp = tt.shape
            t = np.moveaxis(K.mvm(tt.reshape(ng[i], -1)).reshape(shp), 0, axn)
        else:
            t = np.moveaxis(np.tensordot(K, t, axes=([1], [axn])), 0, axn)
    out = t.reshape(m, -1).T
    return out[:, 0] if one else out


# ----------------------------------------------------------------------------
# conjugate gradients (apx.m:281-302)
# ----------------------------------------------------------------------------
def conjgrad(A, b, tol=1e-10, maxit=None):
    b = np.asarray(b, dtype=np.float64)
    one = b.ndim == 1
    if one:
        b = b[:, None]
    if maxit is None:
        maxit = min(b.shape[0], 20)
    x = np.zeros_like(b)
    r = b - A(x); r2 = (r * r).sum(0); r2new = r2.copy()
    nb = np.sqrt((b * b).sum(0)); flag = 0; it = 1
    with np.errstate(all='ignore'):
        relres = np.sqrt(r2) / nb
    todo = relres >= tol
    if not todo.any():
        return (x[:, 0] if one else x), 1, relres, it
    r = r[:, todo]; d = r.copy()
    for it in range(2, int(maxit) + 1):
        z = A(d)
        a = r2[todo] / (d * z).sum(0)
        x[:, todo] += a[None, :] * d
        r = r - a[None, :] * z
        r2new[todo] = (r * r).sum(0)
        relres = np.sqrt(r2new) / nb
        cnv = relres[todo] < tol; todo = relres >= tol
        d = d[:, ~cnv]; r = r[:, ~cnv]
        if not todo.any():
            flag = 1
            break
        bb = r2new / r2
        d = r + bb[todo][None, :] * d
        r2 = r2new.copy()
    return (x[:, 0] if one else x), flag, relres, it


class CGNotConverged(Exception):
    pass


# ----------------------------------------------------------------------------
# Fiedler log-det bound and its derivatives (apxGrid.m:571-588, apx.m:246-269,312-318)
# ----------------------------------------------------------------------------
def eigkron(Ks, s=3):
    V, ee = [], []
    for K in Ks:
        if isinstance(K, ToepFactor):                     # Whittle-circulant eigenvalues, V = Fourier matrix (apxGrid.m:576-580)
            ej = np.real(np.fft.fft(circ_whittle(K, s)[0]))
            V.append(None); ee.append(np.maximum(ej, 0.0))
            continue
        ej, Vj = np.linalg.eigh((K + K.T) / 2.0)
        V.append(Vj); ee.append(np.maximum(ej, 0.0))
    e = np.ones(1)
    for ej in ee:
        e = np.kron(ej, e)
    return V, ee, e


def logdet_fiedler(E, W):
    ie = np.argsort(-E, kind='stable'); iw = np.argsort(-W, kind='stable')
    Es, Ws = E[ie], W[iw]
    N, n = E.size, W.size
    k = n / N * Es
    k = np.concatenate([k, np.zeros(n - N)]) if n > N else k[:n]
    kw1 = k * Ws + 1.0
    ub = np.log(kw1).sum() / 2.0
    dW = np.zeros(n); dW[iw] = k / (2.0 * kw1)
    m = min(n, N)
    dE = np.zeros(N); dE[ie[:m]] = Ws[:m] / (N * 2.0 / n * kw1[:m])
    return ub, dE, dW


def fiedler_dhyp(Ks, dKs, V, ee, de, s=3):
    """dhyp_i = dK_i( V_i diag(z_i) V_i' ),  z_i = de contracted with the other factors' eigenvalues."""
    ng = [_fsize(K) for K in Ks]; p = len(Ks)
    out = []
    for i in range(p):
        z = de.reshape(ng[::-1])                          # axes: dim p ... dim 1
        for j in range(p):
            if j != i:
                axn = z.ndim - 1 - j
                z = np.moveaxis(np.tensordot(ee[j][None, :], z, axes=([1], [axn])), 0, axn)
        z = z.reshape(-1)
        if isinstance(Ks[i], ToepFactor):                 # apx.m:256-266: sum over the 2 s n Whittle lags
            Fz = np.real(np.fft.fft(z))
            _, xw = circ_whittle(Ks[i], s)
            out.append(Ks[i].dkii(xw, np.tile(Fz, 2 * s)))
            continue
        Q = (V[i] * z[None, :]) @ V[i].T
        out.append(dKs[i](Q))
    return np.concatenate(out)


def dirder(Ks, dKs, Mx, a, b):
    """d (a' Mx Kg Mx' b) / d hyp   (apxGrid.m:631-652), dense factors."""
    ng = [_fsize(K) for K in Ks]; p = len(Ks)
    Mta = Mx.T @ a; Mtb = Mx.T @ b
    out = []
    for i in range(p):
        v = Mta.reshape(ng[::-1])
        for j in range(p):
            if j != i:
                axn = v.ndim - 1 - j
                if isinstance(Ks[j], ToepFactor):
                    vv = np.moveaxis(v, axn, 0); shp = vv.shape
                    v = np.moveaxis(Ks[j].mvm(vv.reshape(ng[j], -1)).reshape(shp), 0, axn)
                else:
                    v = np.moveaxis(np.tensordot(Ks[j], v, axes=([1], [axn])), 0, axn)
        axi = len(ng) - 1 - i
        A = np.moveaxis(v, axi, 0).reshape(ng[i], -1)
        B = np.moveaxis(Mtb.reshape(ng[::-1]), axi, 0).reshape(ng[i], -1)
        if isinstance(Ks[i], ToepFactor):                 # apxGrid.m:642-650: lag sums of A B' through padded FFTs
            L = 2 * ng[i] - 1
            Fvb = np.fft.fft((np.fft.fft(A, L, axis=0) * np.fft.ifft(B, L, axis=0)).sum(1))
            out.append(Ks[i].dkii(Ks[i].xc, np.real(Fvb)))
            continue
        out.append(dKs[i](A @ B.T))
    return np.concatenate(out)


# ----------------------------------------------------------------------------
# Chebyshev-Hutchinson log-det (apx.m:336-391); probes V (n x m) and emax are inputs
# ----------------------------------------------------------------------------
def logdet_sample(W, Kmvm, dK, V, d, emax, want_deriv=True):
    sW = np.sqrt(W); n = W.size; emin = 1.0; m = V.shape[1]
    B0 = lambda x: x + sW[:, None] * Kmvm(sW[:, None] * x) if x.ndim == 2 else x + sW * Kmvm(sW * x)
    a = emin + emax; dl = 1 - emax / a; ldB2 = n * np.log(a) / 2.0
    sf = 2.0 / a / (1 - 2 * dl)
    B = lambda x: sf * B0(x) - 1.0 / (1 - 2 * dl) * x
    xk = np.cos(np.pi * (np.arange(d + 1) + 0.5) / (d + 1))
    fk = np.log(((1 - 2 * dl) / 2.0) * xk + 0.5)
    Tk = np.ones((d + 1, d + 1)); Tk[:, 1] = xk
    for i in range(2, d + 1):
        Tk[:, i] = 2 * xk * Tk[:, i - 1] - Tk[:, i - 2]
    c = 2.0 / (d + 1) * (Tk.T @ fk); c[0] /= 2
    p1 = np.zeros(d + 1); p1[0] = 1; p2 = np.zeros(d + 1); p2[1] = 1
    p = c[0] * p1 + c[1] * p2
    for i in range(2, d + 1):
        p3 = np.concatenate([[0.0], 2 * p2[:d]]) - p1
        p = p + c[i] * p3; p1, p2 = p2, p3
    dhyp = 0.0; dWv = np.zeros(n)
    for i in range(m):
        v = V[:, i]
        Bv = np.zeros((n, d + 1)); Bv[:, 0] = v
        for j in range(d):
            Bv[:, j + 1] = B(Bv[:, j])
        ldB2 += (v @ Bv @ p) / (2.0 * m)
        if not want_deriv:
            continue
        sWBv = sW[:, None] * Bv
        for j in range(1, d + 1):
            for k in range(1, int(np.ceil(j / 2)) + 1):
                akj = sf * p[j] / m
                if k == int(np.ceil(j / 2)) and j % 2 == 1:
                    akj /= 2
                dhyp = dhyp + akj * dK(sWBv[:, j - k], sWBv[:, k - 1])
        u = Bv / sW[:, None]; w = Kmvm(sWBv)
        for j in range(1, d + 1):
            dWj = (u[:, j - 1::-1][:, :j] * w[:, :j] + w[:, j - 1::-1][:, :j] * u[:, :j]).sum(1)
            dWv += sf * p[j] / (4.0 * m) * dWj
    return ldB2, dhyp, dWv


# ----------------------------------------------------------------------------
# Stochastic Lanczos quadrature log-det (NOT in GPML 4.0 -- an extension, checked against the
# dense identities it estimates, never reported as reference parity).  Probes V (n x m) are inputs.
#   ldB2 = 1/(2m) sum_i |v_i|^2 e1' log(T_i) e1 ,  T_i from `steps` Lanczos steps on B = I + sW K sW started at v_i
#   d ldB2/d theta = 1/(2m) sum_i dK(sW g_i, sW v_i),  g_i = B^-1 v_i   (Hutchinson estimate of tr(B^-1 dB)/2)
#   dW             = 1/(2m) sum_i (g_i ./ sW) .* (K (sW .* v_i))
# ----------------------------------------------------------------------------
def lanczos_quadrature(B, v, steps):
    """v' log(B) v by Gauss quadrature on the Lanczos tridiagonal; the basis is kept and every new vector is
    re-orthogonalised against it (one classical Gram-Schmidt pass), which makes the recurrence reproducible."""
    nv2 = float(v @ v)
    q = v / np.sqrt(nv2); qp = np.zeros_like(q); beta = 0.0
    al, be = [], []; Qm = [q]
    for j in range(steps):
        w = B(q)
        a = float(q @ w)
        w = w - a * q - beta * qp
        Qa = np.stack(Qm, axis=1)
        w = w - Qa @ (Qa.T @ w)
        b = float(np.sqrt(w @ w))
        al.append(a)
        if j + 1 == steps or not (b > 1e-12 * max(abs(a), 1.0)):
            break
        be.append(b); qp = q; q = w / b; beta = b; Qm.append(q)
    k = len(al)
    T = np.diag(al) + np.diag(be[:k - 1], 1) + np.diag(be[:k - 1], -1)
    th, S = np.linalg.eigh(T)
    return nv2 * float((S[0] ** 2) @ np.log(th)), k


def logdet_slq(W, Kmvm, dK, V, steps, solve, want_deriv=True):
    sW = np.sqrt(W); n = W.size; m = V.shape[1]
    B = lambda x: x + sW * Kmvm(sW * x)
    ldB2 = 0.0; dhyp = 0.0; dWv = np.zeros(n)
    for i in range(m):
        v = V[:, i]
        val, _ = lanczos_quadrature(B, v, steps)
        ldB2 += val / (2.0 * m)
        if want_deriv:
            g = solve(v)                                   # B^-1 v  (same conjgrad as the alpha solve)
            dhyp = dhyp + dK(sW * g, sW * v) / (2.0 * m)
            dWv += (g / sW) * Kmvm(sW * v) / (2.0 * m)
    return ldB2, dhyp, dWv


# ----------------------------------------------------------------------------
# infGrid with likGauss (infGrid.m:53-89 -> infGaussLik.m:19-33 -> apx.m grid branch)
# ----------------------------------------------------------------------------
def inf_grid(hyp, mean, kind, xg, x, y, opt=None, Q=0, want_grad=True, cheby=None, slq=None):
    """returns post {alpha, Kalpha (grid vector), cg_iter, cg_relres}, nlZ, dnlZ {cov, lik, mean}.

    opt: cg_tol (1e-6), cg_maxit (min(n,20)).  cheby = dict(V=..., d=..., emax=...) selects the
    Chebyshev estimator with injected probes (opt.ldB2_cheby), else the Fiedler bound.
    slq = dict(V=..., steps=...) selects stochastic Lanczos quadrature (extension)."""
    opt = opt or {}
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    y = np.asarray(y, dtype=np.float64).reshape(-1)
    n, D = x.shape; p = len(xg)
    assert p == D, 'Grid and data dimension are different.'
    tol = opt.get('cg_tol', 1e-6); maxit = opt.get('cg_maxit', min(n, 20))
    hs = split_hyp(kind, hyp['cov'], p, Q)
    Ks, dKs = [], []
    for i in range(p):
        if is_toep(xg[i]):
            Ks.append(ToepFactor(kind, hs[i], axis_of(xg[i]), Q)); dKs.append(None)
        else:
            K_, dK_ = factor_cov(kind, hs[i], xg[i], Q); Ks.append(K_); dKs.append(dK_)
    xg = flat_grid(xg)
    Mx = interpgrid(xg, x, deg=opt.get('deg', 3))                   # apx.m:122,131
    MxT = Mx.T.tocsr()
    m, dm = mean_eval(mean, hyp.get('mean', []), x)
    sn2 = np.exp(2.0 * float(hyp['lik'])); W = np.full(n, 1.0 / sn2); sW = np.sqrt(W)
    Kmvm = lambda v: Mx @ kronmvm(Ks, MxT @ v)

    def mvmB(v):
        return (sW[:, None] * Kmvm(sW[:, None] * v) + v) if v.ndim == 2 else sW * Kmvm(sW * v) + v

    def solveKiW(r):
        sw_ = sW[:, None] if np.ndim(r) == 2 else sW
        q, flag, relres, it = conjgrad(mvmB, sw_ * r, tol, maxit)
        if not flag:
            raise CGNotConverged('Not converged after %d iterations, r=%1.2e' % (it, np.max(relres)))
        solveKiW.iters, solveKiW.relres = it, float(np.max(relres))
        return sw_ * q

    dK_ab = lambda a, b: dirder(Ks, dKs, Mx, a, b)
    if slq is not None:
        def solveB(r):
            q, flag, relres, it = conjgrad(mvmB, r, tol, maxit)
            if not flag:
                raise CGNotConverged('Not converged after %d iterations, r=%1.2e' % (it, np.max(relres)))
            return q
        ldB2, dhyp_cov, dW = logdet_slq(W, Kmvm, dK_ab, slq['V'], slq['steps'], solveB, want_grad)
    elif cheby is None:
        V, ee, e = eigkron(Ks)
        ldB2, de, dW = logdet_fiedler(e, W)
        de = de * (e > 0)
        dhyp_cov = fiedler_dhyp(Ks, dKs, V, ee, de) if want_grad else None
    else:
        ldB2, dhyp_cov, dW = logdet_sample(W, Kmvm, dK_ab, cheby['V'], cheby['d'], cheby['emax'], want_grad)
    alpha = solveKiW(y - m)
    post = {'alpha': alpha, 'Kalpha': kronmvm(Ks, MxT @ alpha), 'cg_iter': solveKiW.iters,
            'cg_relres': solveKiW.relres, 'ldB2': ldB2,
            '_Ks': Ks, '_Mx': Mx, '_sW': sW, '_solveKiW': solveKiW}        # handles for pred_var_grid
    nlZ = (y - m) @ alpha / 2.0 + ldB2 + n * np.log(2 * np.pi * sn2) / 2.0
    if not want_grad:
        return post, nlZ, None
    dnlZ = {'cov': dhyp_cov - dK_ab(alpha, alpha) / 2.0,               # apx.m:271-272
            'mean': -dm(alpha),
            'lik': -sn2 * (alpha @ alpha) - 2.0 * dW.sum() / sn2 + n}
    return post, nlZ, dnlZ


def dlik_grid(hyp, mean, kind, xg, x, y, opt=None, Q=0, h=1e-5):
    """dnlZ/dX (n x D): dlikGrid.m:8-40 -> deriv_x.m:4-26 (no projection)."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    post, _, _ = inf_grid(hyp, mean, kind, xg, x, y, opt, Q, want_grad=False)
    alpha = post['alpha']
    xg = flat_grid(xg)
    _, dM = interpgrid(xg, x, want_deriv=True, deg=(opt or {}).get('deg', 3))           # dlikGrid.m:23, deriv_x.m:7
    beta = post['Kalpha']
    cols = []
    for i in range(x.shape[1]):
        wi = (xg[i].max() - xg[i].min()) if equi(xg[i]) else 1.0     # deriv_x.m:11-15
        xm = x.copy(); xm[:, i] -= h
        xp = x.copy(); xp[:, i] += h            # reference: -h then +2h (deriv_x.m:16-19)
        dmi = (mean_eval(mean, hyp.get('mean', []), xp)[0] - mean_eval(mean, hyp.get('mean', []), xm)[0]) / (2 * h)
        cols.append(-alpha * (dmi + dM[i] @ beta / wi))
    return np.stack(cols, axis=1)


def pred_var_grid(post, Rg, Rn):
    """Explained variance on the grid by perturb-and-MAP (infGrid.m:91-101 with sample(), infGrid.m:147-152).

    The reference draws ``randn(N,ns)`` and ``randn(n,ns)`` unseeded; here they are the explicit inputs Rg (N x ns)
    and Rn (n x ns).  a ~ N(0, Ku) through the clamped eigen-square-root, A = M a + Rn / sW, A <- post.L(A)
    (= -solveKiW, batched conjgrad), z = Ku M' A, vg = mean(z^2)."""
    Ks, Mx, sW, solve = post['_Ks'], post['_Mx'], post['_sW'], post['_solveKiW']
    Rg = np.asarray(Rg, dtype=np.float64); Rn = np.asarray(Rn, dtype=np.float64)
    ns = Rg.shape[1]
    V, ee, e = eigkron(Ks)
    A = kronmvm(V, np.sqrt(e)[:, None] * kronmvm([v.T for v in V], Rg))      # a ~ N(0, Ku)
    A = Mx @ A + Rn / sW[:, None]
    A = -solve(A)
    z = kronmvm(Ks, Mx.T @ A)
    return (z * z).sum(1) / ns


def predict_grid(hyp, mean, kind, xg, post, xs, Q=0, vg=None, deg=3):
    """[fmu fs2 ymu ys2] = post.predict(xs)  (infGrid.m:135-145; vg = 0 unless opt.pred_var; deg = opt.deg, infGrid.m:66,131)."""
    xs = np.asarray(xs, dtype=np.float64)
    if xs.ndim == 1:
        xs = xs[:, None]
    xg = flat_grid(xg)
    Ms = interpgrid(xg, xs, deg=deg)
    ms, _ = mean_eval(mean, hyp.get('mean', []), xs)
    fmu = ms + Ms @ post['Kalpha']
    hs = split_hyp(kind, hyp['cov'], len(xg), Q)
    ks = np.prod([factor_diag(kind, h_, Q) for h_ in hs]) * np.ones(xs.shape[0])
    ve = Ms @ vg if (vg is not None and np.abs(vg).sum() > 1e-10) else 0.0      # infGrid.m:141
    fs2 = np.maximum(ks - ve, 0.0)
    sn2 = np.exp(2.0 * float(hyp['lik']))
    return fmu, fs2, fmu.copy(), fs2 + sn2


# ---- learned input projection hyp.proj (infGrid.m:56-65,82-87,155-186; dlikGrid.m:13-22,39) ----------------------
def eff_proj(P, mode='none'):
    """hP of infGrid.m:58-65: 'orth' -> sqrtm(P P' + eps I) \\ P, 'norm' -> diag(1/sqrt(diag(P P') + eps)) P."""
    P = np.atleast_2d(np.asarray(P, dtype=np.float64))
    eps = np.finfo(np.float64).eps
    if str(mode).lower().startswith('orth'):
        w, V = np.linalg.eigh(P @ P.T + eps * np.eye(P.shape[0]))
        return np.linalg.solve((V * np.sqrt(w)) @ V.T, P)
    if str(mode).lower().startswith('norm'):
        return P / np.sqrt(np.sum(P * P, axis=1) + eps)[:, None]
    return P


def chain_orth(P, dP):
    """infGrid.m:176-180: chain rule for Q = sqrtm(P P') \\ P."""
    F, V = np.linalg.eigh(P @ P.T)
    sf = np.sqrt(F); S = (V * sf) @ V.T
    H = np.linalg.solve(S.T, dP).T                      # dP' / S
    G = H.T @ np.linalg.solve(S.T, P).T                 # H' * (P' / S)
    o = np.ones((dP.shape[0], 1))
    return (H - P.T @ V @ ((V.T @ (G + G.T) @ V) / (sf[:, None] @ o.T + o @ sf[None, :])) @ V.T).T


def chain_norm(P, dP):
    """infGrid.m:183-185: chain rule for Q = diag(1 / sqrt(diag(P P'))) P."""
    p = 1.0 / np.sqrt(np.sum(P * P, axis=1))
    return p[:, None] * dP - (np.sum(dP * P, axis=1) * p ** 3)[:, None] * P


def inf_grid_proj(hyp, mean, kind, xg, x, y, opt=None, Q=0, mode='none', h=1e-4):
    """infGrid with hyp['proj'] (d x D): inference on x hP'; dnlZ['proj'] by deriv_proj (infGrid.m:155-167: the mean's
    input derivative is a FORWARD difference with h = 1e-4) followed by the orth / norm chain rule."""
    x = np.asarray(x, dtype=np.float64)
    P = np.atleast_2d(np.asarray(hyp['proj'], dtype=np.float64)); hP = eff_proj(P, mode)
    xP = x @ hP.T
    post, nlZ, d = inf_grid(hyp, mean, kind, xg, xP, y, opt, Q)
    alpha = post['alpha']; beta = post['Kalpha']
    fg = flat_grid(xg)
    _, dM = interpgrid(fg, xP, want_deriv=True)
    m0 = mean_eval(mean, hyp.get('mean', []), xP)[0]
    dP = np.zeros_like(hP)
    for i in range(hP.shape[0]):
        wi = (fg[i].max() - fg[i].min()) if equi(fg[i]) else 1.0
        xq = xP.copy(); xq[:, i] += h
        dmi = (mean_eval(mean, hyp.get('mean', []), xq)[0] - m0) / h
        betai = dmi + dM[i] @ beta / wi
        for j in range(hP.shape[1]):
            dP[i, j] = -alpha @ (x[:, j] * betai)
    if str(mode).lower().startswith('orth'):
        dP = chain_orth(P, dP)
    elif str(mode).lower().startswith('norm'):
        dP = chain_norm(P, dP)
    d = dict(d); d['proj'] = dP
    return post, nlZ, d, hP


def dlik_grid_proj(hyp, mean, kind, xg, x, y, opt=None, Q=0, mode='none'):
    """dlikGrid.m with hyp.proj: deriv_x on x hP', then dx = dxP * hP (deriv_x.m:24)."""
    hP = eff_proj(hyp['proj'], mode)
    return dlik_grid(hyp, mean, kind, xg, np.asarray(x, dtype=np.float64) @ hP.T, y, opt, Q) @ hP
