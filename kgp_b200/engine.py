"""B200GP -- the GP backend object keras-gp expects, over sm_100a CUDA kernels.

Drop-in for ``kgp.backend.gpml.GPML`` (kgp/backend/gpml.py:34-229): the same
constructor and the same eight methods with the same argument meaning, return
types and error behaviour, so ``kgp.layers.GP`` (layers.py:63-75,144),
``kgp.callbacks.UpdateGP`` (callbacks.py:56-67,88) and ``kgp.models.Model``
(models.py:147-152,184,216-229) drive it unchanged.  Where GPML formats MATLAB
command strings for an Octave process, this class calls the C ABI of
libkgp_b200.so through ``_cabi``; NumPy arrays are the host-side carrier exactly
as in the reference.  There is no CPU fallback.
"""
from __future__ import absolute_import, division, print_function

import warnings
from pprint import pprint

import numpy as np

from . import _cabi
from .minimize import minimize

_MEANS = {'meanZero': 0, 'meanLinear': 1, 'meanConst': 2}


# ---- hyp.proj re-parametrisations (GPML/inf/infGrid.m:56-65) and their chain rules (infGrid.m:176-185): d x D host algebra
def _eff_proj(P, mode):
    eps = np.finfo(np.float64).eps
    if mode.startswith('orth'):                                    # sqrtm(P P' + eps I) \ P
        w, V = np.linalg.eigh(P.dot(P.T) + eps * np.eye(P.shape[0]))
        return np.linalg.solve((V * np.sqrt(w)).dot(V.T), P)
    if mode.startswith('norm'):                                    # diag(1 / sqrt(diag(P P') + eps)) P
        return P / np.sqrt(np.sum(P * P, axis=1) + eps)[:, None]
    return P


def _chain_proj(P, dP, mode):
    if mode.startswith('orth'):
        F, V = np.linalg.eigh(P.dot(P.T))
        sf = np.sqrt(F); S = (V * sf).dot(V.T)
        H = np.linalg.solve(S, dP).T
        G = H.T.dot(np.linalg.solve(S, P).T)
        return (H - P.T.dot(V).dot(V.T.dot(G + G.T).dot(V) / (sf[:, None] + sf[None, :])).dot(V.T)).T
    if mode.startswith('norm'):
        p = 1.0 / np.sqrt(np.sum(P * P, axis=1))
        return p[:, None] * dP - (np.sum(dP * P, axis=1) * p ** 3)[:, None] * P
    return dP


def _mean_mask(mean, mean_args):
    """'meanZero' | 'meanConst' | 'meanLinear' | 'meanSum' + ['{@meanLinear, @meanConst}']"""
    if mean == 'meanSum':
        spec = ' '.join(str(a) for a in (mean_args or []))
        names = [t.strip('@{}, ') for t in spec.replace(',', ' ').split() if t.strip('@{}, ')]
        if names not in (['meanLinear', 'meanConst'], ['meanLinear'], ['meanConst'], ['meanZero']):
            raise ValueError('meanSum is supported as {@meanLinear, @meanConst}; got %r' % (mean_args,))
        mask = 0
        for nm in names:
            mask |= _MEANS[nm]
        return mask
    if mean not in _MEANS:
        raise ValueError('Unknown mean function: %s' % mean)
    return _MEANS[mean]


class B200GP(object):
    """GP backend with GPML's interface, computed on one B200.

    Arguments (kept for signature compatibility with GPML.__init__,
    kgp/backend/gpml.py:42): ``engine`` may be None/'b200' or an int device
    index; ``engine_kwargs`` may hold {'device': i}; ``gpml_path`` is ignored.
    """

    def __init__(self, engine=None, engine_kwargs=None, gpml_path=None):
        engine_kwargs = dict(engine_kwargs or {})
        device = engine_kwargs.get('device', engine if isinstance(engine, int) else 0)
        self.eng = _cabi.Context(device)
        self.config = {}
        self.using_grid = False
        self._shapes = {}
        self._n_hyp = 0

    # -- configure (gpml.py:76-150) -----------------------------------------
    def configure(self, input_dim, hyp, opt, inf, mean, cov, lik, dlik,
                  grid_kwargs=None, cov_args=None, mean_args=None, verbose=1):
        if lik != 'likGauss':
            raise ValueError('Exact inference only possible with Gaussian likelihood')   # infGaussLik.m:12
        if inf not in _cabi.INF:
            raise ValueError('Unknown inference method: %s' % inf)
        if cov not in _cabi.COV:
            raise ValueError('Unknown covariance function: %s' % cov)
        self.input_dim = int(input_dim)
        self.config = {'lik': '{@%s}' % lik, 'inf': inf, 'dlik': dlik, 'mean': mean, 'cov': cov}
        self._inf, self._cov = inf, cov
        self._mean_mask = _mean_mask(mean, mean_args)
        self._sm_Q = int(cov_args[0]) if (cov == 'covSM' and cov_args) else 0
        self._opt = dict(opt or {})
        self._proj_mode = str(self._opt.get('proj', 'none')).lower()
        if self._proj_mode[:4] not in ('none', 'orth', 'norm'):
            raise ValueError("opt.proj must be 'none', 'orth' or 'norm' (infGrid.m:56-65)")
        if hyp.get('proj') is not None and inf != 'infGrid':
            raise ValueError('hyp.proj belongs to infGrid (infGrid.m:56-65)')
        self.using_grid = (inf == 'infGrid')
        if self.using_grid and np.ndim(self._opt.get('deg', 3)) > 0:          # one degree per grid dimension (apxGrid.m:448)
            p_grid = np.atleast_2d(hyp['proj']).shape[0] if hyp.get('proj') is not None else self.input_dim
            if np.size(self._opt['deg']) not in (1, p_grid):
                raise ValueError('opt.deg needs one entry or one per grid dimension (%d); got %r' % (p_grid, self._opt['deg']))
        if self.using_grid:
            assert grid_kwargs is not None, (
                "GPML: No arguments provided for grid generation for infGrid.")
            self._grid_k = float(grid_kwargs['k'])
            self._grid_eq = int(float(grid_kwargs['eq']))
            if 'xg' in grid_kwargs:
                xg = grid_kwargs['xg']
                # a nested entry [axis] (GPML's cell {axis}) selects a Toeplitz factor; plain arrays are dense factors
                self.eng.grid_set([([np.asarray(a[0], dtype=np.float64).reshape(-1)] if isinstance(a, (list, tuple)) else
                                    np.asarray(a, dtype=np.float64).reshape(-1)) for a in xg])
            # same in-place tiling as gpml.py:134; GPML then reads it column-major
            hyp['cov'] = np.tile(hyp['cov'], (1, self.input_dim))
        elif cov == 'covSM' and isinstance(hyp['cov'], (list, tuple)):
            # multivariate covSM: [log w (Q); log m (D x Q); log sqrt v (D x Q)], each block column-major (covSM.m:127-129)
            hyp['cov'] = np.concatenate([np.asarray(a, dtype=np.float64).reshape(-1, order='F') for a in hyp['cov']])[:, None]
        else:
            hyp['cov'] = np.asarray(hyp['cov'])
        self.hyp = {'cov': np.asarray(hyp['cov'], dtype=np.float64),
                    'lik': float(np.asarray(hyp['lik']).reshape(-1)[0]),
                    'mean': np.asarray(hyp.get('mean', []), dtype=np.float64)}
        if hyp.get('proj') is not None:                           # learned projection, d x D (infGrid.m:58)
            self.hyp['proj'] = np.atleast_2d(np.asarray(hyp['proj'], dtype=np.float64))
        self._push_model()
        if verbose:
            print("GP configuration:")
            pprint(self.config)

    def _hyp_vec(self):
        """unwrap(hyp): fields in alphabetical order cov, lik, mean (minimize.m:173-186)."""
        v = [self.hyp['cov'].flatten(order='F'), [self.hyp['lik']], self.hyp['mean'].flatten(order='F')]
        if 'proj' in self.hyp:
            v.append(self.hyp['proj'].flatten(order='F'))
        return np.concatenate(v)

    def _set_hyp_vec(self, v):
        nc, nm = self.hyp['cov'].size, self.hyp['mean'].size
        self.hyp['cov'] = np.asarray(v[:nc]).reshape(self.hyp['cov'].shape, order='F')
        self.hyp['lik'] = float(v[nc])
        self.hyp['mean'] = np.asarray(v[nc + 1:nc + 1 + nm]).reshape(self.hyp['mean'].shape, order='F')
        if 'proj' in self.hyp:
            self.hyp['proj'] = np.asarray(v[nc + 1 + nm:], dtype=np.float64).reshape(self.hyp['proj'].shape, order='F')

    def _push_model(self):
        self.eng.set_model(self._inf, self._cov, self.hyp['cov'].flatten(order='F'), self.hyp['lik'],
                           self.hyp['mean'].flatten(order='F'), mean_mask=self._mean_mask,
                           sm_Q=self._sm_Q, opt=self._opt)
        self._n_hyp = self.hyp['cov'].size + 1 + self.hyp['mean'].size
        if 'proj' in self.hyp:                                    # loaded slots are re-projected by the engine
            self.eng.set_proj(_eff_proj(self.hyp['proj'], self._proj_mode))

    # -- data / grid (gpml.py:152-165) --------------------------------------
    def update_data(self, which_set, X, y=None, verbose=0):
        assert which_set in {'tr', 'tst', 'val', 'tmp'}
        X = np.asarray(X)
        if X.ndim == 1:
            X = X[:, None]
        self.eng.set_data(_cabi.SLOTS[which_set], X, y)
        self._shapes[which_set] = X.shape

    def update_grid(self, which_set, verbose=0):
        assert which_set in {'tr', 'tst', 'val', 'tmp'}
        self.eng.grid_create(_cabi.SLOTS[which_set], self._grid_eq, self._grid_k)

    # -- evaluate (gpml.py:167-178; gp.m:93-126) ----------------------------
    def _eval(self, which_set, want):
        """gp.m training mode: an inference failure becomes NaN + zero gradient + warning."""
        n_out = self._n_hyp + (self.hyp['proj'].size if 'proj' in self.hyp else 0)
        try:
            nlZ, g = self.eng.eval(_cabi.SLOTS[which_set], want, self._n_hyp)
            if 'proj' in self.hyp and want >= 1:                  # dnlZ.proj (infGrid.m:82-87): deriv_proj + chain rule
                P = self.hyp['proj']
                dP = self.eng.dnlZ_proj(_cabi.SLOTS[which_set], P.shape[0], P.shape[1])
                g = np.concatenate([g, _chain_proj(P, dP, self._proj_mode).flatten(order='F')])
            return nlZ, g
        except _cabi.KgpError as e:
            if e.code in (_cabi.NOT_PD, _cabi.CG_NOT_CONVERGED):
                warnings.warn('Inference method failed [%s] .. attempting to continue' % e)   # gp.m:121
                return float('nan'), np.zeros(n_out)
            raise

    def evaluate(self, which_set, X=None, y=None, verbose=0):
        assert which_set in {'tr', 'tst', 'val', 'tmp'}
        if X is not None and y is not None:
            self.update_data(which_set, X, y)
        nlZ, _ = self._eval(which_set, 1)
        return nlZ

    # -- predict (gpml.py:180-195) ------------------------------------------
    def predict(self, X, X_tr=None, y_tr=None, return_var=False, verbose=0):
        self.update_data('tst', X)
        if X_tr is not None and y_tr is not None:
            self.update_data('tr', X_tr, y_tr)
        ymu, ys2 = self.eng.predict(np.asarray(X), want_var=bool(return_var))
        preds = ymu[:, None]
        if return_var:
            preds = (preds, ys2[:, None])
        return preds

    # -- train (gpml.py:197-205; minimize.m) --------------------------------
    def train(self, n_iter, X_tr=None, y_tr=None, verbose=0):
        if X_tr is not None and y_tr is not None:
            self.update_data('tr', X_tr, y_tr)

        def f(v):
            self._set_hyp_vec(v)
            self._push_model()
            return self._eval('tr', 1)

        v, fX, nev = minimize(self._hyp_vec(), f, -int(n_iter))
        self._set_hyp_vec(v)
        self._push_model()
        self.last_train = {'fX': fX, 'n_evals': nev}
        out = {'cov': self.hyp['cov'].copy(), 'lik': self.hyp['lik'], 'mean': self.hyp['mean'].copy()}
        if 'proj' in self.hyp:
            out['proj'] = self.hyp['proj'].copy()
        return out

    # -- dlik/dx (gpml.py:207-215) ------------------------------------------
    def get_dlik_dx(self, which_set, verbose=0):
        assert which_set in {'tr', 'tst', 'val', 'tmp'}
        n, D = self._shapes[which_set]
        return self.eng.dlik_dx(_cabi.SLOTS[which_set], n, D)     # raises on failure like the reference

    # -- eval_predict (gpml.py:217-229) -------------------------------------
    def eval_predict(self, X, X_tr=None, y_tr=None, verbose=0):
        self.update_data('tst', X)
        if X_tr is not None and y_tr is not None:
            self.update_data('tr', X_tr, y_tr)
        nlZ, _ = self._eval('tr', 1)
        ymu, _ = self.eng.predict(np.asarray(X), want_var=False)       # only ymu is pulled (gpml.py:224-229)
        return nlZ, ymu[:, None]

    # -- extras the reference keeps in the Octave workspace -----------------
    def stats(self):
        return self.eng.stats()

    def alpha(self):
        return self.eng.alpha(self._shapes['tr'][0])

    def grid(self):
        return self.eng.grid_get()


GP_BACKEND = B200GP
