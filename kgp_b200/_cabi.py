"""ctypes binding of libkgp_b200.so (C ABI in include/kgp_b200.h).

Replaces the reference's engine layer (kgp/backend/engines.py:7-99: push / pull /
eval through oct2py).  There is no CPU fallback: if the library or a CUDA device is
missing, importing works but creating a context raises.
"""
import ctypes as C
import os
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('KGP_B200_LIB') or os.path.join(_HERE, 'libkgp_b200.so')   # (override: A/B builds of the library)

OK, BAD_ARG, CUDA_ERR, NOT_PD, CG_NOT_CONVERGED, NO_GRID, NO_DATA = range(7)
SLOTS = {'tr': 0, 'tst': 1, 'val': 2, 'tmp': 3}
INF = {'infExact': 0, 'infGrid': 1}
COV = {'covSEiso': 0, 'covSEard': 1, 'covSM': 2}


class KgpOpt(C.Structure):
    _fields_ = [('cg_tol', C.c_double), ('cg_maxit', C.c_int), ('ldB2_cheby', C.c_int),
                ('ldB2_cheby_hutch', C.c_int), ('ldB2_cheby_degree', C.c_int),
                ('ldB2_cheby_maxit', C.c_int), ('ldB2_cheby_seed', C.c_uint64),
                ('ard_reference_form', C.c_int), ('cg_precond', C.c_int), ('pred_var', C.c_int),
                ('ldB2_slq', C.c_int), ('ldB2_slq_steps', C.c_int), ('cg_deterministic', C.c_int), ('deg', C.c_int), ('deg_axis', C.c_int * 4)]


class KgpStats(C.Structure):
    _fields_ = [('ms_total', C.c_double), ('ms_kbuild', C.c_double), ('ms_potrf', C.c_double),
                ('ms_inverse', C.c_double), ('ms_reduce', C.c_double), ('ms_solve', C.c_double),
                ('ms_gemm', C.c_double), ('gemm_flops', C.c_double), ('gemm_launches', C.c_long),
                ('kernel_launches', C.c_long), ('cg_iters', C.c_int), ('cg_relres', C.c_double),
                ('potrf_info', C.c_int), ('n', C.c_long), ('np', C.c_long),
                ('gemm_class_ms', C.c_double * 8), ('gemm_class_flops', C.c_double * 8)]

    def as_dict(self):
        return {k: (list(getattr(self, k)) if k.startswith('gemm_class') else getattr(self, k)) for k, _ in self._fields_}


class KgpError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__('kgp_b200 error %d: %s' % (code, msg))
        self.code = code


_lib = None
_DP = C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol include/kgp_b200.h declares
SIGNATURES = {
    'kgp_ctx_create': (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    'kgp_ctx_destroy': (C.c_int, [C.c_void_p]),
    'kgp_last_error': (C.c_char_p, [C.c_void_p]),
    'kgp_version': (C.c_char_p, []),
    'kgp_opt_default': (None, [C.POINTER(KgpOpt)]),
    'kgp_abi_sizeof': (C.c_long, [C.c_int]),
    'kgp_set_profiling': (C.c_int, [C.c_void_p, C.c_int]),
    'kgp_set_data_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_void_p]),
    'kgp_set_data_dev': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_long, C.c_int, C.c_void_p]),
    'kgp_set_model': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int,
                                C.c_double, C.c_void_p, C.c_int, C.POINTER(KgpOpt)]),
    'kgp_eval': (C.c_int, [C.c_void_p, C.c_int, C.c_int, _DP, C.c_void_p]),
    'kgp_dlik_dx_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    'kgp_dlik_dx_dev': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    'kgp_predict_host': (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    'kgp_grid_create': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_double]),
    'kgp_grid_set_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]),
    'kgp_grid_set_toeplitz': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    'kgp_grid_axis_equispaced': (C.c_int, [C.c_void_p, C.c_int]),
    'kgp_kmeans1d_host': (C.c_int, [C.c_void_p, C.c_long, C.c_long, C.c_void_p]),
    'kgp_grid_get_host': (C.c_int, [C.c_void_p, C.POINTER(C.c_int), C.c_void_p, C.c_void_p]),
    'kgp_set_proj_host': (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_int]),
    'kgp_predict_mean_noalloc_dev': (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_void_p]),
    'kgp_dnlZ_proj_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p]),
    'kgp_get_alpha_host': (C.c_int, [C.c_void_p, C.c_void_p]),
    'kgp_last_stats': (C.c_int, [C.c_void_p, C.POINTER(KgpStats)]),
    'kgp_bench_dmma_peak': (C.c_int, [C.c_void_p, C.c_int, _DP]),
    'kgp_bench_dgemm': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, _DP, _DP]),
    'kgp_bench_syrk': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_int, _DP, _DP]),
    'kgp_set_probes_host': (C.c_int, [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_double]),
    'kgp_grid_solve_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]),
    'kgp_grid_pred_var_host': (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_void_p]),
    'kgp_nccl_unique_id': (C.c_int, [C.c_char_p, C.c_char_p]),
    'kgp_dist_last_error': (C.c_char_p, []),
    'kgp_comm_init': (C.c_int, [C.c_void_p, C.c_int, C.c_int, C.c_char_p, C.c_char_p]),
    'kgp_dist_col_owner': (C.c_int, [C.c_int, C.c_int]),
    'kgp_dist_zrow_begin': (C.c_long, [C.c_int, C.c_int, C.c_long]),
    'kgp_dist_row_begin': (C.c_long, [C.c_int, C.c_int, C.c_long]),
    'kgp_dist_rows_sharded': (C.c_int, [C.c_int, C.c_long, C.c_long]),
}


def lib():
    """Load the shared library (built by __graft_entry__.build / csrc/Makefile)."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise ImportError('libkgp_b200.so not built: run `python -c "import __graft_entry__ as g; g.build()"`')
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        # the ctypes mirrors must have the layout the library was compiled with (a mismatch would corrupt memory silently)
        if l.kgp_abi_sizeof(0) != C.sizeof(KgpOpt) or l.kgp_abi_sizeof(1) != C.sizeof(KgpStats):
            raise ImportError('libkgp_b200.so and kgp_b200/_cabi.py disagree on kgp_opt / kgp_stats (%d / %d vs %d / %d bytes): rebuild'
                              % (l.kgp_abi_sizeof(0), l.kgp_abi_sizeof(1), C.sizeof(KgpOpt), C.sizeof(KgpStats)))
        _lib = l
    return _lib


def nccl_lib_path():
    """libnccl.so.2 as shipped with PyTorch (nvidia-nccl wheel); None -> the loader's default search."""
    try:
        import nvidia.nccl as _n
        for base in list(getattr(_n, '__path__', [])):
            cand = os.path.join(base, 'lib', 'libnccl.so.2')
            if os.path.isfile(cand):
                return cand
    except Exception:
        pass
    return None


def nccl_unique_id():
    """128-byte NCCL id (rank 0 calls this; the host layer ships it to the other ranks)."""
    buf = C.create_string_buffer(128)
    p = nccl_lib_path()
    rc = lib().kgp_nccl_unique_id(buf, p.encode() if p else None)
    if rc != OK:
        raise KgpError(rc, lib().kgp_dist_last_error().decode())
    return buf.raw


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def f64(a, shape=None):
    a = np.ascontiguousarray(a, dtype=np.float64)      # float32 -> float64 like engines.py:84-88
    if shape is not None:
        a = a.reshape(shape)
    return a


class Context(object):
    """One engine context = one CUDA device, stream and workspace set."""

    def __init__(self, device=0):
        self._l = lib()
        h = C.c_void_p()
        rc = self._l.kgp_ctx_create(int(device), C.byref(h))
        if rc != OK:
            raise KgpError(rc, 'cannot create a CUDA context on device %d (no GPU / driver?) -- '
                               'this engine has no CPU path' % device)
        self._h = h

    def close(self):
        if getattr(self, '_h', None):
            self._l.kgp_ctx_destroy(self._h)
            self._h = None

    __del__ = close

    def _chk(self, rc):
        if rc != OK:
            raise KgpError(rc, self._l.kgp_last_error(self._h).decode())

    def set_data(self, slot, X, y=None):
        X = f64(X)
        if X.ndim == 1:
            X = X[:, None]
        n, D = X.shape
        yy = None if y is None else f64(y, (-1,))
        if yy is not None and yy.size != n:
            raise ValueError('X and y disagree: %d vs %d rows' % (n, yy.size))
        self._chk(self._l.kgp_set_data_host(self._h, slot, _ptr(X), n, D, _ptr(yy)))

    def set_data_dev(self, slot, X_ptr, n, D, y_ptr=None):
        self._chk(self._l.kgp_set_data_dev(self._h, slot, C.c_void_p(X_ptr), n, D,
                                           C.c_void_p(y_ptr) if y_ptr else None))

    def set_model(self, inf, cov, hyp_cov, hyp_lik, hyp_mean=(), mean_mask=0, sm_Q=0, opt=None):
        hc = f64(hyp_cov, (-1,)); hm = f64(hyp_mean, (-1,))
        o = KgpOpt(); self._l.kgp_opt_default(C.byref(o))
        for k, v in (opt or {}).items():
            if k == 'pred_var':
                v = int(np.ceil(abs(float(v))))                  # infGrid.m:94
            if k == 'deg':
                # one degree for all axes or one per grid dimension (apxGrid.m:448); d > 3 selects GP interpolation
                # (apxGrid.m:501-504), which is not built
                dv = [int(t) for t in np.atleast_1d(v).ravel()]
                if not dv or len(dv) > 4 or any(t not in (0, 1, 3) for t in dv) or any(float(t) != int(t) for t in np.atleast_1d(v).ravel()):
                    raise ValueError('opt.deg=%r is not supported by the B200 engine (degrees 0, 1 or 3)' % (v,))
                if len(dv) > 1:
                    for i, t in enumerate(dv):
                        o.deg_axis[i] = t
                    continue
                v = dv[0]
            if hasattr(o, k):
                setattr(o, k, type(getattr(o, k))(v))
            elif k == 'stat' and not v:
                pass
            elif k == 'proj':
                pass                                             # handled above the C ABI (engine.py: hyp.proj, orth / norm)
            elif k in ('ndcovs', 'stat'):
                # GPML options that change the numbers (infGrid.m:56,70): refusing beats returning results as if the option
                # had been honoured
                raise ValueError('opt.%s=%r is not supported by the B200 engine' % (k, v))
            else:
                import warnings
                warnings.warn('opt.%s is not a known GPML option of this path; ignored' % k)
        self._chk(self._l.kgp_set_model(self._h, INF[inf], COV[cov], int(sm_Q), int(mean_mask), _ptr(hc), hc.size,
                                        float(hyp_lik), _ptr(hm), hm.size, C.byref(o)))

    def eval(self, slot, want, n_hyp):
        nlZ = C.c_double()
        g = np.zeros(max(n_hyp, 1))
        self._chk(self._l.kgp_eval(self._h, slot, want, C.byref(nlZ), _ptr(g) if want else None))
        return nlZ.value, g[:n_hyp]

    def dlik_dx(self, slot, n, D):
        out = np.empty((n, D))
        self._chk(self._l.kgp_dlik_dx_host(self._h, slot, _ptr(out)))
        return out

    def predict(self, Xs, want_var=True):
        Xs = f64(Xs)
        if Xs.ndim == 1:
            Xs = Xs[:, None]
        ns = Xs.shape[0]
        ymu = np.empty(ns); ys2 = np.empty(ns) if want_var else None
        self._chk(self._l.kgp_predict_host(self._h, _ptr(Xs), ns, _ptr(ymu), _ptr(ys2), None, None))
        return ymu, ys2

    def alpha(self, n):
        a = np.empty(n)
        self._chk(self._l.kgp_get_alpha_host(self._h, _ptr(a)))
        return a

    def grid_create(self, slot, eq, k):
        self._chk(self._l.kgp_grid_create(self._h, slot, int(eq), float(k)))

    def grid_set(self, axes):
        """axes: one array per grid dimension; a list / tuple holding one array (GPML's nested cell {axis}) marks the
        dimension as a Toeplitz factor (apxGrid.m:258-281)."""
        toep = np.array([1 if isinstance(a, (list, tuple)) else 0 for a in axes], dtype=np.int32)
        axes = [a[0] if isinstance(a, (list, tuple)) else a for a in axes]
        self._grid_set_plain(axes)
        self._chk(self._l.kgp_grid_set_toeplitz(self._h, len(axes), _ptr(toep)))

    def _grid_set_plain(self, axes):
        axes = [f64(a, (-1,)) for a in axes]
        lens = np.array([a.size for a in axes], dtype=np.int32)
        cat = np.concatenate(axes)
        self._chk(self._l.kgp_grid_set_host(self._h, len(axes), _ptr(lens), _ptr(cat)))

    def predict_mean_noalloc(self, xs_dev_ptr, ns, ymu_dev_ptr):
        """kgp_predict_mean_noalloc_dev: device pointers in and out, static scratch inside the library, no allocation."""
        self._chk(self._l.kgp_predict_mean_noalloc_dev(self._h, C.c_void_p(int(xs_dev_ptr)), int(ns), C.c_void_p(int(ymu_dev_ptr))))

    def set_proj(self, P):
        """hyp.proj: the effective projector hP (d x D), or None to clear (kgp_set_proj_host)."""
        if P is None:
            self._chk(self._l.kgp_set_proj_host(self._h, None, 0, 0)); return
        P = f64(np.atleast_2d(P))
        self._chk(self._l.kgp_set_proj_host(self._h, _ptr(P), P.shape[0], P.shape[1]))

    def dnlZ_proj(self, slot, d, D):
        out = np.empty((d, D))
        self._chk(self._l.kgp_dnlZ_proj_host(self._h, slot, _ptr(out)))
        return out

    def grid_get(self):
        na = C.c_int(); lens = np.zeros(16, dtype=np.int32)
        self._chk(self._l.kgp_grid_get_host(self._h, C.byref(na), _ptr(lens), None))
        cat = np.empty(int(lens[:na.value].sum()))
        self._chk(self._l.kgp_grid_get_host(self._h, C.byref(na), _ptr(lens), _ptr(cat)))
        out, o = [], 0
        for a in range(na.value):
            out.append(cat[o:o + lens[a]].copy()); o += lens[a]
        return out

    def comm_init(self, nranks, rank, unique_id):
        """Join the NCCL communicator of a multi-GPU job (one Context per rank)."""
        p = nccl_lib_path()
        self._chk(self._l.kgp_comm_init(self._h, int(nranks), int(rank), unique_id, p.encode() if p else None))

    def set_probes(self, V=None, emax=0.0):
        """Inject Hutchinson probes (V: n x m, +-1) and emax for opt.ldB2_cheby; V=None -> generated probes."""
        if V is None:
            self._chk(self._l.kgp_set_probes_host(self._h, None, 0, 0, 0.0)); return
        V = np.asarray(V, dtype=np.float64)
        Vt = np.ascontiguousarray(V.T)                    # one probe per row
        self._chk(self._l.kgp_set_probes_host(self._h, _ptr(Vt), Vt.shape[1], Vt.shape[0], float(emax)))

    def grid_solve(self, slot, R):
        """post.L on the columns of R (n x m): one batched CG with per-column convergence masks."""
        R = np.asarray(R, dtype=np.float64)
        if R.ndim == 1:
            R = R[:, None]
        Rt = np.ascontiguousarray(R.T); out = np.empty_like(Rt)
        self._chk(self._l.kgp_grid_solve_host(self._h, slot, _ptr(Rt), Rt.shape[0], _ptr(out)))
        return np.ascontiguousarray(out.T)

    def grid_pred_var(self, slot, N, Rg=None, Rn=None, ns=20, seed=0):
        """Explained variance on the grid (infGrid.m:91-101).  Rg (N x ns), Rn (n x ns): the reference's randn draws,
        or None for device-generated ones.  Returns vg (N,), first axis fastest."""
        vg = np.empty(int(N))
        if Rg is not None:
            Rgt = np.ascontiguousarray(np.asarray(Rg, dtype=np.float64).T)
            Rnt = np.ascontiguousarray(np.asarray(Rn, dtype=np.float64).T)
            assert Rgt.shape[1] == int(N) and Rgt.shape[0] == Rnt.shape[0]
            self._chk(self._l.kgp_grid_pred_var_host(self._h, slot, _ptr(Rgt), _ptr(Rnt), Rgt.shape[0], 0, _ptr(vg)))
        else:
            self._chk(self._l.kgp_grid_pred_var_host(self._h, slot, None, None, int(ns), int(seed), _ptr(vg)))
        return vg

    def set_profiling(self, on):
        self._chk(self._l.kgp_set_profiling(self._h, int(bool(on))))

    def stats(self):
        s = KgpStats()
        self._chk(self._l.kgp_last_stats(self._h, C.byref(s)))
        return s.as_dict()

    def bench_dmma_peak(self, iters=20000):
        t = C.c_double()
        self._chk(self._l.kgp_bench_dmma_peak(self._h, iters, C.byref(t)))
        return t.value

    def bench_dgemm(self, n=8192, k=8192, reps=3, syrk=False):
        t = C.c_double(); ms = C.c_double()
        self._chk((self._l.kgp_bench_syrk if syrk else self._l.kgp_bench_dgemm)(self._h, n, k, reps, C.byref(t), C.byref(ms)))
        return t.value, ms.value
