"""CPU: the C-ABI library builds, loads and exports every symbol include/kgp_b200.h declares;
without a GPU the product path fails loudly (no CPU fallback)."""
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'kgp_b200.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(kgp_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol(built_lib):
    from kgp_b200 import _cabi
    syms = header_symbols()
    assert len(syms) >= 18
    for s in syms:
        assert hasattr(built_lib, s), 'missing export %s' % s
        assert s in _cabi.SIGNATURES, 'no ctypes signature for %s' % s
    assert built_lib.kgp_version().decode().startswith('kgp_b200')


def test_no_cpu_fallback(built_lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from kgp_b200 import B200GP, _cabi
    with pytest.raises(_cabi.KgpError):
        B200GP()


def test_product_never_imports_oracle():
    for dp, _, fs in os.walk(os.path.join(ROOT, 'kgp_b200')):
        for f in fs:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                s = open(os.path.join(dp, f)).read()
                assert 'gpml_exact' not in s and 'gpml_grid' not in s and 'import oracle' not in s, f


def test_mean_mask_parsing():
    from kgp_b200.engine import _mean_mask
    assert _mean_mask('meanZero', None) == 0
    assert _mean_mask('meanConst', None) == 2
    assert _mean_mask('meanSum', ['{@meanLinear, @meanConst}']) == 3
    with pytest.raises(ValueError):
        _mean_mask('meanPoly', None)


def test_product_minimize_reproduces_kat2():
    """KAT-2 of the GPML manual (GPML/doc/index.html:1019-1021: nlml2 = 14.13, exp(hyp2.lik) = 0.15 after
    minimize(hyp2, @gp, -100, ...)) run directly on the engine's host-side optimiser -- the one copy of minimize."""
    import gpml_exact as orc
    from kgp_b200.minimize import minimize as pmin
    from test_oracle_kat import demo_data
    x, y, _ = demo_data()

    def f(v):
        _, nl, d = orc.inf_exact({'cov': v[:2], 'lik': v[2], 'mean': []}, ['meanZero'], 'iso', x, y)
        return nl, np.concatenate([d['cov'], [d['lik']]])
    v, fX, nev = pmin(np.array([0.0, 0.0, np.log(0.1)]), f, -100)
    assert '%.2f' % fX[-1] == '14.13' and '%.2f' % np.exp(v[2]) == '0.15'
    assert nev <= 100


def test_grid_axis_equispaced_rule():
    """kgp_grid_set_host refuses axes that fail GPML's equi() test (apxGrid.m:684-695) instead of treating them as
    equispaced; the rule itself is host logic and callable without a GPU."""
    import ctypes as C
    from kgp_b200 import _cabi
    l = _cabi.lib()
    def eq(a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        return l.kgp_grid_axis_equispaced(a.ctypes.data_as(C.c_void_p), a.size)
    assert eq(np.linspace(-1, 2, 50)) == 1
    assert eq(np.linspace(-1, 2, 50) ** 3) == 0                       # non-uniform (-> neqinterp, or refused if not increasing)
    assert eq(np.linspace(2, -1, 50)) == 0                            # decreasing
    a = np.linspace(0, 1, 20); a[7] += 1e-8
    assert eq(a) == 0                                                 # beyond the 1e-9 tolerance
    a = np.linspace(0, 1, 20); a[7] += 1e-11
    assert eq(a) == 1


def test_unsupported_opt_keys_are_refused():
    from kgp_b200 import _cabi
    class Dummy(_cabi.Context):
        def __init__(self):
            self._l = _cabi.lib(); self._h = None
        def _chk(self, rc):
            pass
    d = Dummy()
    d._l = type('L', (), {'kgp_opt_default': staticmethod(_cabi.lib().kgp_opt_default), 'kgp_set_model': staticmethod(lambda *a: 0)})()
    d.set_model('infGrid', 'covSEiso', [0., 0.], 0.0, opt={'cg_tol': 1e-4, 'deg': 3})
    d.set_model('infGrid', 'covSEiso', [0., 0.], 0.0, opt={'proj': 'orth'})     # handled above the C ABI (engine.py)
    d.set_model('infGrid', 'covSEiso', [0., 0.], 0.0, opt={'deg': 1})           # linear / nearest-node interpolation are built
    d.set_model('infGrid', 'covSEiso', [0., 0.], 0.0, opt={'deg': [3, 1]})      # one degree per grid dimension (apxGrid.m:448)
    for bad in ({'deg': 5}, {'deg': [3, 2]}, {'deg': 1.5}, {'ndcovs': 5}):
        with pytest.raises(ValueError):
            d.set_model('infGrid', 'covSEiso', [0., 0.], 0.0, opt=bad)


def test_creategrid_fractional_k():
    """apxGrid.m:404-409: k > 1 stays fractional in du = range/k, only linspace floors the point count."""
    import gpml_grid as gg
    x = np.linspace(0.0, 10.0, 101)[:, None]
    ax = gg.creategrid(x, 1, 12.5)[0]
    du = 10.0 / 12.5
    assert ax.size == 12 and abs(ax[0] - (0.0 - 5 * du)) < 1e-15 and abs(ax[-1] - (10.0 + 5 * du)) < 1e-15
    assert np.allclose(np.diff(ax), (10.0 + 10 * du) / 11)


def test_struct_layouts_match_the_library():
    """kgp_abi_sizeof: the ctypes mirrors of kgp_opt / kgp_stats have the size the library was compiled with (lib() refuses to load
    otherwise), and kgp_opt_default fills the documented defaults."""
    import ctypes as C
    from kgp_b200 import _cabi
    l = _cabi.lib()
    assert l.kgp_abi_sizeof(0) == C.sizeof(_cabi.KgpOpt) and l.kgp_abi_sizeof(1) == C.sizeof(_cabi.KgpStats) and l.kgp_abi_sizeof(7) == -1
    o = _cabi.KgpOpt(); l.kgp_opt_default(C.byref(o))
    assert o.deg == 3 and list(o.deg_axis) == [-1, -1, -1, -1] and o.cg_deterministic == 1 and o.cg_precond == 0 and o.pred_var == 0
