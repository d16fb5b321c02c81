"""kgp_predict_mean_noalloc_dev: the allocation-free predictive mean (static device scratch inside the library) against the
existing entry point kgp_predict_host, through the C ABI, on two calls in a row.  The file sorts last on purpose:
a failure here must not hide another test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def rel(a, b):
    a = np.asarray(a, dtype=float).ravel(); b = np.asarray(b, dtype=float).ravel()
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def make_gp(cov, n, D, seed):
    from kgp_b200 import B200GP
    rng = np.random.RandomState(seed)
    X = rng.randn(n, D); y = np.sin(X.sum(1) / 2) + 0.3 * X[:, 0] + 0.1 * rng.randn(n)
    ell = [[0.3 + 0.1 * d] for d in range(D)] if cov == 'covSEard' else [[0.4]]
    hyp = {'lik': np.log(0.2), 'mean': [[0.05 * (d + 1)] for d in range(D)] + [[0.3]], 'cov': ell + [[0.1]]}
    gp = B200GP()
    gp.configure(D, hyp, {}, 'infExact', 'meanSum', cov, 'likGauss', 'dlikExact', mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    gp.update_data('tr', X, y[:, None])
    return gp, rng


@pytest.mark.parametrize('cov,n,D', [('covSEard', 1500, 5), ('covSEiso', 700, 12)])
def test_noalloc_predict_mean_matches_predict_twice(built_lib, cov, n, D):
    import torch
    from kgp_b200 import _cabi
    gp, rng = make_gp(cov, n, D, 11)
    # the new entry point never factorises: without a current posterior it must refuse
    x0 = torch.zeros(4, D, dtype=torch.float64, device='cuda:0'); o0 = torch.zeros(4, dtype=torch.float64, device='cuda:0')
    with pytest.raises(_cabi.KgpError):
        gp.eng.predict_mean_noalloc(x0.data_ptr(), 4, o0.data_ptr())
    nlZ = gp.evaluate('tr')
    assert np.isfinite(nlZ)
    # two calls in a row, the second smaller than the first (stale scratch of call 1 must not leak into call 2)
    for ns in (300, 77):
        Xs = rng.randn(ns, D) * 1.2
        ref = gp.predict(Xs)[:, 0]                                   # kgp_predict_host
        xs_d = torch.as_tensor(Xs, dtype=torch.float64, device='cuda:0').contiguous()
        out_d = torch.full((ns,), float('nan'), dtype=torch.float64, device='cuda:0')
        torch.cuda.synchronize()
        gp.eng.predict_mean_noalloc(xs_d.data_ptr(), ns, out_d.data_ptr())
        got = out_d.cpu().numpy()
        assert np.all(np.isfinite(got))
        assert rel(got, ref) < 1e-10, (ns, rel(got, ref))
        # and once more on the same inputs: bit-identical (fixed summation order, shared scratch rewritten)
        out2 = torch.full((ns,), float('nan'), dtype=torch.float64, device='cuda:0')
        gp.eng.predict_mean_noalloc(xs_d.data_ptr(), ns, out2.data_ptr())
        assert np.array_equal(out2.cpu().numpy(), got)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(built_lib):
    import torch
    from kgp_b200 import _cabi
    gp, rng = make_gp('covSEiso', 1100, 3, 12)                       # 3 chunks of 512 training points
    gp.evaluate('tr')
    x0 = torch.zeros(8, 3, dtype=torch.float64, device='cuda:0'); o0 = torch.zeros(8, dtype=torch.float64, device='cuda:0')
    too_many = 262144 // 3 + 1
    with pytest.raises(_cabi.KgpError) as ei:
        gp.eng.predict_mean_noalloc(x0.data_ptr(), too_many, o0.data_ptr())     # refused before any memory is touched
    assert 'static array' in str(ei.value)
    gp.eng.predict_mean_noalloc(x0.data_ptr(), 8, o0.data_ptr())     # the context is still usable
    assert np.all(np.isfinite(o0.cpu().numpy()))
