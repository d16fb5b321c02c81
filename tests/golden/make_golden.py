"""Generates tests/golden/*.npz with the CPU oracle (oracle/gpml_exact.py).

The reference itself cannot run here (GPML needs Octave/MATLAB; probed, absent), so
these vectors come from the NumPy restatement, which is pinned on GPML's documented
known-answer values by tests/test_oracle_kat.py.  Re-run: python tests/golden/make_golden.py
"""
import os, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, 'oracle')); sys.path.insert(0, ROOT)
import numpy as np
import gpml_exact as orc
from bench import synth_c2


def exact_case(name, X, y, hyp, mode, mean, Xs):
    post, nlZ, d = orc.inf_exact(hyp, mean, mode, X, y)
    dX = orc.dlik_exact(hyp, mean, mode, X, y)
    ymu, ys2, _, _ = orc.gp_predict(hyp, mean, mode, X, y, Xs, post)
    g = np.concatenate([d['cov'], [d['lik']], d['mean']])
    idx = np.linspace(0, X.shape[0] - 1, 64).astype(int)
    np.savez_compressed(os.path.join(HERE, name + '.npz'), nlZ=nlZ, dnlZ=g, dX_rows=idx, dX=dX[idx],
                        alpha_rows=idx, alpha=post['alpha'][idx], Xs=Xs, ymu=ymu, ys2=ys2,
                        hyp_cov=hyp['cov'], hyp_lik=hyp['lik'], hyp_mean=np.asarray(hyp['mean'], dtype=float))
    print(name, nlZ, g)


if __name__ == '__main__':
    # C1: README GP-RNN shape (n=512, D=32, covSEiso), SURVEY 8(d)
    rng = np.random.RandomState(42)
    X = np.tanh(rng.randn(512, 32)); y = np.sin(X.sum(1) / 4) + 0.1 * rng.randn(512)
    exact_case('exact_c1', X, y, {'cov': np.array([-0.7, 0.0]), 'lik': -2.0, 'mean': []}, 'iso', ['meanZero'],
               np.tanh(rng.randn(100, 32)))
    # C2 at n=8192 (full-size C2 is checked through properties)
    X, y, hyp = synth_c2(8192)
    exact_case('exact_c2_8192', X, y, hyp, 'ard', ['meanZero'], X[:200] + 0.05)
    # meanSum{meanLinear, meanConst}, ragged n (not a multiple of 128), D=3
    rng = np.random.RandomState(7)
    X = rng.randn(777, 3); y = 0.5 * X[:, 0] - 0.2 * X[:, 2] + 1.0 + np.sin(2 * X[:, 1]) + 0.2 * rng.randn(777)
    exact_case('exact_mean_777', X, y, {'cov': np.array([0.2, -0.1, 0.4, 0.3]), 'lik': np.log(0.2),
                                       'mean': np.array([0.4, 0.0, -0.1, 0.8])}, 'ard',
               ['meanLinear', 'meanConst'], rng.randn(50, 3))
