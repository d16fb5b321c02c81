"""End-to-end epochs of two BASELINE configs through the PyTorch restatement of the caller:
C1 README GP-RNN (SimpleRNN(32) -> exact GP covSEiso, (10,2) sequences, 512 train, batch 32) and a small
msgp_mlp (MLP -> KISS-GP covSEiso on a 2-D feature grid).  The objective the GP reports must improve."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip('torch')


def test_c1_gp_rnn_epochs(built_lib):
    from kgp_b200 import harness
    torch.manual_seed(0); rng = np.random.RandomState(42)
    X = rng.randn(512, 10, 2).astype(np.float32); y = np.sin(X.sum((1, 2)) / 4)[:, None] + 0.1 * rng.randn(512, 1)
    net = harness.make_net('rnn', (10, 2), hidden=32).cuda()
    head = harness.GPHead(32, {'lik': -2.0, 'mean': [], 'cov': np.array([[-0.7], [0.0]])}, inf='infExact', cov='covSEiso')
    Xt = torch.from_numpy(X).cuda()
    hist = harness.fit(net, head, Xt, y, torch.optim.RMSprop(net.parameters(), lr=1e-3), epochs=4, batch_size=32, gp_n_iter=5)
    assert all(np.isfinite(h['nlml']) for h in hist)
    assert hist[-1]['nlml'] < hist[0]['nlml']
    ymu, ys2 = harness.predict(net, head, Xt[:64], return_var=True)
    assert ymu.shape == (64, 1) and np.all(ys2 > 0)


def test_msgp_mlp_epochs(built_lib):
    from kgp_b200 import harness
    torch.manual_seed(1); rng = np.random.RandomState(44)
    X = rng.randn(3000, 8).astype(np.float32); y = (np.sin(X[:, 0] * X[:, 1]) + X[:, 2])[:, None] + 0.1 * rng.randn(3000, 1)
    net = harness.make_net('mlp', (8,), hidden=16, out_dim=2).cuda()
    head = harness.GPHead(2, {'lik': np.log(0.3), 'mean': [], 'cov': np.array([[0.5], [1.0]])}, inf='infGrid', cov='covSEiso',
                          opt={'cg_maxit': 2000, 'cg_tol': 1e-6}, grid_kwargs={'eq': 1, 'k': 30.}, update_grid=1)
    Xt = torch.from_numpy(X).cuda()
    hist = harness.fit(net, head, Xt, y, torch.optim.Adam(net.parameters(), lr=1e-3), epochs=3, batch_size=128, gp_n_iter=3)
    assert all(np.isfinite(h['nlml']) for h in hist)
    assert hist[-1]['nlml'] < hist[0]['nlml']
