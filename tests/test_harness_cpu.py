"""The PyTorch restatement of the caller (kgp/callbacks.py:44-70, kgp/losses.py:11-17) against a recording
stand-in for the backend: same calls in the same order, and the surrogate loss back-propagates dlik_dh."""
import numpy as np
import pytest

torch = pytest.importorskip('torch')


class FakeBackend(object):
    def __init__(self):
        self.calls = []; self.n = 0; self.D = 0

    def configure(self, *a, **k): self.calls.append('configure')
    def update_data(self, which, X, y=None, verbose=0):
        self.calls.append('update_data:' + which); self.n, self.D = X.shape
    def update_grid(self, which, verbose=0): self.calls.append('update_grid:' + which)
    def train(self, n_iter, verbose=0): self.calls.append('train:%d' % n_iter); return {'cov': np.zeros(2), 'lik': 0.0, 'mean': np.zeros(0)}
    def get_dlik_dx(self, which, verbose=0):
        self.calls.append('get_dlik_dx:' + which)
        return np.arange(self.n * self.D, dtype=np.float64).reshape(self.n, self.D) / 10.0
    def eval_predict(self, X, verbose=0): self.calls.append('eval_predict'); return 1.5, np.zeros((X.shape[0], 1))
    def predict(self, X, X_tr=None, y_tr=None, return_var=False, verbose=0): self.calls.append('predict'); return np.zeros((X.shape[0], 1))


def test_update_gp_call_sequence_matches_updategp():
    from kgp_b200 import harness
    fb = FakeBackend()
    head = harness.GPHead(3, {'lik': 0.0, 'mean': [], 'cov': [[0.0], [0.0]]}, inf='infGrid', grid_kwargs={'eq': 1, 'k': 10.}, update_grid=2, backend=fb)
    net = torch.nn.Linear(4, 3)
    X = torch.randn(50, 4); y = np.random.RandomState(0).randn(50, 1)
    harness.update_gp(net, head, X, y, epoch=0, batch_size=16, gp_n_iter=5)
    assert fb.calls == ['configure', 'update_data:tr', 'update_grid:tr', 'train:5', 'get_dlik_dx:tr', 'eval_predict']
    fb.calls.clear()
    harness.update_gp(net, head, X, y, epoch=1, batch_size=16, gp_n_iter=5)        # update_grid=2: odd epochs keep the grid
    assert fb.calls == ['update_data:tr', 'train:5', 'get_dlik_dx:tr', 'eval_predict']
    assert head.nlml == 1.5 and head.dlik_dh.shape == (50, 3)


def test_surrogate_loss_backpropagates_dlik_dh():
    from kgp_b200 import harness
    H = torch.randn(8, 3, dtype=torch.float64, requires_grad=True)
    dl = torch.randn(8, 3, dtype=torch.float64)
    harness.gp_surrogate_loss(H, dl).backward()
    assert torch.allclose(H.grad, dl / 8)                      # Keras: mean over the batch of sum(H * dlik_dh, axis=1)


def test_fit_runs_epoch_loop_and_steps_the_network():
    from kgp_b200 import harness
    fb = FakeBackend()
    head = harness.GPHead(3, {'lik': 0.0, 'mean': [], 'cov': [[0.0], [0.0]]}, backend=fb)
    net = harness.make_net('rnn', (10, 2), hidden=3)
    X = torch.randn(40, 10, 2); y = np.zeros((40, 1))
    w0 = [p.detach().clone() for p in net.parameters()]
    hist = harness.fit(net, head, X, y, torch.optim.SGD(net.parameters(), lr=0.1), epochs=2, batch_size=16)
    assert len(hist) == 2 and fb.calls.count('train:1') == 2 and 'update_grid:tr' not in fb.calls
    assert any(not torch.equal(a, b) for a, b in zip(w0, net.parameters()))
