"""The caller of the hot path, restated in PyTorch: the per-epoch GP update of
``kgp.callbacks.UpdateGP`` (callbacks.py:44-89), the surrogate objective ``gen_gp_loss``
(losses.py:11-17: ``sum(H * dlik_dh[batch_ids], axis=1)``) and ``Model.fit / predict / evaluate``'s use of
the backend (models.py:63-234), so that the five BASELINE configs run end to end without Keras.

Nothing here is on the measured path: the networks are ordinary torch modules (SimpleRNN(32) / MLP /
LSTM as in the reference's examples), the GP is a ``B200GP`` backend object driven through exactly the
eight methods the reference calls.  The hand-off the reference pays through oct2py (features to Octave,
gradients back) is a host<->device copy here.
"""
from __future__ import absolute_import, division, print_function

import time

import numpy as np

from .engine import B200GP


def gp_surrogate_loss(H, dlik_dh_batch):
    """gen_gp_loss (kgp/losses.py:11-17): the value is meaningless, its gradient w.r.t. H is dnlZ/dH, which is
    all back-propagation needs (tutorials/README.md:56-81).  Keras averages the per-sample losses over the
    batch (mean of sum(.., axis=1)); this keeps that scale."""
    return (H * dlik_dh_batch).sum(dim=1).mean()


class GPHead(object):
    """Output GP layer state (kgp/layers.py:16-151): the backend object plus what UpdateGP stores on the layer."""

    def __init__(self, input_dim, hyp, inf='infExact', mean='meanZero', cov='covSEiso', opt=None, grid_kwargs=None,
                 cov_args=None, mean_args=None, update_grid=1, device=0, backend=None):
        self.backend = backend if backend is not None else B200GP(engine_kwargs={'device': device})
        dlik = 'dlikGrid' if inf == 'infGrid' else 'dlikExact'                       # layers.py:66-69
        self.backend.configure(input_dim, hyp, opt or {}, inf, mean, cov, 'likGauss', dlik, grid_kwargs=grid_kwargs,
                               cov_args=cov_args, mean_args=mean_args, verbose=0)
        self.update_grid = update_grid if inf == 'infGrid' else 0
        self.hyp = None; self.dlik_dh = None; self.nlml = None; self.mse = None


def transform(net, X, batch_size):
    """Model.transform (models.py:118-131): forward pass of the network in batches, features on the host."""
    import torch
    outs = []
    with torch.no_grad():
        for i in range(0, X.shape[0], batch_size):
            outs.append(net(X[i:i + batch_size]).double().cpu())
    return torch.cat(outs).numpy()


def update_gp(net, head, X, y, epoch, batch_size=128, gp_n_iter=1):
    """UpdateGP.on_epoch_begin (callbacks.py:44-70), in the reference's order."""
    t0 = time.perf_counter()
    H = transform(net, X, batch_size)
    head.backend.update_data('tr', H, y)
    if head.update_grid and (epoch % head.update_grid == 0):
        head.backend.update_grid('tr')
    head.hyp = head.backend.train(gp_n_iter)
    head.dlik_dh = head.backend.get_dlik_dx('tr')
    nlml, preds = head.backend.eval_predict(H)
    head.nlml = nlml; head.mse = float(np.mean((np.asarray(y).reshape(-1, 1) - preds) ** 2))
    return time.perf_counter() - t0


def fit(net, head, X, y, optimizer, epochs=1, batch_size=128, gp_n_iter=1, shuffle=True, seed=0, log=None):
    """Model.fit with the UpdateGP callback (models.py:63-116): GP update at the beginning of every epoch, then
    mini-batch SGD on the surrogate loss with the batch's rows of dlik_dh."""
    import torch
    rng = np.random.RandomState(seed)
    n = X.shape[0]
    history = []
    for epoch in range(epochs):
        gp_s = update_gp(net, head, X, y, epoch, batch_size, gp_n_iter)
        dl = torch.from_numpy(head.dlik_dh).to(X.device, torch.float32)
        idx = rng.permutation(n) if shuffle else np.arange(n)
        net.train()
        for i in range(0, n, batch_size):
            b = torch.from_numpy(idx[i:i + batch_size]).to(X.device)
            optimizer.zero_grad()
            loss = gp_surrogate_loss(net(X[b]), dl[b])
            loss.backward()
            optimizer.step()
        rec = {'epoch': epoch, 'nlml': head.nlml, 'mse': head.mse, 'gp_update_elapsed': gp_s}
        history.append(rec)
        if log:
            log(rec)
    return history


def finetune(net, head, X, y, batch_size=128, gp_n_iter=10):
    """Model.finetune (models.py:133-152): re-fit the GP hyper-parameters on the final features."""
    H = transform(net, X, batch_size)
    head.backend.update_data('tr', H, y)
    if head.update_grid:
        head.backend.update_grid('tr')
    head.hyp = head.backend.train(gp_n_iter)
    return head.hyp


def predict(net, head, X, X_tr=None, y_tr=None, batch_size=128, return_var=False):
    """Model.predict (models.py:186-234)."""
    H = transform(net, X, batch_size)
    if X_tr is not None and y_tr is not None:
        H_tr = transform(net, X_tr, batch_size)
        return head.backend.predict(H, H_tr, y_tr, return_var=return_var)
    return head.backend.predict(H, return_var=return_var)


def make_net(kind, input_shape, hidden=32, out_dim=None):
    """The feature extractors of the reference's examples: 'rnn' SimpleRNN(32) (README.md:38-48), 'mlp'
    (examples/mlp_kin40k.py), 'lstm' (examples/msgp_lstm_actuator.py)."""
    import torch
    import torch.nn as nn

    class Seq(nn.Module):
        def __init__(self, cell, d_in, hid, d_out):
            super().__init__()
            self.rnn = cell(d_in, hid, batch_first=True)
            self.proj = nn.Linear(hid, d_out) if d_out else None

        def forward(self, x):
            o, _ = self.rnn(x)
            h = o[:, -1]
            return self.proj(h) if self.proj is not None else h

    if kind == 'rnn':
        return Seq(torch.nn.RNN, input_shape[-1], hidden, out_dim)
    if kind == 'lstm':
        return Seq(torch.nn.LSTM, input_shape[-1], hidden, out_dim)
    if kind == 'mlp':
        d = input_shape[-1]
        return nn.Sequential(nn.Linear(d, 4 * hidden), nn.ReLU(), nn.Linear(4 * hidden, 2 * hidden), nn.ReLU(),
                             nn.Linear(2 * hidden, out_dim or hidden), nn.ReLU() if out_dim is None else nn.Identity())
    raise ValueError(kind)
