"""GPU tests of the sharded exact path (SURVEY App. D "Multi-GPU invariance": 1e-10 relative to the 1-GPU result).

* the tile-symmetric gradient pass + ownership-filtered Z Z^T launch also run on ONE GPU (KGP_TILE_REDUCE=1):
  parity against the oracle at 1e-8 like the default path;
* with >= 2 GPUs visible: torchrun world_size 2 over NCCL, sharded vs unsharded results."""
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SINGLE = r'''
import sys, os, numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, 'oracle'))
from kgp_b200 import B200GP
import gpml_exact as orc
rng = np.random.RandomState(7); n, D = 1100, 5
X = rng.randn(n, D); y = np.sin(X.sum(1)) + 0.1 * rng.randn(n)
ell = np.log(np.array([1.0, 2.0, 0.7, 1.5, 3.0]))
hyp = {'lik': -1.5, 'mean': [], 'cov': np.concatenate([ell, [0.3]])[:, None]}
gp = B200GP(); gp.configure(D, dict(hyp), {}, 'infExact', 'meanZero', 'covSEard', 'likGauss', 'dlikExact', verbose=0)
gp.update_data('tr', X, y[:, None])
nlZ, g = gp.eng.eval(0, 2, gp._n_hyp); dX = gp.get_dlik_dx('tr')
oh = {'cov': np.concatenate([ell, [0.3]]), 'lik': -1.5, 'mean': []}
post, nlZ0, d0 = orc.inf_exact(oh, ['meanZero'], 'ard', X, y)
dX0 = orc.dlik_exact(oh, ['meanZero'], 'ard', X, y, ard_reference_form=False)
g0 = np.concatenate([d0['cov'], [d0['lik']]])
rel = lambda a, b: float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(b)))
print('REL', rel(nlZ, nlZ0), rel(g, g0), rel(dX, dX0))
'''


@pytest.mark.parametrize('tile_reduce', ['1', '0'])
def test_tile_reduce_path_single_gpu(built_lib, tile_reduce):
    """both gradient passes: the symmetric tile pass (default) and the full-square pass (KGP_TILE_REDUCE=0)"""
    env = dict(os.environ, KGP_TILE_REDUCE=tile_reduce)
    out = subprocess.check_output([sys.executable, '-c', _SINGLE % {'root': ROOT}], env=env, timeout=600).decode()
    vals = [float(v) for v in [l for l in out.splitlines() if l.startswith('REL')][0].split()[1:]]
    assert max(vals) < 1e-8, out


@pytest.mark.parametrize('zpush', ['default', '1'])
def test_sharded_equals_single_world2(built_lib, zpush):
    """zpush = '1': the all-gather of Z = L^-T as pipelined peer copies into the other ranks' W2 (the default for more than two
    ranks) instead of the default of two ranks (pipelined broadcasts under a staged Z Z^T)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29533' if zpush == 'default' else '29536', os.path.join(ROOT, 'scripts', 'dist_exact.py'), '1000', '2600']
    env = dict(os.environ) if zpush == 'default' else dict(os.environ, KGP_DIST_ZPUSH=zpush)
    out = subprocess.check_output(cmd, timeout=900, stderr=subprocess.STDOUT, env=env).decode()
    lines = [l for l in out.splitlines() if l.startswith('n=')]
    assert len(lines) == 2, out
    for l in lines:
        rels = [float(tok) for tok in l.split('rel')[1].replace('dnlZ', ' ').replace('dX', ' ').replace('ymu', ' ').replace('ys2', ' ').split()]
        assert max(rels) < 1e-10, l
    # a non-positive-definite input: both ranks return NOT_PD (status 3), none is left with NaNs and status OK
    npd = [l for l in out.splitlines() if l.startswith('NOTPD')]
    assert npd and 'min 3 max 3' in npd[0] and 'nan' in npd[0], out


def test_sharded_probes_equal_single_world2(built_lib):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29534', os.path.join(ROOT, 'scripts', 'dist_grid.py'), '6000']
    out = subprocess.check_output(cmd, timeout=900, stderr=subprocess.STDOUT).decode()
    line = [l for l in out.splitlines() if l.startswith('n=')][0]
    rels = [float(t) for t in line.split('rel')[1].replace('dnlZ', ' ').split()]
    assert max(rels) < 1e-9, line


def test_row_sharded_kissgp_cg_world2(built_lib):
    """SURVEY 8(e) KISS-GP (ii): points row-sharded over two ranks, the per-iteration reductions crossing the ranks
    inside the persistent CG kernel over peer memory.  Same result as one GPU to the CG tolerance (1e-10 here ->
    1e-7 on vectors), and both ranks must hold bit-identical values (they take the same convergence decisions)."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs 2 GPUs')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', '29535', os.path.join(ROOT, 'scripts', 'dist_grid_cg.py'), '20000']
    out = subprocess.check_output(cmd, timeout=900, stderr=subprocess.STDOUT).decode()
    line = [l for l in out.splitlines() if l.startswith('n=')][0]
    toks = line.split('|')[0].replace('rel', ' ').split()
    vals = {toks[i]: float(toks[i + 1]) for i in range(len(toks) - 1) if toks[i] in ('nlZ', 'dnlZ', 'dX', 'alpha', 'ranks_agree')}
    assert vals['nlZ'] < 1e-9 and vals['dnlZ'] < 1e-7 and vals['dX'] < 1e-7 and vals['alpha'] < 1e-7 and vals['ranks_agree'] == 0.0, line
