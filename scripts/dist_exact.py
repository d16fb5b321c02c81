"""torchrun check of the sharded exact path: every rank evaluates the same problem; results must equal the
single-GPU run (rank 0 also runs it unsharded on its own GPU) and the oracle at small n."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import numpy as np
import torch, torch.distributed as td
from kgp_b200 import B200GP, dist

rank, world, local = dist.env_rank()
torch.cuda.set_device(local)
td.init_process_group('nccl', device_id=torch.device('cuda', local))
def rel(a, b): return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))
for n in [int(a) for a in sys.argv[1:]] or [3000]:
    rng = np.random.RandomState(43); D = 8
    X = rng.randn(n, D); y = np.sin(X[:, 0] * X[:, 1]) + X[:, 2] ** 2 - X[:, 3] + 0.5 * np.cos(X.sum(1)); y = (y - y.mean()) / y.std() + 0.1 * rng.randn(n)
    hyp = {'lik': np.log(0.1), 'mean': [], 'cov': np.concatenate([np.log(2.0) * np.ones(D), [0.0]])[:, None]}
    def mk(attach):
        gp = B200GP(engine_kwargs={'device': local})
        gp.configure(D, dict(hyp), {}, 'infExact', 'meanZero', 'covSEard', 'likGauss', 'dlikExact', verbose=0)
        if attach: dist.attach(gp)
        gp.update_data('tr', X, y[:, None])
        return gp
    gpd = mk(True)
    for rep in range(2):
        td.barrier(); torch.cuda.synchronize(); t0 = time.time()
        nlZ, g = gpd.eng.eval(0, 2, gpd._n_hyp)
        torch.cuda.synchronize(); td.barrier(); t1 = time.time()
    dX = gpd.get_dlik_dx('tr'); st = gpd.stats()
    ymu, ys2 = gpd.predict(X[:300] + 0.1, return_var=True)
    if rank == 0:
        gps = mk(False)
        for rep in range(2):
            t2 = time.time(); nlZ1, g1 = gps.eng.eval(0, 2, gps._n_hyp); t3 = time.time()
        dX1 = gps.get_dlik_dx('tr'); ymu1, ys21 = gps.predict(X[:300] + 0.1, return_var=True)
        print('n=%d world=%d: sharded %.4fs (dev %.1f ms, potrf %.1f inv %.1f red %.1f) single %.4fs | nlZ %.12g vs %.12g rel %.2e dnlZ %.2e dX %.2e ymu %.2e ys2 %.2e' % (
            n, world, t1 - t0, st['ms_total'], st['ms_potrf'], st['ms_inverse'], st['ms_reduce'], t3 - t2, nlZ, nlZ1, rel(nlZ, nlZ1), rel(g, g1), rel(dX, dX1), rel(ymu, ymu1), rel(ys2, ys21)), flush=True)
        print('   stats:', {k: (round(v, 2) if isinstance(v, float) else v) for k, v in st.items()}, flush=True)
        del gps
    del gpd
# a matrix that is not positive definite (duplicated rows, vanishing noise): EVERY rank must report NOT_PD -- the owner of
# the failing panel is one rank only, the status is made global inside the library (all-reduce of the pivot flag)
if os.environ.get('KGP_DIST_NOTPD', '1') != '0':
    import warnings
    from kgp_b200 import _cabi
    rng = np.random.RandomState(5); n, D = 1500, 3
    X = rng.randn(n, D); X[700:1400] = X[:700]; y = rng.randn(n)
    gp = B200GP(engine_kwargs={'device': local})
    gp.configure(D, {'lik': -40.0, 'mean': [], 'cov': np.array([[0.0], [0.0]])}, {}, 'infExact', 'meanZero', 'covSEiso', 'likGauss', 'dlikExact', verbose=0)
    dist.attach(gp); gp.update_data('tr', X, y[:, None])
    code = 0
    try:
        gp.eng.eval(0, 1, gp._n_hyp)
    except _cabi.KgpError as e:
        code = e.code
    t = torch.tensor([code], dtype=torch.int64, device='cuda'); lo = t.clone(); hi = t.clone()
    td.all_reduce(lo, op=td.ReduceOp.MIN); td.all_reduce(hi, op=td.ReduceOp.MAX)
    with warnings.catch_warnings():
        warnings.simplefilter('ignore')
        nl = gp.evaluate('tr')                              # gp.m training mode: NaN + warning on every rank
    if rank == 0:
        print('NOTPD status min %d max %d (NOT_PD = %d) nlZ %s' % (int(lo), int(hi), _cabi.NOT_PD, nl), flush=True)
td.destroy_process_group()
