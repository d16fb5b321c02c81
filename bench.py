#!/usr/bin/env python
"""bench.py -- GP nlZ + full-gradient evaluations per second (BASELINE.json metric).

A *step* is one pass of the hot path over one synthetic batch: one exact-GP
evaluation [nlZ, dnlZ] plus dnlZ/dX -- what `train` + `get_dlik_dx` consume per
function evaluation.  Workload at N=1: BASELINE.json configs[1],
"mlp_kin40k" scale-up: exact GP, covSEard, n=40 000, d=8, FP64, on one B200.

  python bench.py [--gpus N --steps K --warmup W] [--n 40000]      # this engine
  python bench.py --impl reference ...                             # CPU arm (oracle port)

N>1 (torchrun): ONE n=40 000 problem is sharded over the ranks (block columns of the
Cholesky with NCCL panel broadcasts, rows of L^-T, tiles of B^-1) -> "scaling": "strong"; value is
evaluations per second of that one problem.  `--replicas` keeps one independent GP per GPU instead.
At N=1 the line also carries `kissgp`: the KISS-GP (infGrid) evaluation at n = 1 000 000
(BASELINE configs[4] scaled), with the CG kernel's algorithmic bytes against the HBM roofline.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# The CPU arm must use every host core at every N: torchrun exports OMP_NUM_THREADS=1 to its workers, which would hand
# the oracle's BLAS a single thread (round 1: `cores` 16 at N=1, 1 at N=2/4/8).  Set the BLAS pools BEFORE numpy loads.
_NCPU = os.cpu_count() or 1
if int(os.environ.get('RANK', 0)) == 0:
    for _k in ('OMP_NUM_THREADS', 'OPENBLAS_NUM_THREADS', 'MKL_NUM_THREADS'):
        os.environ[_k] = str(_NCPU)

import numpy as np


def synth_c2(n, D=8, seed=43):
    """kin40k-shaped synthetic inputs, standardised targets."""
    rng = np.random.RandomState(seed)
    X = rng.randn(n, D)
    y = np.sin(X[:, 0] * X[:, 1]) + X[:, 2] ** 2 - X[:, 3] + 0.5 * np.cos(X.sum(1))
    y = (y - y.mean()) / y.std() + 0.1 * rng.randn(n)
    hyp = {'cov': np.concatenate([np.log(2.0) * np.ones(D), [0.0]]), 'lik': float(np.log(0.1)), 'mean': np.zeros(0)}
    return X, y, hyp


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,'
         'clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,'
         'clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu = gpu_index; self.samples = []; self.stop_flag = False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.check_output(['nvidia-smi', '-i', str(self.gpu), '--query-gpu=' + self.Q,
                                               '--format=csv,noheader,nounits'], timeout=5).decode().strip()
                self.samples.append([t.strip() for t in out.split(',')])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.samples:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [nm for k, nm in enumerate(names) if any(s[3 + k].lower().startswith('active') for s in self.samples)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.samples[0][1]),
                'power_w_max': max(float(s[2]) for s in self.samples), 'reasons': reasons, 'samples': len(sm)}


def cpu_model():
    try:
        for ln in open('/proc/cpuinfo'):
            if ln.startswith('model name'):
                return ln.split(':', 1)[1].strip()
    except Exception:
        pass
    return 'unknown'


def use_all_cores():
    """BLAS pools at os.cpu_count() threads (belt and braces beside the environment variables above)."""
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=_NCPU)
    except Exception:
        pass


def cpu_fit(sizes, n_target, D=8):
    """Oracle evaluation time MEASURED at each of `sizes` on all host cores, split into the seconds inside Cholesky and
    the n-right-hand-side triangular solves (t3, grows as n^3) and the rest (t2: kernel matrix, exp, element-wise passes,
    the n^2 D gradient products -- grows as n^2).  The cost at n_target is t3(n_max) (n/n_max)^3 + t2(n_max) (n/n_max)^2:
    each component extrapolated with its own exponent from the LARGEST measured size.  `loo_error` applies the same rule
    from the second-largest size to the largest and compares with its measurement."""
    use_all_cores()
    cpu_eval_seconds(256)                                           # imports, BLAS thread pools
    ts = [cpu_eval_seconds(n, split=True) for n in sizes]
    tot = [t[0] for t in ts]; t3 = [t[1] for t in ts]; t2 = [a - b for a, b in zip(tot, t3)]
    r = float(n_target) / float(sizes[-1])
    out = {'model': 't(n) = t3(n_max) (n/n_max)^3 + t2(n_max) (n/n_max)^2, t3 = seconds inside Cholesky / n-RHS triangular solves',
           'n': list(map(int, sizes)), 'seconds': [float(v) for v in tot], 'seconds_blas3': [float(v) for v in t3],
           'predicted_seconds_at_target': t3[-1] * r ** 3 + t2[-1] * r ** 2, 'n_target': int(n_target),
           'n3_from_largest_seconds': float(tot[-1]) * r ** 3, 'n2_from_largest_seconds': float(tot[-1]) * r ** 2}
    if len(sizes) >= 2:
        q = float(sizes[-1]) / float(sizes[-2])
        pred = t3[-2] * q ** 3 + t2[-2] * q ** 2
        out['loo_error'] = {'n': int(sizes[-1]), 'predicted_s': pred, 'measured_s': float(tot[-1]), 'rel': pred / float(tot[-1]) - 1.0}
        ln, lt = np.log(np.asarray(sizes, dtype=np.float64)), np.log(np.asarray(tot))
        out['local_exponent_last_two'] = float((lt[-1] - lt[-2]) / (ln[-1] - ln[-2]))
    return out


_BLAS3 = {'s': 0.0}


def _timed_blas3(orc):
    """Wrap the oracle's LAPACK/BLAS-3 entry points (Cholesky, triangular solves with n right-hand sides) with timers, so
    that an evaluation's time splits into its O(n^3) part and the O(n^2 D) rest (kernel matrix, exp, reductions)."""
    sla = orc.sla
    if getattr(sla, '_kgp_timed', False):
        return
    def wrap(fn):
        def inner(*a, **k):
            t0 = time.perf_counter()
            try:
                return fn(*a, **k)
            finally:
                _BLAS3['s'] += time.perf_counter() - t0
        return inner
    class Timed(object):
        pass
    t = Timed()
    for name in dir(sla):
        if not name.startswith('_'):
            setattr(t, name, getattr(sla, name))
    t.cholesky = wrap(sla.cholesky); t.solve_triangular = wrap(sla.solve_triangular); t._kgp_timed = True
    orc.sla = t


def cpu_eval_seconds(n_s, D=8, reps=1, split=False):
    """Oracle (CPU FP64 restatement of the GPML path) on a bounded sample: one evaluation =
    gp() [nlZ, dnlZ] + dlikExact (which re-runs the inference, dlikExact.m:14).  split=True also returns the seconds
    spent inside Cholesky / n-RHS triangular solves."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import gpml_exact as orc
    _timed_blas3(orc)
    X, y, hyp = synth_c2(n_s, D)
    best = (float('inf'), 0.0)
    for _ in range(reps):
        _BLAS3['s'] = 0.0
        t0 = time.perf_counter()
        orc.inf_exact(hyp, ['meanZero'], 'ard', X, y, want_grad=True)
        orc.dlik_exact(hyp, ['meanZero'], 'ard', X, y)
        tot = time.perf_counter() - t0
        if tot < best[0]:
            best = (tot, _BLAS3['s'])
    return best if split else best[0]


def synth_c5(n, seed=46, Q=6):
    """spectral-mixture KISS-GP on a 2-D feature grid."""
    rng = np.random.RandomState(seed)
    H = rng.rand(n, 2); y = np.sin(2 * np.pi * 3 * H).sum(1) + 0.5 * H[:, 0] + 0.3 * rng.randn(n)
    hyp_cov = [[np.log(np.ones((Q, 1)))], [np.log(np.maximum(0.05 * rng.rand(Q, 1), 1e-8))], [np.log(np.abs(rng.randn(Q, 1) + 1))]]
    return H, y, hyp_cov


def kissgp_cpu_sample(n, cg_iters_needed, sample_iters=500):
    """CPU arm of the KISS-GP leg: the oracle's pieces (scipy.sparse CSR Mx like the reference's sparse path, dense
    Kronecker factors) on the host cores.  A full solve needs thousands of CG iterations, so the sample times the set-up
    (grid, interpolation matrix, factors) once and `sample_iters` iterations of B = I + sW Mx Kg Mx' sW, and
    extrapolates to the iteration count the engine needed.  One evaluation = gp() + dlikGrid() = two inferences
    (dlikGrid.m:29-36 re-runs the inference)."""
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    import gpml_grid as gg
    H, y, hc = synth_c5(n)
    hv = np.tile(np.concatenate([np.asarray(a, dtype=np.float64).reshape(-1) for a in hc])[:, None], (1, 2)).flatten(order='F')
    t0 = time.perf_counter()
    xg = gg.creategrid(H, 1, 70.)
    Mx = gg.interpgrid(xg, H); MxT = Mx.T.tocsr()
    hs = gg.split_hyp('covSM', hv, 2, 6)
    Ks = [gg.factor_cov('covSM', hs[i], xg[i], 6)[0] for i in range(2)]
    t_setup = time.perf_counter() - t0
    sw = 1.0 / 0.3
    v = y.copy()
    t0 = time.perf_counter()
    for _ in range(sample_iters):
        z = v + sw * (Mx @ gg.kronmvm(Ks, MxT @ (sw * v)))
        a = (v @ v) / (v @ z); v = v - a * z                  # the vector work of one CG iteration (apx.m:290-300)
    t_iter = (time.perf_counter() - t0) / sample_iters
    per_inference = t_setup + cg_iters_needed * t_iter
    return {'value': 1.0 / (2.0 * per_inference), 'unit': 'evals/s', 'cores': cpu_threads(), 'kind': 'port',
            'sample': 'oracle set-up (grid + CSR Mx + factors) %.2f s once + %d CG iterations at %.1f ms each, extrapolated to %d iterations; x2 inferences per evaluation (gp + dlikGrid)'
                      % (t_setup, sample_iters, 1e3 * t_iter, int(cg_iters_needed))}


def kissgp_leg(n, device, steps, warmup, with_cpu=True):
    """Secondary headline: KISS-GP (infGrid, covSM Q=6 x 2, meanSum, 70 x 70 grid, Fiedler log-det) nlZ + dnlZ + dnlZ/dX
    at n points.  The dominant kernel is the persistent CG kernel; its roofline is HBM: algorithmic bytes per
    iteration n*(16*D + 88) + 32*N over the measured copy bandwidth."""
    from kgp_b200 import B200GP
    H, y, hc = synth_c5(n)
    gp = B200GP(engine_kwargs={'device': device})
    maxit = 30000       # the reference's 2000 cannot converge at n = 1M without a preconditioner (cond ~ 1e7)
    gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc}, {'cg_maxit': maxit, 'cg_tol': 1e-6}, 'infGrid',
                 'meanSum', 'covSM', 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6],
                 mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    N = int(np.prod([a.size for a in gp.grid()]))
    for _ in range(max(1, warmup)):
        gp.eng.eval(0, 2, gp._n_hyp)
    ms = 0.0; cg_ms = 0.0; iters = 0
    for _ in range(steps):
        nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
        st = gp.stats(); ms += st['ms_total']; cg_ms += st['ms_solve']; iters += st['cg_iters'] - 1
    t0 = time.perf_counter()
    for _ in range(steps):
        gp.update_data('tr', H, y[:, None])
        nl = gp.evaluate('tr'); dX = gp.get_dlik_dx('tr')
    e2e = steps / (time.perf_counter() - t0)
    # the same solve with FP64 atomics in the scatter (opt.cg_deterministic = 0): faster, not bit-reproducible
    gpf = B200GP(engine_kwargs={'device': device})
    gpf.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc.copy()}, {'cg_maxit': maxit, 'cg_tol': 1e-6, 'cg_deterministic': 0},
                  'infGrid', 'meanSum', 'covSM', 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6],
                  mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    gpf.update_data('tr', H, y[:, None]); gpf.update_grid('tr')
    gpf.eng.eval(0, 2, gpf._n_hyp)
    fms = 0.0; fcg = 0.0; fit_ = 0; fiters = []
    for _ in range(steps):
        nlZf, _g = gpf.eng.eval(0, 2, gpf._n_hyp)
        st = gpf.stats(); fms += st['ms_total']; fcg += st['ms_solve']; fit_ += st['cg_iters'] - 1; fiters.append(int(st['cg_iters']))
    fp_atomics = {'value': steps * 1e3 / fms, 'unit': 'evals/s', 'ms_per_eval': fms / steps, 'cg_us_per_iteration': 1e3 * fcg / max(fit_, 1),
                  'cg_iterations_seen': fiters, 'rel_diff_nlZ_vs_deterministic': abs(nlZf - nlZ) / abs(nlZ),
                  'note': 'opt.cg_deterministic=0: FP64 REDs instead of the order-independent fixed-point integer REDs; iteration counts vary from run to run'}
    del gpf
    # same problem through the grid-space Woodbury preconditioner (opt.cg_precond, an extension: SURVEY a21(i)); same
    # stopping rule, so the result is the same alpha to cg_tol -- only the iteration count changes
    gp2 = B200GP(engine_kwargs={'device': device})
    gp2.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc.copy()}, {'cg_maxit': maxit, 'cg_tol': 1e-6, 'cg_precond': 1},
                  'infGrid', 'meanSum', 'covSM', 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6],
                  mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    gp2.update_data('tr', H, y[:, None]); gp2.update_grid('tr')
    for _ in range(max(1, warmup)):
        gp2.eng.eval(0, 2, gp2._n_hyp)
    pms = 0.0; pit = 0
    v0 = gp2._hyp_vec().copy()
    for i in range(steps):
        # hyper-parameters move a little between evaluations, as inside an optimiser run: the eigen-decompositions are
        # warm-started from the previous evaluation and must not be handed an unchanged problem
        gp2._set_hyp_vec(v0 + 2e-3 * (1 if i % 2 else -1)); gp2._push_model()
        gp2.eng.eval(0, 2, gp2._n_hyp)
        st = gp2.stats(); pms += st['ms_total']; pit += st['cg_iters'] - 1
    gp2._set_hyp_vec(v0); gp2._push_model()
    nlZp, gpre = gp2.eng.eval(0, 2, gp2._n_hyp)
    pre = {'value': steps * 1e3 / pms, 'unit': 'evals/s', 'ms_per_eval': pms / steps, 'cg_iterations_per_eval': pit / steps,
           'nlZ': nlZp, 'rel_diff_nlZ_vs_reference_cg': abs(nlZp - nlZ) / abs(nlZ),
           'rel_diff_dnlZ_vs_reference_cg': float(np.max(np.abs(gpre - g)) / np.max(np.abs(g))),
           'note': 'opt.cg_precond=1: B^-1 ~ I - s2 M F C^-1 F\' M\' with C on the kept Kronecker modes, factored by the exact path\'s DMMA Cholesky each evaluation; hyper-parameters perturbed by +-2e-3 between timed evaluations'}
    del gp2
    bytes_iter = n * (16 * 2 + 88) + 32 * N
    peak = 6455.3; src = 'B200_PROFILING.md fallback'
    try:
        peak = float(json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json')))['hbm_gbs']); src = 'MEASURED_PEAKS.json hbm_gbs'
    except Exception:
        pass
    ach = bytes_iter * iters / (cg_ms * 1e-3) / 1e9
    return {'workload': 'KISS-GP infGrid covSM(Q=6)x2 + meanSum, n=%d, 70x70 grid (N=%d), Fiedler log-det, cg_tol 1e-6: nlZ + dnlZ + dnlZ/dX' % (n, N),
            'value': steps * 1e3 / ms, 'unit': 'evals/s', 'ms_per_eval': ms / steps, 'e2e_evals_per_s': e2e,
            'cg_iterations_per_eval': iters / steps, 'cg_us_per_iteration': 1e3 * cg_ms / max(iters, 1), 'cg_share_of_eval': cg_ms / ms,
            'roofline': {'bound': 'hbm', 'kernel': 'cg_persistent2_kernel (whole CG solve, one cooperative launch)', 'achieved': ach, 'peak': peak,
                         'unit': 'GB/s', 'frac': ach / peak, 'bytes_per_iteration': bytes_iter, 'peak_source': src,
                         'note': 'algorithmic bytes; the working set (48 MB at n=1M) stays in L2 / shared memory, DRAM traffic per iteration is ~0 (profiles/)'},
            'nlZ': nlZ, 'deterministic': True, 'fp_atomics': fp_atomics, 'preconditioned': pre,
            'parity': 'unpinned by the reference: no infGrid number is recorded anywhere in the reference tree (SURVEY 8(c)); the engine is '
                      'compared with the CPU restatement, which is tied to the KAT-pinned exact oracle on a fine grid (tests/test_oracle_grid.py)',
            'real_bound': 'issue / dependent-FP64 latency and block barriers, not HBM: ncu (profiles/r02_cg_persistent2_ncu.md) shows 42 MB of DRAM '
                          'traffic for the WHOLE solve (~8 kB per iteration; the 48 MB working set lives in shared memory / L2), issue slots 49 % busy, '
                          'FP64 pipe 16 %, dominant stall = CTA barrier (3.3 point units per warp -> a ragged last round)',
            'cpu_baseline': kissgp_cpu_sample(n, iters / steps) if with_cpu else None}


def small_config_legs(device):
    """The latency-bound reference-sized configurations: C1 exact n=512 d=32 (README demo) and C4
    KISS-GP n=460 1-D, k=100: device microseconds and kernel launches per nlZ + dnlZ + dnlZ/dX evaluation."""
    from kgp_b200 import B200GP
    out = {}
    rng = np.random.RandomState(42)
    X = np.tanh(rng.randn(512, 32)); y = np.sin(X.sum(1) / 4) + 0.1 * rng.randn(512)
    gp = B200GP(engine_kwargs={'device': device})
    gp.configure(32, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, {}, 'infExact', 'meanZero', 'covSEiso', 'likGauss', 'dlikExact', verbose=0)
    gp.update_data('tr', X, y[:, None])
    for _ in range(5):
        gp.eng.eval(0, 2, gp._n_hyp)
    t = []; 
    for _ in range(20):
        gp.eng.eval(0, 2, gp._n_hyp); st = gp.stats(); t.append(st['ms_total'])
    t0 = time.perf_counter()
    for _ in range(20):
        gp.eng.eval(0, 2, gp._n_hyp)
    out['c1_exact_n512_d32'] = {'device_us_per_eval': 1e3 * float(np.median(t)), 'wall_us_per_eval': 1e6 * (time.perf_counter() - t0) / 20,
                                'kernel_launches_per_eval': int(st['kernel_launches'])}
    del gp
    rng = np.random.RandomState(45)
    H = np.tanh(rng.randn(460, 1)); y = np.sin(4 * H[:, 0]) + np.exp(-2) * rng.randn(460)
    gp = B200GP(engine_kwargs={'device': device})
    gp.configure(1, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, {'cg_maxit': 500, 'cg_tol': 1e-4}, 'infGrid', 'meanZero', 'covSEiso', 'likGauss',
                 'dlikGrid', grid_kwargs={'eq': 1, 'k': 100.}, verbose=0)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    for _ in range(5):
        gp.eng.eval(0, 2, gp._n_hyp)
    t = []; v0 = gp._hyp_vec().copy()
    for i in range(20):
        gp._set_hyp_vec(v0 + 2e-3 * (1 if i % 2 else -1)); gp._push_model()        # (see kissgp_leg: no unchanged problem for the warm start)
        gp.eng.eval(0, 2, gp._n_hyp); st = gp.stats(); t.append(st['ms_total'])
    t0 = time.perf_counter()
    for i in range(20):
        gp._set_hyp_vec(v0 + 2e-3 * (1 if i % 2 else -1)); gp._push_model()
        gp.eng.eval(0, 2, gp._n_hyp)
    out['c4_kissgp_n460_k100'] = {'device_us_per_eval': 1e3 * float(np.median(t)), 'wall_us_per_eval': 1e6 * (time.perf_counter() - t0) / 20,
                                  'kernel_launches_per_eval': int(st['kernel_launches']), 'cg_iterations': int(st['cg_iters'])}
    return out


def epoch_leg(device):
    """training epochs of the reference-sized BASELINE configs through the PyTorch restatement of UpdateGP +
    gen_gp_loss (kgp_b200/harness.py): C1 README GP-RNN (SimpleRNN(32) -> exact GP covSEiso, (10,2) sequences, 512 train,
    batch 32), C3 msgp_mlp_kin40k (MLP -> KISS-GP covSEiso on a 2-D feature grid, n = 40 000, k = 70, cg 2000/1e-6) and
    C4 msgp_lstm_actuator (LSTM(16) -> KISS-GP, n = 460, k = 100, cg 500/1e-4); gp_n_iter = 5 each.  Reports the epoch
    time and the GP-update share, i.e. what the reference pays through oct2py per epoch."""
    import torch
    from kgp_b200 import harness
    out = {}

    def run(name, net, head, Xt, y, opt, bs, desc):
        try:
            run_(name, net, head, Xt, y, opt, bs, desc)
        except Exception as ex:                           # e.g. CG not converged inside get_dlik_dx: raised like dlikGrid.m would
            out[name] = {'config': desc, 'error': str(ex)}

    def run_(name, net, head, Xt, y, opt, bs, desc):
        harness.fit(net, head, Xt, y, opt, epochs=1, batch_size=bs, gp_n_iter=5)            # warm-up epoch
        torch.cuda.synchronize(); t0 = time.perf_counter()
        hist = harness.fit(net, head, Xt, y, opt, epochs=3, batch_size=bs, gp_n_iter=5)
        torch.cuda.synchronize(); per = (time.perf_counter() - t0) / 3
        out[name] = {'config': desc, 'epoch_ms': 1e3 * per, 'gp_update_ms': 1e3 * float(np.mean([h['gp_update_elapsed'] for h in hist])),
                     'nlml_last': float(hist[-1]['nlml'])}

    torch.manual_seed(0); rng = np.random.RandomState(42)
    X = rng.randn(512, 10, 2).astype(np.float32); y = np.sin(X.sum((1, 2)) / 4)[:, None] + 0.1 * rng.randn(512, 1)
    net = harness.make_net('rnn', (10, 2), hidden=32).cuda(device)
    head = harness.GPHead(32, {'lik': -2.0, 'mean': [], 'cov': np.array([[-0.7], [0.0]])}, inf='infExact', cov='covSEiso', device=device)
    run('c1_gp_rnn', net, head, torch.from_numpy(X).cuda(device), y, torch.optim.RMSprop(net.parameters(), lr=1e-3), 32,
        'README GP-RNN n=512, 16 mini-batches, gp_n_iter=5')
    torch.manual_seed(1); rng = np.random.RandomState(44)
    X = rng.randn(40000, 8).astype(np.float32); y = (np.sin(X[:, 0] * X[:, 1]) + X[:, 2])[:, None] + 0.1 * rng.randn(40000, 1)
    net = harness.make_net('mlp', (8,), hidden=16, out_dim=2).cuda(device)
    head = harness.GPHead(2, {'lik': np.log(0.3), 'mean': [], 'cov': np.array([[0.5], [1.0]])}, inf='infGrid', cov='covSEiso',
                          opt={'cg_maxit': 20000, 'cg_tol': 1e-6}, grid_kwargs={'eq': 1, 'k': 70.}, update_grid=1, device=device)
    run('c3_msgp_mlp', net, head, torch.from_numpy(X).cuda(device), y, torch.optim.Adam(net.parameters(), lr=1e-3), 128,
        'msgp_mlp_kin40k n=40000, 2-D grid 70x70, 313 mini-batches, gp_n_iter=5 (cg_maxit 20000: with the script\'s 2000 the un-preconditioned '
        'CG does not converge on these synthetic features and dlikGrid raises, as it would in the reference)')
    torch.manual_seed(3); rng = np.random.RandomState(45)
    u = np.cumsum(rng.randn(1024 + 33)) * 0.1; yy = np.tanh(np.convolve(u, np.ones(8) / 8, 'same'))
    X = np.stack([u[i:i + 32] for i in range(460)])[:, :, None].astype(np.float32); y = yy[32:32 + 460][:, None]
    net = harness.make_net('lstm', (32, 1), hidden=16, out_dim=1).cuda(device)
    head = harness.GPHead(1, {'lik': -2.0, 'mean': [], 'cov': [[-0.7], [0.0]]}, inf='infGrid', cov='covSEiso',
                          opt={'cg_maxit': 500, 'cg_tol': 1e-4}, grid_kwargs={'eq': 1, 'k': 100.}, update_grid=1, device=device)
    run('c4_gp_lstm', net, head, torch.from_numpy(X).cuda(device), y, torch.optim.Adam(net.parameters(), lr=1e-2), 128,
        'msgp_lstm_actuator n=460, grid k=100, 4 mini-batches, gp_n_iter=5')
    return out


def kissgp_sharded_leg(n_total, device, steps, warmup, world):
    """N > 1: the same KISS-GP evaluation with the points ROW-SHARDED over the ranks (SURVEY 8(e) KISS-GP (ii)), weak
    scaling: n_total = kissgp_n x N points.  The two reductions of every CG iteration (grid vector, four scalars)
    cross the ranks inside the persistent kernel over peer memory (NVLink); every rank calls this."""
    import torch
    import torch.distributed as td
    from kgp_b200 import B200GP, dist as kdist
    H, y, hc = synth_c5(n_total)
    gp = B200GP(engine_kwargs={'device': device})
    gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc}, {'cg_maxit': 200000, 'cg_tol': 1e-6}, 'infGrid',
                 'meanSum', 'covSM', 'likGauss', 'dlikGrid', grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6],
                 mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    kdist.attach(gp)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    for _ in range(max(1, min(warmup, 2))):
        gp.eng.eval(0, 2, gp._n_hyp)
    ms = 0.0; cg_ms = 0.0; iters = 0
    for _ in range(steps):
        td.barrier(); torch.cuda.synchronize()
        nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
        st = gp.stats(); ms += st['ms_total']; cg_ms += st['ms_solve']; iters += st['cg_iters'] - 1
    t = torch.tensor([ms, cg_ms], dtype=torch.float64, device='cuda')
    td.all_reduce(t, op=td.ReduceOp.MAX)
    ms, cg_ms = float(t[0]), float(t[1])
    N = int(np.prod([a.size for a in gp.grid()]))
    return {'workload': 'KISS-GP infGrid covSM(Q=6)x2 + meanSum, n=%d (= %d per GPU), 70x70 grid (N=%d), Fiedler log-det, cg_tol 1e-6' % (n_total, n_total // world, N),
            'scaling': 'weak', 'parallelism': 'points row-sharded over %d GPUs; per CG iteration one grid-vector and one 4-scalar all-reduce inside the persistent kernel over NVLink peer memory (LL-style tagged packets), alpha all-reduced once' % world,
            'value': steps * 1e3 / ms, 'unit': 'evals/s', 'ms_per_eval': ms / steps, 'cg_iterations_per_eval': iters / steps,
            'cg_us_per_iteration': 1e3 * cg_ms / max(iters, 1), 'point_iterations_per_s': n_total * iters / (cg_ms * 1e-3), 'nlZ': nlZ}


def kissgp_probe_leg(n, device, steps, warmup, world):
    """SURVEY 8(e) KISS-GP (i): the path north_star says shards naturally -- the log-det probes.  Same C5 problem with the
    Chebyshev-Hutchinson log-det (opt.ldB2_cheby, m = 64 Rademacher probes, degree 15) and the preconditioned alpha solve;
    the probes are split contiguously over the ranks, every rank runs its Krylov recurrences with zero communication, the
    grid-sized accumulators are all-reduced once.  STRONG scaling: the same problem at every N."""
    import torch
    import torch.distributed as td
    from kgp_b200 import B200GP, dist as kdist
    H, y, hc = synth_c5(n)
    gp = B200GP(engine_kwargs={'device': device})
    opt = {'cg_maxit': 30000, 'cg_tol': 1e-6, 'cg_precond': 1, 'ldB2_cheby': 1, 'ldB2_cheby_hutch': 64, 'ldB2_cheby_degree': 15, 'ldB2_cheby_seed': 7}
    gp.configure(2, {'lik': np.log(0.3), 'mean': [[0.], [0.], [0.]], 'cov': hc}, opt, 'infGrid', 'meanSum', 'covSM', 'likGauss', 'dlikGrid',
                 grid_kwargs={'eq': 1, 'k': 70.}, cov_args=[6], mean_args=['{@meanLinear, @meanConst}'], verbose=0)
    if world > 1:
        kdist.attach(gp)
    gp.update_data('tr', H, y[:, None]); gp.update_grid('tr')
    for _ in range(max(1, min(warmup, 2))):
        gp.eng.eval(0, 2, gp._n_hyp)
    ms = 0.0; ld_ms = 0.0
    for _ in range(steps):
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()
        nlZ, g = gp.eng.eval(0, 2, gp._n_hyp)
        st = gp.stats(); ms += st['ms_total']; ld_ms += st['ms_reduce']
    if world > 1:
        t = torch.tensor([ms, ld_ms], dtype=torch.float64, device='cuda'); td.all_reduce(t, op=td.ReduceOp.MAX); ms, ld_ms = float(t[0]), float(t[1])
    return {'workload': 'KISS-GP n=%d, 70x70 grid, Chebyshev-Hutchinson log-det with 64 probes (degree 15) sharded over %d GPU(s), '
                        'Woodbury-preconditioned alpha solve replicated' % (n, world), 'scaling': 'strong',
            'value': steps * 1e3 / ms, 'unit': 'evals/s', 'ms_per_eval': ms / steps, 'logdet_and_derivatives_ms': ld_ms / steps, 'nlZ': nlZ,
            'note': 'efficiency = ms_per_eval(N=1) / (N x ms_per_eval(N)) from the per-N lines; the replicated part (alpha solve, eigen-decompositions, dirder) bounds it'}


def cpu_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([p.get('num_threads', 1) for p in threadpool_info()] + [1])
    except Exception:
        return os.cpu_count() or 1


def run_reference(args, rank, world):
    """CPU arm: the oracle port (the reference is Octave .m code and cannot run here) on ALL host cores.
    The stated config (n = 40 000) costs ~4.7 n^3 = 3e14 flop per evaluation -- ~15 minutes on this box's 16 cores, and
    the reference itself could not run it at all (dlikExact's n x n x D tensor is 102 GB) -- so the arm MEASURES:
      * a size ladder n in --cpu-fit-sizes once (default 2048..16384), fits t(n) = a n^3 + b n^2 and reports the fit's
        leave-one-out error;
      * K timed steps, each one full evaluation at --cpu-sample-n (the bounded sample of the workload).
    `value` = evaluations/s at the stated n from the fitted model, re-anchored on the timed steps
    (value = 1 / (t_step * model(n) / model(n_sample))); `ms_per_step` is the MEASURED time of a step at the sample size,
    so steps x ms_per_step is the real duration of the timed region.  profiles/ holds one direct measurement at
    n = 40 000 to check the model against."""
    if rank != 0:
        return
    use_all_cores()
    n, n_s = args.n, args.cpu_sample_n
    sizes = [int(v) for v in args.cpu_fit_sizes.split(',') if v]
    t0 = time.perf_counter()
    fit = cpu_fit(sizes, n)
    fit['ladder_wall_s'] = time.perf_counter() - t0
    for _ in range(args.warmup):
        cpu_eval_seconds(n_s)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_eval_seconds(n_s)
    per = (time.perf_counter() - t0) / max(args.steps, 1)
    sec_n = fit['predicted_seconds_at_target']
    val = 1.0 / sec_n
    sample = ('one step = oracle gp()+dlikExact at n=%d (d=8, covSEard), measured %.2f s; evals/s at n=%d from the measured cost '
              'model (BLAS-3 seconds x (n/n_max)^3 + the rest x (n/n_max)^2, measured on n in %s; the same rule predicts n=%d from the size below it with %+.1f%% error), i.e. %.0f s per evaluation'
              % (n_s, per, n, sizes, sizes[-1], 100 * fit.get('loo_error', {}).get('rel', float('nan')), sec_n))
    line = {'impl': 'reference', 'metric': 'gp_nlZ_grad_evals_per_sec', 'value': val, 'unit': 'evals/s',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': per * 1e3,
            'higher_is_better': True, 'scaling': 'weak' if args.replicas else 'strong', 'vs_baseline': None,
            'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'exact GP covSEard infExact n=%d d=8 (BASELINE configs[1])' % n,
                       'note': 'reference GPML/Octave unavailable on this box; CPU arm = NumPy/SciPy port (oracle/); each step is a '
                               'bounded sample (n=%d); value is for n=%d through the measured cost model' % (n_s, n)},
            'value_basis': 'measured-fit', 'seconds_per_eval_at_n': sec_n, 'fit': fit,
            'cpu_baseline': {'value': val, 'unit': 'evals/s', 'cores': cpu_threads(), 'os_cpu_count': _NCPU, 'cpu_model': cpu_model(),
                             'kind': 'port', 'sample': sample},
            'e2e': {'value': val, 'unit': 'evals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--n', type=int, default=40000)
    ap.add_argument('--cpu-sample-n', type=int, default=4096)
    ap.add_argument('--cpu-fit-sizes', default='2048,4096,8192,16384', help='--impl reference: sizes of the measured cost ladder')
    ap.add_argument('--no-cpu', action='store_true', help='skip the cpu_baseline leg')
    ap.add_argument('--replicas', action='store_true', help='N>1: one independent GP per GPU instead of sharding one problem')
    ap.add_argument('--no-kissgp', action='store_true', help='skip the secondary KISS-GP n=1M measurement')
    ap.add_argument('--kissgp-n', type=int, default=1000000)
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0)); world = int(os.environ.get('WORLD_SIZE', 1))
    local = int(os.environ.get('LOCAL_RANK', 0))

    if args.impl == 'reference':
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import __graft_entry__ as g
    if not os.path.isfile(os.path.join(ROOT, 'kgp_b200', 'libkgp_b200.so')):
        g.build()
    from kgp_b200 import B200GP
    if world > 1:
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        torch.cuda.set_device(local)
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))

    n, D = args.n, 8
    sharded = world > 1 and not args.replicas
    X, y, hyp = synth_c2(n, D, seed=43 + (0 if sharded else rank))
    gp = B200GP(engine_kwargs={'device': local})
    gp.configure(D, {'lik': hyp['lik'], 'mean': [], 'cov': hyp['cov'][:, None].copy()}, {}, 'infExact', 'meanZero',
                 'covSEard', 'likGauss', 'dlikExact', verbose=0)
    if sharded:
        from kgp_b200 import dist as kdist
        kdist.attach(gp)
    gp.update_data('tr', X, y[:, None])
    eng = gp.eng
    nh = D + 2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing (value): K evaluations, per-launch profiling OFF (look-ahead Cholesky active) ------------
    for _ in range(args.warmup):
        eng.eval(0, 2, nh)
    sampler = ClockSampler(local); sampler.start()
    barrier()
    dev_ms = 0.0; launches = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        nlZ, grad = eng.eval(0, 2, nh)                 # blocking; device time from CUDA events inside
        st = eng.stats()
        dev_ms += st['ms_total']; launches += st['kernel_launches']
    phases = {k: st[k] for k in ('ms_kbuild', 'ms_potrf', 'ms_inverse', 'ms_solve', 'ms_reduce')}      # (last timed step, this rank)
    barrier()
    wall = time.perf_counter() - t0
    # ---- the same evaluation with an event pair around every GEMM launch (roofline of the dominant kernel); still inside
    #      the clock-sampled region, but not part of `value`: the event pairs serialise the look-ahead overlap
    eng.set_profiling(True)
    gemm_ms = 0.0; gemm_fl = 0.0; gemm_l = 0; prof_ms = 0.0; cls_ms = np.zeros(8); cls_fl = np.zeros(8)
    nprof = 2 if args.steps >= 2 else 1
    for _ in range(nprof):
        eng.eval(0, 2, nh)
        st = eng.stats()
        gemm_ms += st['ms_gemm']; gemm_fl += st['gemm_flops']; gemm_l += st['gemm_launches']; prof_ms += st['ms_total']
        cls_ms += np.asarray(st['gemm_class_ms']); cls_fl += np.asarray(st['gemm_class_flops'])
    eng.set_profiling(False)
    sampler.stop_flag = True; sampler.join(timeout=3)
    t = torch.tensor([dev_ms], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms_max = float(t.item())
    ms_per_step = dev_ms_max / args.steps
    value = (1 if sharded else world) * 1e3 / ms_per_step

    # ---- multi-GPU invariance, driver-visible: rank 0 evaluates the same problem unsharded on its own GPU ---------------
    invariance = None
    if sharded:
        dX_sh = gp.get_dlik_dx('tr')
        if rank == 0:
            g1 = B200GP(engine_kwargs={'device': local})
            g1.configure(D, {'lik': hyp['lik'], 'mean': [], 'cov': hyp['cov'][:, None].copy()}, {}, 'infExact', 'meanZero',
                         'covSEard', 'likGauss', 'dlikExact', verbose=0)
            g1.update_data('tr', X, y[:, None])
            nl1, gr1 = g1.eng.eval(0, 2, nh); dX1 = g1.get_dlik_dx('tr')
            rl = lambda a, b: float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / np.max(np.abs(b)))
            invariance = {'vs': 'the same n=%d problem evaluated unsharded on rank 0\'s GPU' % n, 'rel_nlZ': rl(nlZ, nl1),
                          'rel_dnlZ': rl(grad, gr1), 'rel_dX': rl(dX_sh, dX1), 'ms_single': g1.stats()['ms_total']}
            del g1
        barrier()

    # ---- end to end through the public API with host buffers -------------------------------
    Xh = torch.from_numpy(X).pin_memory().numpy(); yh = torch.from_numpy(y[:, None].copy()).pin_memory().numpy()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        gp.update_data('tr', Xh, yh)                  # H2D of this step's inputs
        nl = gp.evaluate('tr')                        # nlZ (+ dnlZ, dX cached on device)
        dX = gp.get_dlik_dx('tr')                     # D2H n x D
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = (1 if sharded else world) * args.steps / float(t.item())

    kiss_sh = None
    if sharded and not args.no_kissgp:
        del gp, eng
        gp = eng = None
        kiss_sh = kissgp_sharded_leg(args.kissgp_n * world, local, args.steps, args.warmup, world)
        kiss_sh['probe_sharded'] = kissgp_probe_leg(args.kissgp_n, local, args.steps, args.warmup, world)
        gp = B200GP(engine_kwargs={'device': local}); eng = gp.eng          # (for the peak micro-benchmark below)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (dgemm_nt_kernel, FP64 DMMA) -----------------------
    peak_tf = eng.bench_dmma_peak(20000)
    try:
        a = torch.randn(8192, 8192, dtype=torch.float64, device='cuda')
        torch.matmul(a, a); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
        e0.record(); torch.matmul(a, a); torch.matmul(a, a); e1.record(); torch.cuda.synchronize()
        cublas_tf = 2 * 2 * 8192.0 ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
        del a
    except Exception:
        cublas_tf = None
    achieved = gemm_fl / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else None
    traffic = None; traffic_note = None
    tp = os.path.join(ROOT, 'profiles', 'r02_dgemm_traffic.json')
    if os.path.isfile(tp):
        try:
            tj = json.load(open(tp))
            traffic = tj.get('dram_bytes_per_launch'); traffic_note = tj.get('note')
        except Exception:
            traffic = None
    names = ['potrf panel update', 'potrf trailing SYRK', 'Z-sweep GEMM', 'trsm block update', 'Z Z^T', 'predict']
    classes = {names[i]: {'ms_per_eval': cls_ms[i] / nprof, 'tflops': cls_fl[i] / cls_ms[i] / 1e9} for i in range(6) if cls_ms[i] > 0}
    roofline = {'bound': 'tensor', 'kernel': 'dgemm_nt_tma_kernel (FP64 DMMA.8x8x4 fed by TMA + mbarrier ring)', 'achieved': achieved, 'peak': peak_tf,
                'unit': 'TFLOP/s', 'frac': (achieved / peak_tf) if achieved else None, 'traffic': traffic, 'traffic_note': traffic_note,
                'peak_source': 'FP64 DMMA issue-rate microbenchmark run in this process (MEASURED_PEAKS.json has no FP64 entry)',
                'cublas_dgemm_8192_tflops': cublas_tf,
                'flops_per_launch': gemm_fl / max(gemm_l, 1), 'ms_per_launch': gemm_ms / max(gemm_l, 1),
                'launches_per_step': gemm_l / nprof, 'share_of_step': gemm_ms / prof_ms if prof_ms else None,
                'measured_in': '%d profiled evaluations right after the timed ones (event pair per launch; look-ahead off while profiling)' % nprof,
                'call_sites': classes, 'whole_eval_tflops': (gemm_fl / nprof) / (ms_per_step * 1e-3) / 1e12}

    kiss = kiss_sh; small = None
    if world == 1 and not args.no_kissgp:
        del gp, eng
        kiss = kissgp_leg(args.kissgp_n, local, args.steps, args.warmup, with_cpu=not args.no_cpu)
        kiss['probe_sharded'] = kissgp_probe_leg(args.kissgp_n, local, min(args.steps, 3), args.warmup, 1)
        try:
            small = small_config_legs(local)
        except Exception as ex:                       # secondary information only
            small = {'error': str(ex)}
        try:
            small['epochs'] = epoch_leg(local)
        except Exception as ex:
            small['epochs'] = {'error': str(ex)}

    cpu = None
    if not args.no_cpu:
        # bounded sample (~25 s of CPU work): the oracle measured at three sizes, cost model fitted, evaluated at n
        fit = cpu_fit([2048, 4096, 8192], n)
        sec = fit['predicted_seconds_at_target']
        cpu = {'value': 1.0 / sec, 'unit': 'evals/s', 'cores': cpu_threads(), 'os_cpu_count': _NCPU, 'cpu_model': cpu_model(), 'kind': 'port',
               'sample': 'oracle gp()+dlikExact measured at n=%s: %s s; BLAS-3 part x (n/n_max)^3 + rest x (n/n_max)^2 at n=%d: %.0f s '
                         '(--impl reference runs the longer ladder up to 16384)' % (fit['n'], ['%.2f' % t for t in fit['seconds']], n, sec),
               'fit': fit}

    line = {'metric': 'gp_nlZ_grad_evals_per_sec', 'value': value, 'unit': 'evals/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': ms_per_step, 'higher_is_better': True,
            'scaling': 'weak' if args.replicas else 'strong', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'exact GP covSEard infExact n=%d d=%d (BASELINE configs[1] mlp_kin40k scale-up): '
                                   'nlZ + dnlZ + dnlZ/dX per step' % (n, D),
                       'l2': 'inputs larger than L2 (two %.1f GB matrices per step)' % (n * n * 8 / 1e9),
                       'parallelism': ('one problem sharded over %d GPUs: block-column Cholesky + NCCL panel broadcast, row-sharded '
                                       'L^-T gathered by peer copies under the sweep (P > 2) or pipelined broadcasts (P = 2), '
                                       'tile-sharded B^-1 and gradient pass, all-reduce of the row sums' % world) if sharded
                                      else ('one independent GP per GPU (no collective)' if world > 1 else 'single GPU'),
                       'wall_s_timed_region': wall},
            'roofline': roofline, 'cpu_baseline': cpu,
            'e2e': {'value': e2e_val, 'unit': 'evals/s', 'h2d_bytes_per_step': int(X.nbytes + y.nbytes),
                    'd2h_bytes_per_step': int(n * D * 8 + 8)},
            'kissgp': kiss, 'small_configs': small, 'gpu_launches': int(launches), 'clocks': sampler.summary(),
            'parity_probe': {'nlZ': nlZ, 'nlZ_e2e': nl}, 'multi_gpu_invariance': invariance, 'phases_ms': phases}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
