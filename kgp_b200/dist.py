"""Host-side plumbing of the multi-GPU engine: one process per GPU (torchrun), torch.distributed
only carries the NCCL id and the timing reductions; the data path is NCCL inside libkgp_b200.so.

No reference counterpart (keras-gp runs one Octave process per GP layer, kgp/layers.py:63);
SURVEY.md 8(e) defines the sharding: block columns of the Cholesky, rows of L^-T, tiles of
B^-1 for the exact path; right-hand sides / probes for KISS-GP (`shard_columns`).
"""
import os

import numpy as np

from . import _cabi


def env_rank():
    return int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))


def broadcast_bytes(payload, src=0):
    """Broadcast a bytes object from `src` over the default torch.distributed group (gloo or nccl)."""
    import torch
    import torch.distributed as td
    if not td.is_initialized() or td.get_world_size() == 1:
        return payload
    dev = 'cuda' if td.get_backend() == 'nccl' else 'cpu'
    n = torch.tensor([len(payload) if td.get_rank() == src else 0], dtype=torch.int64, device=dev)
    td.broadcast(n, src)
    buf = torch.zeros(int(n.item()), dtype=torch.uint8, device=dev)
    if td.get_rank() == src:
        buf.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    td.broadcast(buf, src)
    return bytes(buf.cpu().numpy().tobytes())


def attach(gp):
    """Make the engine of `gp` (a B200GP created on this rank's GPU) a member of the job-wide
    communicator.  Every rank must call this; afterwards exact-path evaluations are sharded."""
    import torch.distributed as td
    if not td.is_initialized() or td.get_world_size() == 1:
        return gp
    rank, world = td.get_rank(), td.get_world_size()
    uid = _cabi.nccl_unique_id() if rank == 0 else b''
    uid = broadcast_bytes(uid, 0)
    gp.eng.comm_init(world, rank, uid)
    gp.nranks, gp.rank = world, rank
    return gp


def shard_columns(m, rank, world):
    """Columns (CG right-hand sides / log-det probes) of a batch of m that `rank` solves:
    contiguous, sizes differing by at most one (zero communication inside CG)."""
    base, extra = divmod(int(m), int(world))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def col_owner(block, world):
    return _cabi.lib().kgp_dist_col_owner(int(block), int(world))


def zrow_ranges(n, world):
    """Row ranges [a_r, b_r) of Z = L^-T per rank (equal flops: cube-root split, multiples of 512)."""
    l = _cabi.lib()
    b = [l.kgp_dist_zrow_begin(r, int(world), int(n)) for r in range(int(world) + 1)]
    return [(b[r], b[r + 1]) for r in range(int(world))]


def row_ranges(n, world):
    """[(lo, hi)] of the data rows each rank owns in the row-sharded KISS-GP solve (contiguous shares)."""
    l = _cabi.lib()
    b = [l.kgp_dist_row_begin(r, int(world), int(n)) for r in range(int(world) + 1)]
    return [(b[r], b[r + 1]) for r in range(int(world))]


def rows_sharded(n, N, world):
    """True when a KISS-GP problem of n points on N grid nodes runs the row-sharded CG on `world` ranks."""
    return bool(_cabi.lib().kgp_dist_rows_sharded(int(world), int(n), int(N)))


def allreduce_sum(values):
    """Sum a small float64 vector over ranks (scalar / hyper-gradient terms of sharded probes)."""
    import torch
    import torch.distributed as td
    v = np.asarray(values, dtype=np.float64)
    if not td.is_initialized() or td.get_world_size() == 1:
        return v
    dev = 'cuda' if td.get_backend() == 'nccl' else 'cpu'
    t = torch.from_numpy(v.copy()).to(dev)
    td.all_reduce(t)
    return t.cpu().numpy()
