"""Host-side logic of the multi-GPU engine on CPU: ownership maps exported by the C ABI, column sharding of
right-hand sides / probes, and the torch.distributed plumbing (gloo, world_size 2) that carries the NCCL id
and the scalar reductions.  The NCCL data path itself needs GPUs (scripts/dist_exact.py, tests -m gpu)."""
import os
import socket

import numpy as np
import pytest


def test_block_column_owner_is_cyclic(built_lib):
    from kgp_b200 import dist
    for world in (1, 2, 4, 8):
        owners = [dist.col_owner(k, world) for k in range(79)]
        assert owners == [k % world for k in range(79)]
        counts = np.bincount(owners, minlength=world)
        assert counts.max() - counts.min() <= 1


@pytest.mark.parametrize('n', [3000, 8192, 40000])
@pytest.mark.parametrize('world', [1, 2, 4, 8])
def test_zrow_ranges_cover_and_balance(built_lib, n, world):
    from kgp_b200 import dist
    np_ = ((n + 127) // 128) * 128
    r = dist.zrow_ranges(n, world)
    assert r[0][0] == 0 and r[-1][1] == np_
    for (a, b), (a2, b2) in zip(r[:-1], r[1:]):
        assert b == a2 and a <= b
    for a, b in r[:-1]:
        assert b % 512 == 0                      # block aligned (NB = 512)
    # flops of row i of Z = L^-T are ~ (n-i)^2: the cube-root split equalises them up to block rounding
    if n >= 40000:
        i = np.arange(np_, dtype=np.float64)
        w = (np_ - i) ** 2
        share = np.array([w[a:b].sum() for a, b in r]) / w.sum()
        assert share.max() < 1.25 / world


def test_shard_columns_partitions_every_batch(built_lib):
    from kgp_b200 import dist
    for m in (1, 7, 10, 11, 64):
        for world in (1, 2, 3, 8):
            got = []
            for rk in range(world):
                lo, hi = dist.shard_columns(m, rk, world)
                got += list(range(lo, hi))
            assert got == list(range(m))
            sizes = [dist.shard_columns(m, rk, world)[1] - dist.shard_columns(m, rk, world)[0] for rk in range(world)]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as td
    from kgp_b200 import dist
    td.init_process_group('gloo', rank=rank, world_size=world)
    payload = bytes(range(128)) if rank == 0 else b''
    got = dist.broadcast_bytes(payload, 0)
    lo, hi = dist.shard_columns(11, rank, world)
    part = np.zeros(3); part[0] = hi - lo; part[1] = sum(range(lo, hi)); part[2] = 1.0
    tot = dist.allreduce_sum(part)
    q.put((rank, got == bytes(range(128)), tot.tolist()))
    td.destroy_process_group()


def test_gloo_world2_id_broadcast_and_probe_reduction(built_lib):
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = [q.get(timeout=120) for _ in ps]
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok, tot in res:
        assert ok                                  # the 128-byte id arrived intact on every rank
        assert tot == [11.0, float(sum(range(11))), 2.0]


@pytest.mark.parametrize('n,world', [(20000, 2), (1000003, 8), (40001, 3)])
def test_row_ranges_partition_the_points(built_lib, n, world):
    """Row-sharded KISS-GP solve (SURVEY 8(e) KISS-GP (ii)): contiguous shares that tile [0, n), sizes within one."""
    from kgp_b200 import dist
    r = dist.row_ranges(n, world)
    assert r[0][0] == 0 and r[-1][1] == n
    assert all(r[i][1] == r[i + 1][0] for i in range(world - 1))
    sizes = [hi - lo for lo, hi in r]
    assert max(sizes) - min(sizes) <= 1


def test_row_sharding_threshold(built_lib):
    from kgp_b200 import dist
    assert not dist.rows_sharded(6000, 1600, 2)          # < 4096 points per rank: the exchange would cost more than it saves
    assert dist.rows_sharded(20000, 1600, 2) and dist.rows_sharded(8_000_000, 4900, 8)
    assert not dist.rows_sharded(10**6, 40000, 2)        # grid larger than the peer window
    assert not dist.rows_sharded(10**6, 4900, 1)
