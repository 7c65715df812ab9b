"""Host-side logic of the multi-GPU path (vega_b200/sharding.py) with world_size 2 over gloo.

The reference's parallel scripts give every MPI rank its own replica and a strided share
of the work with no exchange (bin/run_vega_mc_fits_mpi.py:104-113); here the shards are
contiguous and the per-vector chi2 values are all-gathered.  The evaluator in these tests
is a stand-in closed form (the CUDA engine cannot run on this CPU box): what is covered is
the partition, the ragged / empty shard handling, the gather order and the 1e100 sentinel.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vega_b200 import _lib
from vega_b200.sharding import ShardedChi2, shard_bounds, shard_sizes


def test_shard_bounds_cover_the_batch():
    for n in (0, 1, 2, 7, 1024, 100000):
        for world in (1, 2, 3, 8):
            sizes = shard_sizes(n, world)
            assert sum(sizes) == n and max(sizes) - min(sizes) <= 1
            lo_prev = 0
            for r in range(world):
                lo, hi = shard_bounds(n, world, r)
                assert lo == lo_prev and hi - lo == sizes[r]
                lo_prev = hi
            assert lo_prev == n
    with pytest.raises(ValueError):
        shard_bounds(10, 2, 2)


def _stand_in(theta):
    """Deterministic per-vector 'chi2' and a status word that flags theta[:, 0] < 0."""
    chi2 = theta[:, 1] * 2.0      # exact in FP64: the comparison below is bit-for-bit
    status = torch.where(theta[:, 0] < 0, _lib.STATUS_BOUNDS, 0).to(torch.int32)
    return chi2, status


def _worker(rank, world, port, sizes, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        sharded = ShardedChi2(_stand_in, 'cpu')
        results = {}
        for n in sizes:
            rng = np.random.default_rng(n)
            theta = rng.standard_normal((n, 5))
            if n > 2:
                theta[n // 2, 0] = -abs(theta[n // 2, 0])
            results[n] = sharded.chi2(theta)
            # broadcast from the rank that owns the sampler
            mine = theta if rank == 0 else None
            got = sharded.broadcast_theta(mine, src=0)
            assert np.array_equal(np.asarray(got), theta)
        calls = {'n': 0}

        def counting(t):
            calls['n'] += t.shape[0]
            return _stand_in(t)
        ShardedChi2(counting, 'cpu').chi2(np.ones((7, 5)))
        results['evaluated'] = calls['n']
        out[rank] = results
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


def test_world_size_2_gather_matches_single_process():
    world, sizes = 2, (0, 1, 2, 7, 64)
    manager = mp.Manager()
    out = manager.dict()
    mp.spawn(_worker, args=(world, _free_port(), sizes, out), nprocs=world, join=True)
    assert set(out.keys()) == {0, 1}
    for n in sizes:
        rng = np.random.default_rng(n)
        theta = rng.standard_normal((n, 5))
        if n > 2:
            theta[n // 2, 0] = -abs(theta[n // 2, 0])
        ref = theta[:, 1] * 2.0
        ref[theta[:, 0] < 0] = 1e100       # vega/vega_interface.py:268-279 sentinel
        for rank in range(world):
            assert np.array_equal(out[rank][n], ref), (n, rank)
    # every rank evaluated only its own shard of the 7 vectors
    assert out[0]['evaluated'] == 4 and out[1]['evaluated'] == 3


class _StubFitter:
    """Closed-form stand-in for drivers.MonteCarloFits.fit: mock m has best fit (m, 2m)."""
    names = ['a', 'b']

    def __init__(self):
        self.calls = []

    def result_shape(self, key):
        return (2,) if key in ('theta', 'errors') else ()

    def fit(self, mock_slice=None, **kw):
        ids = np.arange(1000)[mock_slice]
        self.calls.append((ids[0], ids[-1] + 1))
        theta = np.stack([ids * 1.0, ids * 2.0], axis=1)
        return {'theta': theta, 'fval': ids + 0.5, 'errors': theta * 0.1,
                'converged': (ids % 3 != 0)}


def _mc_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    try:
        from vega_b200.sharding import fit_mocks_sharded, gather_rows
        results = {}
        for n in (1, 2, 7):          # n = 1: rank 1 holds an empty shard
            fitter = _StubFitter()
            results[n] = (fit_mocks_sharded(fitter, n, 'cpu'), fitter.calls)
        lo, hi = shard_bounds(5, world, rank)
        results['rows'] = gather_rows(np.arange(5.)[lo:hi, None] * np.ones((1, 3)), 5, 'cpu')
        out[rank] = results
    finally:
        dist.destroy_process_group()


def test_world_size_2_mock_fits_are_sharded_and_gathered():
    """C5 host logic: contiguous shards of the mocks, one padded gather per result array."""
    world = 2
    manager = mp.Manager()
    out = manager.dict()
    mp.spawn(_mc_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    for n in (1, 2, 7):
        ids = np.arange(n)
        for rank in range(world):
            res, calls = out[rank][n]
            assert np.array_equal(res['theta'], np.stack([ids * 1.0, ids * 2.0], axis=1))
            assert np.array_equal(res['fval'], ids + 0.5)
            assert np.array_equal(res['converged'], ids % 3 != 0)
            lo, hi = shard_bounds(n, world, rank)
            assert calls == ([(lo, hi)] if hi > lo else [])
    assert np.array_equal(out[0]['rows'], np.arange(5.)[:, None] * np.ones((1, 3)))


def test_sharded_engine_needs_cuda():
    """No CPU evaluator in the product: building the engine-backed shard evaluator without a
    device fails loudly."""
    if torch.cuda.is_available():
        pytest.skip('CUDA present')
    from vega_b200.utils import VegaDeviceError
    with pytest.raises(VegaDeviceError):
        ShardedChi2.from_vega(object())
