"""Sharding of a batch of parameter vectors over one process per GPU.

The chi2 path has no exchange inside an evaluation: every parameter vector (PolyChord live
point, iminuit stencil point, Monte Carlo mock) is independent, exactly as in the
reference's rank-per-replica MPI scripts (bin/run_vega_mpi.py:24-52,
bin/run_vega_mc_fits_mpi.py:81-159, which split mocks with ``range(rank, n, size)`` and
never communicate).  Here every rank holds a replica of the device plan, takes a
contiguous shard of the batch, and the only collective is ONE all-gather of 16 bytes per
vector (chi2 and its status word side by side in one send buffer, which the chi2 kernels write
into directly) -- NCCL over NVLink on the GPU box, gloo in the CPU tests of the host logic.

The evaluator is the device engine (``VegaInterface.chi2_device`` path); the host-side
partition / gather arithmetic is separated from it so that it can be exercised with
world_size 2 on CPU (tests/test_sharding.py) with a stand-in evaluator.  There is no CPU
evaluator in the product: `ShardedChi2` built from a VegaInterface needs CUDA.
"""
import numpy as np


def shard_bounds(n, world, rank):
    """Contiguous shard [lo, hi) of n items for `rank` of `world`; the first n % world
    ranks hold one extra item (ragged batches are allowed, empty shards too)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f'rank {rank} outside world of {world}')
    base, extra = divmod(int(n), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(n, world):
    return [shard_bounds(n, world, r)[1] - shard_bounds(n, world, r)[0] for r in range(world)]


class ShardedChi2:
    """chi2 of a global batch theta[B, P], evaluated shard-per-rank and all-gathered.

    evaluate(theta_local: torch.Tensor[b, P]) -> (chi2[b] float64, status[b] int32) tensors
    on `device`; with `writes_into` it is called as evaluate(theta_local, chi2_out) and leaves the
    chi2 values in chi2_out (a slice of the send buffer of the collective: no staging copy).
    `group` is a torch.distributed process group (None = default)."""

    def __init__(self, evaluate, device, group=None, writes_into=False):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.evaluate = evaluate
        self.writes_into = writes_into
        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._gather_buf = None

    @classmethod
    def from_vega(cls, vega, names=(), group=None):
        """Shard evaluator backed by the device engine of a VegaInterface replica."""
        import torch
        if not torch.cuda.is_available():
            from ._lib import VegaDeviceError
            raise VegaDeviceError('ShardedChi2 needs a CUDA device: the B200 path has no CPU evaluator.')
        engine = vega._engine_for(tuple(names))
        return cls(lambda t, out=None: engine.chi2_device(t, out), engine.device, group, writes_into=True)

    def _buffers(self, cap):
        """(gathered [world, 2, cap], send [2, cap]) float64: row 0 chi2, row 1 the status word (small
        integers are exact in a double), so ONE collective moves both."""
        torch = self.torch
        if self._gather_buf is None or self._gather_buf[1].shape[1] < cap:
            self._gather_buf = (torch.empty(self.world, 2, cap, dtype=torch.float64, device=self.device),
                                torch.zeros(2, cap, dtype=torch.float64, device=self.device))
        return self._gather_buf

    def local_shard(self, theta):
        lo, hi = shard_bounds(theta.shape[0], self.world, self.rank)
        return theta[lo:hi]

    def chi2_local(self, theta_local, chi2_out=None):
        """(chi2, status) device tensors of this rank's shard (already on the device)."""
        torch = self.torch
        if theta_local.shape[0] == 0:
            return (torch.empty(0, dtype=torch.float64, device=self.device),
                    torch.empty(0, dtype=torch.int32, device=self.device))
        if chi2_out is not None and self.writes_into:
            return self.evaluate(theta_local, chi2_out)
        return self.evaluate(theta_local)

    def gather(self, chi2_local, status_local, n_total):
        """All-gather the per-shard results into global order.  Shards are padded to the
        largest shard so one fixed-size collective serves ragged batches."""
        torch, dist = self.torch, self.dist
        if self.world == 1:
            return chi2_local, status_local
        sizes = shard_sizes(n_total, self.world)
        cap = max(max(sizes), 1)
        gathered, send = self._buffers(cap)
        cap = send.shape[1]
        n = chi2_local.shape[0]
        if n and chi2_local.data_ptr() != send[0].data_ptr():     # (already there when evaluate wrote into it)
            send[0, :n].copy_(chi2_local)
        send[1, :n].copy_(status_local)
        dist.all_gather_into_tensor(gathered.view(-1), send.view(-1), group=self.group)
        chi2 = torch.cat([gathered[r, 0, :s] for r, s in enumerate(sizes)])
        stat = torch.cat([gathered[r, 1, :s] for r, s in enumerate(sizes)]).to(torch.int32)
        return chi2, stat

    def send_slot(self, n_local, n_total):
        """The slice of the send buffer an evaluator can write this rank's n_local chi2 values into."""
        if self.world == 1 or n_local == 0:
            return None
        cap = max(max(shard_sizes(n_total, self.world)), 1)
        return self._buffers(cap)[1][0, :n_local]

    def chi2(self, theta):
        """theta: the GLOBAL batch (host ndarray or tensor, identical on every rank, e.g. the
        sampler's live points after a broadcast).  Returns host ndarray chi2[B] with the
        reference's 1e100 sentinel for model-domain failures, identical on every rank."""
        from ._lib import STATUS_ARINYO, STATUS_BOUNDS
        torch = self.torch
        if not torch.is_tensor(theta):
            theta = torch.from_numpy(np.ascontiguousarray(np.atleast_2d(theta), dtype=np.float64))
        n_total = theta.shape[0]
        local = self.local_shard(theta).contiguous().to(self.device)
        c, s = self.chi2_local(local, self.send_slot(local.shape[0], n_total))
        c, s = self.gather(c, s, n_total)
        c = c.cpu().numpy().copy()
        s = s.cpu().numpy()
        c[(s & (STATUS_ARINYO | STATUS_BOUNDS)) != 0] = 1e100
        return c

    def broadcast_theta(self, theta, src=0):
        """Broadcast a batch from the rank that owns the sampler."""
        torch, dist = self.torch, self.dist
        if self.world == 1:
            return theta
        shape = torch.zeros(2, dtype=torch.int64, device=self.device)
        if self.rank == src:
            theta = torch.as_tensor(np.asarray(theta, dtype=np.float64))
            shape[0], shape[1] = theta.shape
        dist.broadcast(shape, src, group=self.group)
        buf = torch.empty(int(shape[0]), int(shape[1]), dtype=torch.float64, device=self.device)
        if self.rank == src:
            buf.copy_(theta)
        dist.broadcast(buf, src, group=self.group)
        return buf


def gather_rows(local, n_total, device, group=None):
    """All-gather of per-item result rows: rank r holds the rows of its contiguous shard
    (shard_bounds(n_total, world, r)) as a float64 array [n_r, ...]; every rank gets the
    [n_total, ...] array in item order.  One padded fixed-size collective."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    local = np.ascontiguousarray(local, dtype=np.float64)
    if world == 1:
        return local
    sizes = shard_sizes(n_total, world)
    tail = local.shape[1:]
    width = int(np.prod(tail)) if tail else 1
    cap = max(sizes)
    send = torch.zeros(cap * width, dtype=torch.float64, device=device)
    if local.size:
        send[:local.size] = torch.from_numpy(local.reshape(-1)).to(device)
    recv = torch.empty(world * cap * width, dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(recv, send, group=group)
    recv = recv.cpu().numpy().reshape(world, cap, width)
    out = np.concatenate([recv[r, :sizes[r]] for r in range(world)], axis=0)
    return out.reshape((n_total,) + tail)


def fit_mocks_sharded(fitter, n_mocks, device, keys=('theta', 'fval', 'errors', 'converged'), group=None,
                      **fit_kwargs):
    """Monte Carlo mock fits sharded over the ranks (bin/run_vega_mc_fits_mpi.py:104-113 gives
    rank r the mocks range(r, n, size) and writes one file per rank; here the shards are
    contiguous and the best fits are all-gathered).  `fitter.fit(mock_slice=slice(lo, hi), ...)`
    returns a dict of arrays with one row per mock of the slice (drivers.MonteCarloFits.fit)."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_bounds(n_mocks, world, rank)
    res = fitter.fit(mock_slice=slice(lo, hi), **fit_kwargs) if hi > lo else None
    out = {}
    for key in keys:
        if res is not None:
            local = np.asarray(res[key], dtype=np.float64)
        else:          # an empty shard still takes part in the collective
            local = np.zeros((0,) + tuple(fitter.result_shape(key)), dtype=np.float64)
        out[key] = gather_rows(local, n_mocks, device, group)
    if 'converged' in out:
        out['converged'] = out['converged'] > 0.5
    return out
