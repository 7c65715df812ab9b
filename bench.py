"""Benchmark of the batched chi2 path (BASELINE.json metric: chi2 evals/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

Workload at N=1: C2 = BASELINE.json configs[1] -- eBOSS-DR16-shaped Lya x Lya (2500 bins,
15 metal correlations) + Lya x QSO (5000 bins, 4 metal correlations) with dense synthetic
distortion / metal / covariance matrices, a batch of 1024 parameter vectors.  A step is one
chi2 pass over the batch.  For N>1 (torchrun) every rank holds a replica of the plan and its
own 1024 vectors (weak scaling); the per-vector chi2 values are all-gathered over NCCL.

One JSON line on stdout (rank 0).  `value` = device-resident throughput (theta already in
HBM, CUDA events); `e2e` = the same metric through VegaInterface.chi2_batch with HOST
buffers (pinned staging, H2D of theta, D2H of chi2 + status inside the timed region).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / 'tests'))

MAIN_INI = 'configs/c2/main.ini'
WORKLOAD = 'C2 eBOSS-DR16-shaped LyaxLya(2500 bins,15 metal corr)+LyaxQSO(5000 bins,4 metal corr), dense DM/metal/cov'
STAGES = ['pk_ell', 'hankel_gemm', 'xi_gather', 'metal_gemm', 'dm_gemm', 'chi2_gemm', 'elementwise']


def build_state(seed=20261104):
    """Host-side init: VegaInterface + dense synthetic operands (no CUDA touched)."""
    from vega_b200 import VegaInterface, synthetic
    vega = VegaInterface(MAIN_INI)
    synthetic.inject_dense_operands(vega, seed=seed)
    return vega


# ----------------------------------------------------------------------------------------
# CPU baseline: the oracle (a numpy port of the reference's own path) on the host cores,
# one replica per core like the reference's rank-per-replica model (bin/run_vega_mpi.py).
# ----------------------------------------------------------------------------------------
_WORKER = {}


def _worker_init(n_warm):
    os.environ['OMP_NUM_THREADS'] = '1'
    try:
        from threadpoolctl import threadpool_limits
        _WORKER['limit'] = threadpool_limits(1)
    except Exception:
        pass
    oracle, names, theta = _WORKER['oracle'], _WORKER['names'], _WORKER['theta']
    for b in range(n_warm):
        oracle.chi2(dict(zip(names, theta[b])))


def _worker_eval(rows):
    oracle, names, theta = _WORKER['oracle'], _WORKER['names'], _WORKER['theta']
    return [oracle.chi2(dict(zip(names, theta[b]))) for b in rows]


def cpu_baseline(vega, theta, per_core=2, steps=1, warmup=1):
    """evals/s of the CPU oracle on all host cores over a bounded sample of theta."""
    import multiprocessing as mp
    from helpers import static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import synthetic
    cores = len(os.sched_getaffinity(0))
    oracle = Oracle(vega, tables=static_tables())
    synthetic.draw_mock_data(vega, oracle.compute_model())
    names = list(vega.sample_params['limits'])
    n = min(theta.shape[0], cores * per_core)
    _WORKER.update(oracle=oracle, names=names, theta=theta)
    ctx = mp.get_context('fork')
    chunks = [list(range(i, n, cores)) for i in range(cores) if i < n]
    times = []
    with ctx.Pool(len(chunks), initializer=_worker_init, initargs=(1,)) as pool:
        pool.map(_worker_eval, [[0]] * len(chunks))          # all workers up and warm
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            pool.map(_worker_eval, chunks, chunksize=1)
            if it >= warmup:
                times.append(time.perf_counter() - t0)
    dt = float(np.mean(times))
    return {'value': n / dt, 'unit': 'evals/s', 'cores': len(chunks), 'kind': 'port',
            'sample': f'{n} of the batch\'s parameter vectors, {len(chunks)} forked replicas of the numpy '
                      f'oracle (1 BLAS thread each), {dt:.2f} s per pass',
            's_per_eval_per_core': dt * len(chunks) / n}, dt


class ClockSampler:
    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.file = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--id={self.gpu}', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                 '-lms', '100'], stdout=self.file, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        self.proc.wait()
        self.file.flush()
        rows = [r.split(', ') for r in Path(self.file.name).read_text().strip().splitlines() if r.strip()]
        os.unlink(self.file.name)
        sm, mx, reasons = [], [], set()
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(names, r[5:9]):
                if val.strip() == 'Active':
                    reasons.add(name)
        if not sm:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['no samples']}
        return {'sm_mhz': float(np.median(sm)), 'sm_max_mhz': float(max(mx)), 'reasons': sorted(reasons),
                'samples': len(sm)}
