"""Benchmark of the batched chi2 path (BASELINE.json metric: chi2 evals/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl reference]

Workload at N=1: C2 = BASELINE.json configs[1] -- eBOSS-DR16-shaped Lya x Lya (2500 bins,
15 metal correlations) + Lya x QSO (5000 bins, 4 metal correlations) with dense synthetic
distortion / metal / covariance matrices, a batch of 1024 parameter vectors.  A step is one
chi2 pass over the batch.  For N>1 (torchrun) every rank holds a replica of the plan and its
own 1024 vectors (weak scaling); the per-vector chi2 values are all-gathered over NCCL.

One JSON line on stdout (rank 0).  `value` = device-resident throughput (theta already in
HBM, CUDA events); `e2e` = the same metric through VegaInterface.chi2_batch with HOST
buffers (pinned staging, H2D of theta, D2H of chi2 + status inside the timed region; at N > 1
also the gather).  `parity`: the GPU values against the CPU oracle on rows of the very batch that
is timed.  `roofline` / `roofline_kernels`: every kernel of the step against its bound.  After the
headline, in the same process: `c3` (DESI-shaped model: 1 / 2P+1 / P^2+P+1 vectors and one
complete 15-parameter fit, checked on the oracle), `c4` (1e5 vectors sharded over the ranks:
strong scaling) and `c5` (lock-step mock fits, mocks sharded over the ranks).
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
# tags of vg_profile_read (include/vega_b200.h)
STAGES = ['pk_ell', 'hankel_gemm', 'xi_gather', 'metal_gemm', 'dm_gemm', 'chi2_gemm', 'elementwise',
          'moment_tables', 'moment_xi', 'ozaki_slice', 'metal_rp', 'velo_tables']


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
    """chi2 of the listed rows of the batch; for the rows in _WORKER['xi_rows'] also the model xi
    (the parity block compares both with what the GPU computed for the same rows)."""
    oracle, names, theta = _WORKER['oracle'], _WORKER['names'], _WORKER['theta']
    out = []
    for b in rows:
        pars = dict(zip(names, theta[b]))
        xi = oracle.compute_model(pars) if b in _WORKER.get('xi_rows', ()) else None
        out.append((b, oracle.chi2(pars), xi))
    return out


def cpu_baseline(vega, theta, per_core=2, steps=1, warmup=1, xi_rows=(0, 1, 2, 3)):
    """evals/s of the CPU oracle on all host cores over a bounded sample of theta, plus the values
    it computed ({row: (chi2, xi or None)}): the bench line's parity block is built from them."""
    import multiprocessing as mp
    from helpers import static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import synthetic
    cores = len(os.sched_getaffinity(0))
    oracle = Oracle(vega, tables=static_tables())
    synthetic.draw_mock_data(vega, oracle.compute_model())
    names = list(vega.sample_params['limits'])
    n = min(theta.shape[0], cores * per_core)
    _WORKER.update(oracle=oracle, names=names, theta=theta, xi_rows=())
    ctx = mp.get_context('fork')
    chunks = [list(range(i, n, cores)) for i in range(cores) if i < n]
    times, values = [], {}
    with ctx.Pool(len(chunks), initializer=_worker_init, initargs=(1,)) as pool:
        pool.map(_worker_eval, [[0]] * len(chunks))          # all workers up and warm
        for it in range(warmup + steps):
            t0 = time.perf_counter()
            res = pool.map(_worker_eval, chunks, chunksize=1)
            if it >= warmup:
                times.append(time.perf_counter() - t0)
        for part in res:
            for b, chi2, _ in part:
                values[b] = [chi2, None]
    if xi_rows:       # outside the timed passes: the model xi of a few rows, one per worker
        _WORKER['xi_rows'] = tuple(xi_rows)
        with ctx.Pool(min(len(xi_rows), cores), initializer=_worker_init, initargs=(0,)) as pool:
            for part in pool.map(_worker_eval, [[b] for b in xi_rows], chunksize=1):
                for b, chi2, xi in part:
                    values[b] = [chi2, xi]
    dt = float(np.mean(times))
    return {'value': n / dt, 'unit': 'evals/s', 'cores': len(chunks), 'kind': 'port',
            'sample': f'{n} of the batch\'s parameter vectors, {len(chunks)} forked replicas of the numpy '
                      f'oracle (1 BLAS thread each), {dt:.2f} s per pass',
            's_per_eval_per_core': dt * len(chunks) / n}, dt, values


class ClockSampler:
    QUERY = ('index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,'
             'clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
             'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.file = tempfile.NamedTemporaryFile('w+', suffix='.csv', delete=False)
        self.proc = None
        self.first = 0

    def _lines(self):
        try:
            return len([r for r in Path(self.file.name).read_text().splitlines() if r.strip()])
        except OSError:
            return 0

    def mark(self, wait_s=3.0):
        """Samples from here on count (nvidia-smi takes a few hundred ms to deliver its first line: wait
        for it, so that the load that follows is sampled from its start)."""
        if self.proc is None:
            return
        t0 = time.perf_counter()
        while self._lines() == 0 and time.perf_counter() - t0 < wait_s:
            time.sleep(0.02)
        self.first = self._lines()

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ['nvidia-smi', f'--id={self.gpu}', f'--query-gpu={self.QUERY}', '--format=csv,noheader,nounits',
                 '-lms', '50'], stdout=self.file, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        self.proc.terminate()
        self.proc.wait()
        self.file.flush()
        rows = [r.split(', ') for r in Path(self.file.name).read_text().strip().splitlines() if r.strip()][self.first:]
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


def gemm_flops(engine, B, velo_tables=True):
    """Algorithmic FP64 flops of the contraction stages for B vectors:
    Hankel operator 2*n_ell*nk^2 per evaluation, metal matrices 2*Nm^2 each (dense) or 2*Np^2*Nt
    (rp-only new_metals form), distortion matrix 2*Nmask*Nm (masked rows), chi2 2*Nmask^2 (or one
    2*(sum Nmask)^2 under a global covariance)."""
    out = {'hankel_gemm': 0.0, 'metal_gemm': 0.0, 'metal_rp': 0.0, 'dm_gemm': 0.0, 'chi2_gemm': 0.0, 'velo_tables': 0.0}
    nk, n_ell = 814, 4
    total_mask = 0
    for ci, info in enumerate(engine.infos):
        item, data = info['item'], info['data']
        nm = item.model_coordinates.r_grid.size
        nmask = int(data.model_mask.sum())
        total_mask += nmask
        own = [d for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct]
        n_vt, orders, _ = engine.velo_tables(ci) if velo_tables else (0, 0, 0.0)
        n_eval = len(own) - engine.n_moment(ci) - n_vt      # evaluations that reach the Hankel GEMM
        if n_vt:
            # velocity-dispersion tables: one GEMM [B x orders*nm] x [orders*nm x 4*2*816] per evaluation,
            # nm = 3 x (1 + tracers with an HCD term) moment coefficients (csrc/vg_moments.cuh)
            vt = [d for d in own if d.dyn_mask == 64][:n_vt]
            for d in vt:
                n_hcd = (2 * bool(d.hcd_on[0]) if d.same_tracer else bool(d.hcd_on[0]) + bool(d.hcd_on[1])) if d.hcd_model else 0
                out['velo_tables'] += 2.0 * orders * 3 * (1 + n_hcd) * n_ell * 2 * 816 * B
        out['hankel_gemm'] += 2.0 * n_ell * nk * nk * n_eval * B
        # metal matrices applied per vector (fast_metals freezes the metal x metal ones)
        for mat, d in zip(info['metal_mats'], [d for d in info['evals'] if d.out_slot >= 2 and not d.direct]):
            if mat is None or d.frozen:
                continue
            if hasattr(mat, 'rp_mat'):
                out['metal_rp'] += 2.0 * mat.rp_mat.shape[0]**2 * mat.rt_nbins * B
            else:
                out['metal_gemm'] += 2.0 * nm * nm * B
        if not data.distortion_is_identity:
            out['dm_gemm'] += 2.0 * nmask * nm * B
        if not data.cov_is_identity and not getattr(engine.builder.vega, '_use_global_cov', False):
            out['chi2_gemm'] += 2.0 * nmask * nmask * B
    if getattr(engine.builder.vega, '_use_global_cov', False):
        out['chi2_gemm'] = 2.0 * total_mask * total_mask * B
    return out


def hbm_bytes(engine, B):
    """Algorithmic HBM bytes per step of the bandwidth-bound kernels: a knot table is
    4 multipoles x (values + second derivatives) x 816 knots x 8 B = 52 224 B per (vector, evaluation);
    a xi row is 8*Nm B."""
    table = 4 * 2 * 816 * 8.0
    out = {'xi_gather': 0.0, 'moment_xi': 0.0, 'moment_tables': 0.0, 'elementwise': 0.0, 'ozaki_slice': 0.0}
    for ci, info in enumerate(engine.infos):
        item, data = info['item'], info['data']
        nm = item.model_coordinates.r_grid.size
        nmask = int(data.model_mask.sum())
        hot = [d for d in info['evals'] if not d.frozen and not d.direct]
        n_static = engine.n_static_xi(ci)
        n_gather = len(hot) - n_static
        out['xi_gather'] += n_gather * B * (table + 8.0 * nm)       # table read, xi row written
        out['moment_xi'] += n_static * B * 8.0 * nm                  # xi row written (static tables stay in L2)
        out['moment_tables'] += engine.n_moment(ci) * B * table      # knot table written
        # combine (2 cores + metals read, x written) and post (y read, residual written) + theta expansion
        out['elementwise'] += B * (4 * 8.0 * nm + 2 * 8.0 * nmask)
        digits, _ = engine.ozaki_slices(ci)
        if digits:
            n_metal = sum(1 for d in hot if d.out_slot >= 2)
            out['ozaki_slice'] += B * n_metal * nm * (2 * 8.0 + digits)   # two reads of xmet, S digit planes written
        dm_digits, _ = engine.ozaki_dm_slices(ci)
        if dm_digits:
            out['ozaki_slice'] += B * nm * (2 * 8.0 + dm_digits)          # the same for the pre-distortion model
    return out


def fp64_peaks(lib, torch):
    """Measured FP64 denominators (MEASURED_PEAKS.json carries only HBM and bf16):
    cuBLAS DGEMM 8192^3 (tensor/DMMA) and an FMA-chain probe (vector pipe)."""
    a = torch.randn(8192, 8192, dtype=torch.float64, device='cuda')
    b = torch.randn(8192, 8192, dtype=torch.float64, device='cuda')
    best = 1e9
    for it in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        if it:
            best = min(best, e0.elapsed_time(e1))
    dgemm = 2 * 8192.0**3 / (best * 1e-3) / 1e12
    del a, b
    tf = C.c_double()
    scratch = torch.zeros(148 * 8 * 256, dtype=torch.float64, device='cuda')
    lib.vg_probe_dfma(148 * 8, 20000, C.c_void_p(scratch.data_ptr()), C.byref(tf), None)
    return dgemm, tf.value


def measured_peaks():
    path = REPO / 'MEASURED_PEAKS.json'
    if path.exists():
        d = json.loads(path.read_text())
        return {'hbm_gbs': d['hbm_gbs'], 'bf16_tflops': d['bf16_tflops'],
                'bf16_tflops_sustained': d.get('bf16_tflops_sustained', d['bf16_tflops']), 'source': 'MEASURED_PEAKS.json'}
    return {'hbm_gbs': 6650.0, 'bf16_tflops': 1590.0, 'bf16_tflops_sustained': 1400.0,
            'source': 'fallback of B200_PROFILING.md (MEASURED_PEAKS.json absent)'}


def run_reference_arm(args, rank, world, out=sys.stdout):
    if rank != 0:
        return
    vega = build_state()
    from vega_b200 import synthetic
    theta = synthetic.draw_theta(vega, args.batch)
    base, dt, _ = cpu_baseline(vega, theta, per_core=1, steps=args.steps, warmup=max(args.warmup, 1), xi_rows=())
    line = {'impl': 'reference', 'metric': 'chi2 evals/sec', 'value': base['value'], 'unit': 'evals/s',
            'n_gpus': args.gpus, 'steps': args.steps, 'warmup': args.warmup, 'ms_per_step': dt * 1e3,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic', 'config': {'workload': WORKLOAD, 'batch_per_gpu': args.batch},
            'cpu_baseline': base,
            'e2e': {'value': base['value'], 'unit': 'evals/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
            'note': 'the reference is pure Python with no compiled sources; its path is timed through the '
                    'numpy oracle (oracle/vega_oracle.py), one forked replica per host core'}
    print(json.dumps(line), file=out, flush=True)


def _claim_stdout():
    """stdout carries exactly ONE JSON line: libraries that print there (NCCL's version banner,
    torchrun notices) are sent to stderr; the returned file writes to the real stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), 'w')
    os.dup2(2, 1)
    return real


class Dist:
    """The few collectives the bench needs, no-ops at world size 1."""

    def __init__(self, torch, world):
        self.torch, self.world = torch, world
        if world > 1:
            import torch.distributed as dist
            self.dist = dist

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def gather_floats(self, values):
        """[[values of rank 0], [values of rank 1], ...] on every rank."""
        t = self.torch.tensor(values, dtype=self.torch.float64, device='cuda')
        if self.world == 1:
            return [t.cpu().tolist()]
        out = self.torch.empty(self.world * len(values), dtype=self.torch.float64, device='cuda')
        self.dist.all_gather_into_tensor(out, t)
        return out.view(self.world, -1).cpu().tolist()

    def max(self, value):
        t = self.torch.tensor([value], dtype=self.torch.float64, device='cuda')
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def timed_device(torch, dd, step, steps):
    """ms per call of `step` over `steps` calls: CUDA events on the launch stream, barrier +
    synchronize on both sides, max over ranks."""
    dd.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    dd.barrier()
    return dd.max(e0.elapsed_time(e1)) / steps


def parity_block(vega, theta, names, oracle_values):
    """GPU values against what the CPU oracle computed for the SAME rows of the +-3 sigma batch in
    the cpu_baseline leg: chi2 of every sampled row, model xi of a few.  Tolerances of the path
    (BASELINE.json): xi to 1e-10 * max|xi| per correlation; chi2 to 1e-6 absolute where chi2 is of
    the order of the degrees of freedom, i.e. 1e-10 relative on these 1e11-1e14 values."""
    rows = sorted(oracle_values)
    got = vega.chi2_batch(theta[rows], names)
    ref = np.array([oracle_values[b][0] for b in rows])
    rel = np.abs(got - ref) / np.maximum(np.abs(ref), 1.0)
    xi_rows = [b for b in rows if oracle_values[b][1] is not None]
    dxi = 0.0
    if xi_rows:
        models, _ = vega.compute_model_batch(theta[xi_rows], names)
        for i, b in enumerate(xi_rows):
            for name, ref_xi in oracle_values[b][1].items():
                dxi = max(dxi, float(np.max(np.abs(models[name][i] - ref_xi)) / np.max(np.abs(ref_xi))))
    block = {'rows': len(rows), 'xi_rows': len(xi_rows), 'max_rel_dchi2': float(rel.max()),
             'max_dxi_over_maxxi': dxi, 'chi2_range': [float(ref.min()), float(ref.max())],
             'tolerance': {'xi': 1e-10, 'chi2_rel': 1e-10},
             'oracle': 'oracle/vega_oracle.py on the rows of this batch the cpu_baseline leg evaluated'}
    block['ok'] = bool(block['max_rel_dchi2'] <= 1e-10 and dxi <= 1e-10)
    return block


# ----------------------------------------------------------------------------------------------
# C3 / C4 / C5 (BASELINE.json configs[2..4]) on the DESI-DR2-shaped model, after the C2 headline
# ----------------------------------------------------------------------------------------------
C3_NAMES = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'drp_QSO',
            'sigma_velo_disp_lorentz_QSO', 'bao_amp', 'bias_eta_SiII(1260)', 'bias_eta_SiIII(1207)',
            'bias_eta_SiII(1193)', 'bias_eta_SiII(1190)', 'bias_eta_CIV(eff)', 'bias_QSO']


def build_c3(rank):
    """DESI-DR2-shaped workload: 4 correlations, model grids of 5625 / 11250 bins out
    to 300 Mpc/h, rp-only new_metals matrices, fast_metals, one global covariance over the 9540
    masked bins; 15 sampled parameters.  Returns (vega, names, init seconds split)."""
    from helpers import make_c3_config
    from vega_b200 import VegaInterface, synthetic
    t0 = time.perf_counter()
    vega = VegaInterface(make_c3_config(f'/tmp/bench_c3_{os.getpid()}_{rank}'))
    t1 = time.perf_counter()
    synthetic.inject_dense_operands(vega, with_distortion=False, with_metals=False)
    synthetic.inject_global_covariance(vega)
    t2 = time.perf_counter()
    from vega_b200.parameters import get_default_values
    defaults = get_default_values()
    names = [n for n in C3_NAMES if n in vega.params]
    sp = vega.sample_params
    for n in names:
        if n not in sp['limits']:      # what `name = True` in [sample] gives (vega/vega_interface.py:737-816)
            sp['values'][n] = vega.params[n]
            sp['errors'][n] = 0.02 * abs(vega.params[n]) + 1e-4
            sp['limits'][n] = defaults[n]['limits'] if n in defaults else (None, None)
            sp['fix'][n] = False
    for key in ('values', 'errors', 'limits', 'fix'):
        sp[key] = {n: sp[key][n] for n in names}
    t3 = time.perf_counter()
    vega._engine_for()
    t4 = time.perf_counter()
    return vega, names, {'ini_fits_new_metals_s': t1 - t0, 'synthetic_covariance_s': t2 - t1,
                         'plan_build_finalize_s': t4 - t3}


def run_c3(vega, names, torch, dd, rank, lib):
    """Batches an iminuit-style fit consumes (1 vector, a gradient stencil, a Hessian stencil), then
    ONE complete fit of all 15 parameters, checked on the CPU oracle: chi2 at the best fit, and
    Delta chi2 at +-1 sigma along every axis (1 for a parabola)."""
    from vega_b200 import synthetic
    from vega_b200.drivers import stencil
    out = {}
    fiducial = vega.compute_model()
    synthetic.draw_mock_data(vega, fiducial)
    if rank != 0:
        return out, fiducial
    engine = vega._engine_for()
    base = np.array([vega.params[n] for n in names])
    steps = 0.02 * np.abs(base) + 1e-4
    for label, theta in (('single', base[None]), ('gradient_stencil', stencil(base, steps)),
                         ('hessian_stencil', stencil(base, steps, hessian=True))):
        tt = engine.theta_tensor(theta)
        chi2 = torch.empty(theta.shape[0], dtype=torch.float64, device='cuda')
        status = torch.empty(theta.shape[0], dtype=torch.int32, device='cuda')
        def timed(n=10):
            for _ in range(4):
                engine.chi2_device(tt, chi2, status)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                engine.chi2_device(tt, chi2, status)
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / n
        replays0 = lib.vg_plan_graph_replays(engine.handle)
        ms = timed()                                   # batches of <= 96 vectors replay a CUDA graph
        graphed = lib.vg_plan_graph_replays(engine.handle) > replays0
        lib.vg_profile_enable(engine.handle, 1)        # per-stage events: ordinary launches
        ms_launches = timed()
        ms_stage, cnt_stage = (C.c_double * 16)(), (C.c_int64 * 16)()
        lib.vg_profile_read(engine.handle, ms_stage, cnt_stage)
        lib.vg_profile_enable(engine.handle, 0)
        out[label] = {'vectors': int(theta.shape[0]), 'ms_per_pass': ms, 'evals_per_s': theta.shape[0] / (ms * 1e-3),
                      'cuda_graph': bool(graphed), 'ms_per_pass_ordinary_launches': ms_launches,
                      'stages_ms': {s: round(ms_stage[i] / 14, 4) for i, s in enumerate(STAGES) if ms_stage[i] > 0}}
    t0 = time.perf_counter()
    vega.minimize()
    mini = vega.minimizer
    best = np.array([mini.values[n] for n in names])
    errs = np.array([mini.errors[n] for n in names])
    out['fit'] = {'parameters': len(names), 'wall_s': time.perf_counter() - t0, 'chi2_evaluations': int(mini.n_eval),
                  'fval': float(mini.fmin.fval), 'edm': float(mini.fmin.edm), 'valid': bool(mini.fmin.is_valid),
                  'method': mini.method}
    t0 = time.perf_counter()           # the same fit again: no workspace growth, no graph capture left to pay
    vega.minimize()
    out['fit']['wall_s_repeated'] = time.perf_counter() - t0
    out['fit']['fval_repeated'] = float(vega.minimizer.fmin.fval)
    mini = vega.minimizer
    # oracle check of the minimum: bounded CPU work (2P + 1 evaluations)
    from helpers import static_tables
    from oracle.vega_oracle import Oracle
    t0 = time.perf_counter()
    oracle = Oracle(vega, tables=static_tables(), fast_metals=True)
    ref_best = oracle.chi2(dict(zip(names, best)))
    delta = []
    for j in range(len(names)):
        for sign in (1.0, -1.0):
            p = best.copy()
            p[j] += sign * errs[j]
            delta.append(oracle.chi2(dict(zip(names, p))) - ref_best)
    delta = np.array(delta).reshape(len(names), 2)
    out['fit']['oracle_check'] = {
        'chi2_gpu_minus_oracle_at_bestfit': float(mini.fmin.fval - ref_best),
        'delta_chi2_at_plus_minus_1sigma': {n: [round(float(a), 4), round(float(b), 4)] for n, (a, b) in zip(names, delta)},
        'min_delta_chi2': float(delta.min()), 'cpu_s': time.perf_counter() - t0,
        'meaning': 'oracle chi2 at bestfit +- 1 sigma on each axis minus the oracle chi2 at the best fit: all '
                   'positive = a minimum on the oracle too; for a parabola the value on axis i is C_ii (C^-1)_ii: 1 for an '
                   'uncorrelated parameter (ap, at, bao_amp), larger along the bias / beta / HCD degeneracies'}
    return out, fiducial


def run_c4(vega, torch, dd, world, n_vectors=100000):
    """PolyChord-style batched chi2: n_vectors vectors uniform in +-3 sigma, STRONG scaling -- the
    batch is sharded over the ranks (host theta in, chi2 all-gathered, host out)."""
    from vega_b200 import synthetic
    from vega_b200.sharding import ShardedChi2, shard_sizes
    theta = synthetic.draw_theta(vega, n_vectors, seed=20261105)
    sharded = ShardedChi2.from_vega(vega)
    sharded.chi2(theta[:2048 * world])
    dd.barrier()
    t0 = time.perf_counter()
    chi2 = sharded.chi2(theta)
    torch.cuda.synchronize()
    dt = dd.max(time.perf_counter() - t0)
    per_rank = max(shard_sizes(n_vectors, world))
    chunk = 1024
    full, tail = divmod(per_rank, chunk)
    return {'vectors': n_vectors, 'scaling': 'strong', 'wall_s': dt, 'evals_per_s': n_vectors / dt,
            'vectors_per_rank': per_rank, 'chunks_per_rank': full + (1 if tail else 0), 'ragged_last_chunk': tail,
            'checksum': float(np.sum(chi2[chi2 < 1e99])), 'n_sentinel': int(np.sum(chi2 >= 1e99)),
            'path': 'ShardedChi2.chi2(theta_host): H2D of the shard, chi2 kernels, NCCL all-gather, D2H'}


def run_c5(vega, fiducial, torch, dd, world, n_mocks):
    """Monte Carlo mock fits in lock-step, mocks sharded over the ranks (every rank draws only its
    share from the common random stream)."""
    from vega_b200.drivers import MonteCarloFits
    mc = MonteCarloFits(vega, fiducial_model=fiducial)
    dd.barrier()
    t0 = time.perf_counter()
    mc.draw_sharded(n_mocks, seed=1)
    torch.cuda.synchronize()
    t_draw = dd.max(time.perf_counter() - t0)
    dd.barrier()
    t0 = time.perf_counter()
    res = mc.fit_sharded(max_iter=40, keys=('theta', 'fval', 'errors', 'converged', 'n_iter'))
    torch.cuda.synchronize()
    local_s = time.perf_counter() - t0
    dt = dd.max(local_s)
    per_rank = dd.gather_floats([local_s, float(getattr(mc, 'last_n_eval', 0))])
    vega.monte_carlo = False
    theta = res['theta']
    ndof = int(sum(int(d.data_mask.sum()) for d in vega.data.values())) - len(mc.names)
    ok = np.asarray(res['converged'], dtype=bool) & np.all(res['errors'] > 0, axis=1)   # pulls: converged fits only
    if not ok.any():
        ok[:] = True
    return {'mocks': n_mocks, 'parameters': len(mc.names), 'scaling': 'strong', 'draw_s': t_draw, 'fit_wall_s': dt,
            'fits_per_s': n_mocks / dt, 'converged': int(np.sum(res['converged'])),
            'per_rank_fit_s': [r[0] for r in per_rank], 'per_rank_chi2_evaluations': [int(r[1]) for r in per_rank],
            'mean_iterations': float(np.mean(res['n_iter'])), 'mean_fval': float(np.mean(res['fval'])),
            'degrees_of_freedom': ndof,
            'mean_pull_of_bestfit': {n: float(np.mean((theta[ok, j] - vega.params[n]) / np.maximum(res['errors'][ok, j], 1e-300)))
                                     for j, n in enumerate(mc.names)},
            'mocks_in_the_mean_pull': int(ok.sum()),
            'method': 'quasi-Newton lock-step: gradient stencils (2P+1), start Hessian shared by the mocks, one '
                      'finishing Hessian stencil (P^2+P+1) per mock'}


def main():
    real_stdout = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--batch', type=int, default=1024, help='parameter vectors per GPU per step')
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-extra', action='store_true', help='skip the C3 / C4 / C5 sub-records')
    ap.add_argument('--c4-vectors', type=int, default=100000)
    ap.add_argument('--c5-mocks', type=int, default=0, help='default: 1000 at 8 GPUs, 125 per GPU below')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))

    if args.impl == 'reference':
        run_reference_arm(args, rank, world, real_stdout)
        return

    from vega_b200 import synthetic, _lib
    t_init = time.perf_counter()
    vega = build_state()
    init = {'ini_fits_synthetic_operands_s': time.perf_counter() - t_init}
    B = args.batch
    theta = synthetic.draw_theta(vega, B, seed=20261104 + rank)
    names = list(vega.sample_params['limits'])

    base, oracle_values = None, None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process: the replicas are forked.  The mock data are
        # drawn here, around the ORACLE's fiducial model, and kept for the GPU legs.
        base, _, oracle_values = cpu_baseline(vega, theta, per_core=2)

    import torch
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    dd = Dist(torch, world)
    lib = _lib.load()
    peaks = measured_peaks()

    if oracle_values is None:
        synthetic.draw_mock_data(vega, vega.compute_model())
    t0 = time.perf_counter()
    engine = vega._engine_for()
    torch.cuda.synchronize()
    init['plan_build_finalize_s'] = time.perf_counter() - t0
    tt = engine.theta_tensor(theta)
    chi2_t = torch.empty(B, dtype=torch.float64, device='cuda')
    status_t = torch.empty(B, dtype=torch.int32, device='cuda')
    gathered = torch.empty(world * B, dtype=torch.float64, device='cuda') if world > 1 else None

    def step_device():
        engine.chi2_device(tt, chi2_t, status_t)
        if world > 1:
            dist.all_gather_into_tensor(gathered, chi2_t)

    for _ in range(max(args.warmup, 3)):
        step_device()
    dd.barrier()
    dgemm_peak, dfma_peak = fp64_peaks(lib, torch) if rank == 0 else (None, None)

    # ---- timed region 1: device-resident ----
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
        clocks.mark()
    dd.barrier()
    for _ in range(3):            # (the wait for the sampler's first line left the GPU idle)
        step_device()
    lib.vg_launch_count(1)
    ms_per_step = timed_device(torch, dd, step_device, args.steps)
    launches = lib.vg_launch_count(0)
    # ... and the same K steps again with the per-stage events on (the roofline figures): the events
    # keep every correlation on one stream (VG_OPT_LANES overlaps them otherwise, and overlapping
    # stages have no separate durations), so this pass is a few per cent slower than the headline
    lib.vg_profile_enable(engine.handle, 1)
    ms_per_step_staged = timed_device(torch, dd, step_device, args.steps)
    ms_stage = (C.c_double * 16)()
    cnt_stage = (C.c_int64 * 16)()
    lib.vg_profile_read(engine.handle, ms_stage, cnt_stage)
    lib.vg_profile_enable(engine.handle, 0)
    # K steps of ~5 ms are shorter than the sampler's period: the SAME step keeps running (untimed) until
    # the sampler has seen 0.6 s of this load, so the clocks line describes the load that was timed
    t_load = time.perf_counter()
    while time.perf_counter() - t_load < 0.6:
        step_device()
    torch.cuda.synchronize()
    clock_info = clocks.stop() if rank == 0 else None
    if clock_info is not None:
        clock_info['window'] = 'the timed steps plus 0.6 s of the same step, untimed (50 ms sampling period)'

    # ---- timed region 2: end to end through the public API with host buffers; at N > 1 the
    # per-rank chi2 values are gathered (device all-gather, read back) inside the timed region ----
    g_in = torch.empty(B, dtype=torch.float64, device='cuda') if world > 1 else None

    def step_e2e():
        c = vega.chi2_batch(theta)
        if world > 1:
            g_in.copy_(torch.from_numpy(c))
            dist.all_gather_into_tensor(gathered, g_in)
            return gathered.cpu().numpy()
        return c
    for _ in range(2):
        step_e2e()
    dd.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c_all = step_e2e()
    torch.cuda.synchronize()
    e2e_ms = dd.max(time.perf_counter() - t0) * 1e3 / args.steps
    c_host = c_all[rank * B:(rank + 1) * B] if world > 1 else c_all

    # ---- comparison legs, not the headline ----
    digits = [engine.ozaki_slices(ci)[0] for ci in range(len(engine.infos))]
    fused = [engine.fused_metal_operator(ci) for ci in range(len(engine.infos))]
    dm_digits = [engine.ozaki_dm_slices(ci)[0] for ci in range(len(engine.infos))]
    chi2_ref = chi2_t.clone()
    engine.set_option(_lib.OPT_OZAKI, 0)            # metal matrices on the FP64 DMMA kernel ...
    engine.set_option(_lib.OPT_FUSE_DM_METALS, 0)   # ... applied before the distortion matrix, as the reference does
    for _ in range(3):
        step_device()
    dmma_diff = float(((chi2_t - chi2_ref).abs() / chi2_ref.abs().clamp_min(1.0)).max().item())
    dmma_ms = timed_device(torch, dd, step_device, args.steps)
    engine.set_option(_lib.OPT_OZAKI, -1)
    engine.set_option(_lib.OPT_FUSE_DM_METALS, 1)
    if not dmma_diff <= 1e-10:      # the headline against its own FP64 two-step form, in every run
        raise SystemExit(f'bench: the headline path differs from the FP64 DMMA two-step path by {dmma_diff:.3e} (relative chi2)')
    n_folded = sum(engine.n_folded(ci) for ci in range(len(engine.infos)))
    engine.set_option(_lib.OPT_FOLD_METALS, 1)      # metal terms with a static xi stage from M_c X_m vectors
    for _ in range(3):
        step_device()
    fold_diff = float(((chi2_t - chi2_ref).abs() / chi2_ref.abs().clamp_min(1.0)).max().item())
    fold_ms = timed_device(torch, dd, step_device, args.steps)
    engine.set_option(_lib.OPT_FOLD_METALS, 0)
    vt_info = [engine.velo_tables(ci) for ci in range(len(engine.infos))]
    n_vt = sum(v[0] for v in vt_info)
    pp_ms = pp_diff = None
    if n_vt:
        engine.set_option(_lib.OPT_VELO_TABLES, 0)  # the sampled velocity dispersion through the per-point pass
        for _ in range(3):
            step_device()
        pp_diff = float(((chi2_t - chi2_ref).abs() / chi2_ref.abs().clamp_min(1.0)).max().item())
        pp_ms = timed_device(torch, dd, step_device, args.steps)
        engine.set_option(_lib.OPT_VELO_TABLES, 1)

    line = None
    if rank == 0:
        # the reference's scalar call (what iminuit / PolyChord make): one parameter dict -> float,
        # host in, host out, a batch of one through the same plan (reported, not the headline)
        scalar_calls = 50
        for i in range(3):
            vega.chi2(dict(zip(names, theta[i])))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(scalar_calls):
            scalar_value = vega.chi2(dict(zip(names, theta[i % B])))
        scalar_ms = (time.perf_counter() - t0) * 1e3 / scalar_calls
        assert abs(scalar_value - float(c_host[(scalar_calls - 1) % B])) <= 1e-6 * max(1.0, abs(scalar_value) / 1e4)

        flops = gemm_flops(engine, B)
        bytes_ = hbm_bytes(engine, B)
        stage_ms = {s: ms_stage[i] / args.steps for i, s in enumerate(STAGES)}
        n_eval = sum(engine.n_eval)
        n_moment = sum(engine.n_moment(ci) for ci in range(len(engine.infos)))
        n_own = sum(sum(1 for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct)
                    for info in engine.infos)
        n_grid_eval = n_own - n_moment - n_vt
        pk_instr = 0.0          # summed over the evaluations that go through the per-point pass
        for ci, info in enumerate(engine.infos):
            own = [d for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct]
            if engine.n_moment(ci) + vt_info[ci][0] >= len(own):
                continue
            twinned = {d.pk_twin for d in own if d.pk_twin >= 0}
            part = 0.0
            for e, d in enumerate(info['evals']):
                if any(d is o for o in own):
                    hcd = d.hcd_model and (d.hcd_on[0] or d.hcd_on[1])
                    part += 12.0 if (d.pk_twin >= 0 or e in twinned) else (18.5 if hcd else 16.5)
            pk_instr += part * (len(own) - engine.n_moment(ci) - vt_info[ci][0]) / len(own)

        def frac(a, b):
            return a / b if a is not None and b else None
        kernels = {s: {'ms_per_step': stage_ms[s], 'launch_groups_per_step': cnt_stage[i] / args.steps}
                   for i, s in enumerate(STAGES) if cnt_stage[i] > 0}
        roofline_kernels = []
        # -- the dense metal matrices: split-integer GEMM on tcgen05 (or DMMA when the option is off)
        ozaki_on = any(digits)
        metal_ms = stage_ms['metal_gemm']
        if ozaki_on and metal_ms > 0:
            # executed int8 multiply-adds: S(S+1)/2 digit products over the PADDED tiles (128 x 64 x 64)
            ops = 0.0
            for ci, info in enumerate(engine.infos):
                S = digits[ci]
                if not S:
                    continue
                nm = info['item'].model_coordinates.r_grid.size
                n_metal = sum(1 for d in info['evals'] if d.out_slot >= 2 and not d.frozen and not d.direct)
                k = n_metal * ((nm + 15) // 16 * 16)
                rows, S_fused = fused[ci]          # chi2 path: (D_masked M_c) has n_mask rows, its own digits
                n_out = rows if rows else nm
                S = S_fused if rows else S
                ops += S * (S + 1) / 2 * 2.0 * (-(-B // 128) * 128) * (-(-n_out // 128) * 128) * (-(-k // 64) * 64)
            int8_peak = 2.0 * peaks['bf16_tflops']
            tops = ops / (metal_ms * 1e-3) / 1e12
            main_roofline = {
                'bound': 'tensor', 'kernel': 'ozaki_gemm_kernel (tcgen05.mma kind::i8, TMA digit planes, s32 accumulators '
                                             'in TMEM) + its split-K reduction',
                'achieved': tops, 'peak': int8_peak, 'unit': 'TOP/s (int8, executed digit products)', 'frac': tops / int8_peak,
                'peak_source': f"2 x bf16_tflops (burst) of {peaks['source']}: int8 runs at twice the bf16 rate on B200 "
                               '(4.5 vs 2.25 POP/s nominal).  The sustained bf16 figure is a power-capped cuBLAS number that '
                               'this kernel exceeds (int8 digit products draw less power per operation), so the burst figure '
                               'is the denominator; the other two fractions are carried next to it',
                'frac_of_2x_bf16_sustained': tops / (2.0 * peaks['bf16_tflops_sustained']),
                'frac_of_nominal_int8_4500': tops / 4500.0,
                'cluster': 'CTA pairs along the column tiles share the batch operand tile through multicast TMA (VG_OZ_CLUSTER)',
                'algorithmic_fp64_tflops': flops['metal_gemm'] / (metal_ms * 1e-3) / 1e12,
                'fp64_dgemm_peak_tflops': dgemm_peak,
                'frac_of_fp64_dgemm_peak': frac(flops['metal_gemm'] / (metal_ms * 1e-3) / 1e12, dgemm_peak),
                'digits': digits,
                # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch (auto-correlation, M=1024
                # N=2500 K=37680, S=7, 6 splits) from the committed ncu --set full capture; algorithmic bytes
                # of that launch: 7 x (1024 + 2500) x 37680 digit bytes + 6 x 1024 x 2512 x 8 partial sums
                # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch (auto-correlation, fused operator:
                # M = 1024, N = 1590, K = 37680, S = 8, 7 splits, CTA pairs) from the committed ncu --set full
                # capture; algorithmic bytes of that launch: 8 x (1024 + 1590) x 37680 digit bytes + 7 x 1024 x
                # 1600 x 8 of partial sums
                'traffic': 1.283e9 if B == 1024 else None, 'algorithmic_bytes_of_that_launch': 0.880e9,
                'traffic_source': 'profiles/r02_ncu_full_c2_lanes_raw.csv (ozaki_gemm_kernel<8, 2>, grid 8x26x7: 1.195 GB '
                                  'read + 0.088 GB written)',
                'ncu_tensor_pipe_active': [0.726, 0.797],
                'share_of_step': metal_ms / ms_per_step_staged}
        else:
            gemm_ms = sum(stage_ms[s] for s in ('hankel_gemm', 'metal_gemm', 'dm_gemm', 'chi2_gemm'))
            achieved = sum(flops[s] for s in ('hankel_gemm', 'metal_gemm', 'dm_gemm', 'chi2_gemm')) / (gemm_ms * 1e-3) / 1e12
            main_roofline = {'bound': 'tensor', 'kernel': 'dgemm_tn_kernel (DMMA FP64: Hankel, metal, distortion, C^-1)',
                             'achieved': achieved, 'peak': dgemm_peak, 'unit': 'TFLOP/s', 'frac': frac(achieved, dgemm_peak),
                             'traffic': 1.615e9 if B == 1024 else None,
                             'traffic_source': 'profiles/r01_ncu_full_c2_raw.csv (dgemm_tn_kernel<0>, grid 8x20x12)',
                             'share_of_step': gemm_ms / ms_per_step_staged}
        roofline_kernels.append(dict(main_roofline, name='metal_gemm'))
        for s in ('hankel_gemm', 'dm_gemm', 'chi2_gemm', 'velo_tables'):
            if stage_ms[s] > 0:
                tf = flops[s] / (stage_ms[s] * 1e-3) / 1e12
                kernels[s]['tflops'] = tf
                kname = 'dgemm_tn_kernel (DMMA m8n8k4 FP64)'
                if s == 'velo_tables':
                    kname = ('vtab_weights_kernel + dgemm_tn_kernel (DMMA): knot tables of the evaluations with a sampled '
                             f'velocity dispersion from their Chebyshev series, orders {[v[1] for v in vt_info if v[0]]}')
                if s == 'chi2_gemm':
                    kname = ('dgemm_tn_kernel<ROWDOT> (DMMA), symmetric form: one triangle of C^-1 is read, the executed '
                             'flops are (1 + 128/N) / 2 of the algorithmic 2 N^2 the figure counts')
                if s == 'dm_gemm' and any(dm_digits):
                    kname = f'ozaki_gemm_kernel (tcgen05 kind::i8, digits {dm_digits}) + split-K reduction'
                roofline_kernels.append({'name': s, 'kernel': kname, 'bound': 'tensor',
                                         'achieved': tf, 'peak': dgemm_peak, 'unit': 'TFLOP/s (algorithmic FP64)',
                                         'frac': frac(tf, dgemm_peak), 'ms_per_step': stage_ms[s],
                                         'share_of_step': stage_ms[s] / ms_per_step_staged})
        if stage_ms['metal_rp'] > 0:
            tf = flops['metal_rp'] / (stage_ms['metal_rp'] * 1e-3) / 1e12
            roofline_kernels.append({'name': 'metal_rp', 'kernel': 'metal_rp_dmma_kernel', 'bound': 'tensor', 'achieved': tf,
                                     'peak': dgemm_peak, 'unit': 'TFLOP/s (algorithmic FP64)', 'frac': frac(tf, dgemm_peak),
                                     'ms_per_step': stage_ms['metal_rp'], 'share_of_step': stage_ms['metal_rp'] / ms_per_step_staged})
        if stage_ms['pk_ell'] > 0:
            tf = 2 * pk_instr * 814000.0 * B / (stage_ms['pk_ell'] * 1e-3) / 1e12
            kernels['pk_ell']['grid_points_per_s'] = 814000.0 * n_grid_eval * B / (stage_ms['pk_ell'] * 1e-3)
            roofline_kernels.append({'name': 'pk_ell', 'kernel': 'pk_ell_fast_kernel (fused P(k,mu) -> multipoles)',
                                     'bound': 'fp64 alu', 'achieved': tf, 'peak': dfma_peak,
                                     'unit': 'TFLOP/s (FP64 instructions of the SASS main loop x 2)', 'frac': frac(tf, dfma_peak),
                                     'fp64_instr_per_point_vector': pk_instr, 'ncu_fp64_pipe_active': [0.727, 0.675],
                                     'ms_per_step': stage_ms['pk_ell'], 'share_of_step': stage_ms['pk_ell'] / ms_per_step_staged})
        for s, kname in (('xi_gather', 'xi_gather_fast_kernel / xi_gather_kernel'), ('moment_xi', 'moment_xi_kernel'),
                         ('moment_tables', 'moment_tables_kernel'), ('ozaki_slice', 'ozaki_rowexp_kernel + ozaki_slice_kernel'),
                         ('elementwise', 'expand/scatter_theta, combine_pre, post, chi2_partials, prior kernels')):
            if stage_ms.get(s, 0) > 0:
                gbs = bytes_[s] / (stage_ms[s] * 1e-3) / 1e9
                roofline_kernels.append({'name': s, 'kernel': kname, 'bound': 'hbm', 'achieved': gbs, 'peak': peaks['hbm_gbs'],
                                         'unit': 'GB/s (algorithmic bytes, SURVEY 8d)', 'frac': gbs / peaks['hbm_gbs'],
                                         'ms_per_step': stage_ms[s], 'share_of_step': stage_ms[s] / ms_per_step_staged})
        line = {
            'metric': 'chi2 evals/sec', 'value': world * B / (ms_per_step * 1e-3), 'unit': 'evals/s',
            'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms_per_step,
            'ms_per_step_with_stage_events': ms_per_step_staged,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'batch_per_gpu': B, 'n_theta': len(engine.sampled),
                       'evaluations_per_chi2': n_eval,
                       'evaluations_through_pk_ell': n_grid_eval,
                       'evaluations_in_moment_form': n_moment,
                       'evaluations_on_velocity_tables': n_vt,
                       'evaluations_sharing_tables': n_eval - n_grid_eval - n_moment - n_vt,
                       'evaluations_with_static_xi_stage': sum(engine.n_static_xi(ci)
                                                               for ci in range(len(engine.infos))),
                       'metal_gemm_digits': digits, 'distortion_gemm_digits': dm_digits,
                       'fused_distortion_metal_operator': [{'rows': r, 'digits': s} for r, s in fused],
                       'whitened_chi2': [engine.whitened(ci) for ci in range(len(engine.infos))],
                       'lanes': 'correlation i on stream i % 4 (VG_OPT_LANES); one stream in the pass with stage events',
                       'plan_device_bytes': int(engine.device_bytes()),
                       'l2': 'inputs larger than L2: ~2.9 GB of static operands + ~2 GB of workspace '
                             'streamed per step'},
            'clocks': clock_info,
            'e2e': {'value': world * B / (e2e_ms * 1e-3), 'unit': 'evals/s', 'ms_per_step': e2e_ms,
                    'h2d_bytes_per_step': int(theta.nbytes), 'd2h_bytes_per_step': int(B * 12 + (world * B * 8 if world > 1 else 0)),
                    'includes_gather': world > 1,
                    'chi2_checksum': float(np.sum(c_host[c_host < 1e99]))},
            'gpu_launches': int(launches),
            'roofline': main_roofline,
            'roofline_kernels': roofline_kernels,
            'peaks': {'hbm_gbs': peaks['hbm_gbs'], 'bf16_tflops': peaks['bf16_tflops'],
                      'bf16_tflops_sustained': peaks['bf16_tflops_sustained'], 'source': peaks['source'],
                      'fp64_dgemm_tflops_measured': dgemm_peak, 'fp64_dfma_tflops_measured': dfma_peak,
                      'fp64_tflops_nominal': 37.0,
                      'fp64_note': 'measured live in this run (cuBLAS DGEMM 8192^3 through torch.matmul; FMA-chain probe); '
                                   'nominal: 296 TFLOP/s FP64 per 8-GPU HGX B200'},
            'kernels': kernels,
            'init_s': init,
            'with_dmma_metals': {'note': 'VG_OPT_OZAKI = 0, VG_OPT_FUSE_DM_METALS = 0: the dense metal matrices, then the '
                                         'distortion matrix, both on the FP64 DMMA kernel (the parity reference of the '
                                         'split-integer GEMM and of the fused distortion x metal operator; the run fails '
                                         'when the headline differs from it by more than 1e-10)', 'ms_per_step': dmma_ms,
                                 'value': world * B / (dmma_ms * 1e-3), 'unit': 'evals/s',
                                 'max_rel_chi2_difference_to_headline': dmma_diff},
            'with_folded_metals': {'note': 'not the headline: VG_OPT_FOLD_METALS = 1, the metal terms with a '
                                           'static xi stage read from M_c X_m vectors built at finalize',
                                   'folded_metal_terms': n_folded, 'ms_per_step': fold_ms,
                                   'value': world * B / (fold_ms * 1e-3), 'unit': 'evals/s',
                                   'max_rel_chi2_difference_to_headline': fold_diff},
            'with_per_point_velocity': None if pp_ms is None else {
                'note': 'VG_OPT_VELO_TABLES = 0: the evaluations with a sampled velocity dispersion through the per-point '
                        'P(k,mu) pass and the Hankel GEMM (the parity reference of the velocity-dispersion tables)',
                'tables': [{'evaluations': v[0], 'chebyshev_orders': v[1], 'error_at_finalize': v[2]} for v in vt_info],
                'ms_per_step': pp_ms, 'value': world * B / (pp_ms * 1e-3), 'unit': 'evals/s',
                'max_rel_chi2_difference_to_headline': pp_diff},
            'scalar_api': {'call': 'VegaInterface.chi2(params) -> float, batch of one', 'ms_per_call': scalar_ms,
                           'calls': scalar_calls},
            'cpu_baseline': base,
        }
        if oracle_values is not None:
            line['parity'] = parity_block(vega, theta, names, oracle_values)

    # ---- C3 / C4 / C5 on the DESI-shaped model (all ranks take part in C4 and C5) ----
    if not args.no_extra:
        del engine, tt
        vega._engine, vega._engines = None, {}
        torch.cuda.empty_cache()
        vega3, names3, init3 = build_c3(rank)
        c3, fiducial3 = run_c3(vega3, names3, torch, dd, rank, lib)
        dd.barrier()
        c4 = run_c4(vega3, torch, dd, world, args.c4_vectors)
        n_mocks = args.c5_mocks or (1000 if world == 8 else 125 * world)
        c5 = run_c5(vega3, fiducial3, torch, dd, world, n_mocks)
        if rank == 0:
            c3['init_s'] = init3
            line['c3'] = dict(c3, workload='C3 DESI-DR2-shaped: lyaxlya, lyaxlyb, lyaxqso, lybxqso; model grids 5625/11250 bins; '
                                           'rp-only new_metals matrices; fast_metals; global covariance over 9540 masked bins')
            line['c4'] = c4
            line['c5'] = c5

    if rank == 0:
        if line.get('parity') is not None and not line['parity']['ok']:
            print(json.dumps(line), file=real_stdout, flush=True)
            print('bench.py: parity against the oracle FAILED: ' + json.dumps(line['parity']), file=sys.stderr)
            sys.exit(3)
        print(json.dumps(line), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
