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
STAGES = ['pk_ell', 'hankel_gemm', 'xi_gather', 'metal_gemm', 'dm_gemm', 'chi2_gemm', 'elementwise',
          'moment_tables']


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


def gemm_flops(engine, B):
    """Algorithmic FP64 flops of the contraction stages for B vectors:
    Hankel operator 2*n_ell*nk^2 per evaluation, metal matrices 2*Nm^2 each, distortion
    matrix 2*Nmask*Nm (masked rows), chi2 2*Nmask^2."""
    out = {'hankel_gemm': 0.0, 'metal_gemm': 0.0, 'dm_gemm': 0.0, 'chi2_gemm': 0.0}
    nk, n_ell = 814, 4
    for ci, info in enumerate(engine.infos):
        item, data = info['item'], info['data']
        nm = item.model_coordinates.r_grid.size
        nmask = int(data.model_mask.sum())
        own = [d for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct]
        n_eval = len(own) - engine.n_moment(ci)             # evaluations that reach the Hankel GEMM
        # metal matrices applied per vector (fast_metals freezes the metal x metal ones)
        n_metal = sum(m is not None and not d.frozen for m, d in zip(info['metal_mats'], info['evals'][2:]))
        out['hankel_gemm'] += 2.0 * n_ell * nk * nk * n_eval * B
        out['metal_gemm'] += 2.0 * nm * nm * n_metal * B
        if not data.distortion_is_identity:
            out['dm_gemm'] += 2.0 * nmask * nm * B
        if not data.cov_is_identity:
            out['chi2_gemm'] += 2.0 * nmask * nmask * B
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


def run_reference_arm(args, rank, world, out=sys.stdout):
    if rank != 0:
        return
    vega = build_state()
    from vega_b200 import synthetic
    theta = synthetic.draw_theta(vega, args.batch)
    base, dt = cpu_baseline(vega, theta, per_core=1, steps=args.steps, warmup=max(args.warmup, 1))
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


def main():
    real_stdout = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=10)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--batch', type=int, default=1024, help='parameter vectors per GPU per step')
    ap.add_argument('--impl', default='b200')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    args = ap.parse_args()
    rank = int(os.environ.get('RANK', 0))
    world = int(os.environ.get('WORLD_SIZE', 1))
    local_rank = int(os.environ.get('LOCAL_RANK', 0))

    if args.impl == 'reference':
        run_reference_arm(args, rank, world, real_stdout)
        return

    from vega_b200 import synthetic, _lib
    vega = build_state()
    B = args.batch
    theta = synthetic.draw_theta(vega, B, seed=20261104 + rank)

    base = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process: the replicas are forked
        base, _ = cpu_baseline(vega, theta, per_core=2)

    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local_rank))
    lib = _lib.load()

    synthetic.draw_mock_data(vega, vega.compute_model())
    engine = vega._engine_for()
    tt = engine.theta_tensor(theta)
    chi2_t = torch.empty(B, dtype=torch.float64, device='cuda')
    status_t = torch.empty(B, dtype=torch.int32, device='cuda')
    gathered = torch.empty(world * B, dtype=torch.float64, device='cuda') if world > 1 else None

    def step_device():
        engine.chi2_device(tt, chi2_t, status_t)
        if world > 1:
            dist.all_gather_into_tensor(gathered, chi2_t)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    dgemm_peak, dfma_peak = fp64_peaks(lib, torch) if rank == 0 else (None, None)
    barrier()

    # ---- timed region 1: device-resident, per-stage events on ----
    clocks = ClockSampler(local_rank)
    if rank == 0:
        clocks.start()
    lib.vg_profile_enable(engine.handle, 1)
    lib.vg_launch_count(1)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    launches = lib.vg_launch_count(0)
    elapsed = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
    ms_stage = (C.c_double * 8)()
    cnt_stage = (C.c_int64 * 8)()
    lib.vg_profile_read(engine.handle, ms_stage, cnt_stage)
    lib.vg_profile_enable(engine.handle, 0)
    clock_info = clocks.stop() if rank == 0 else None
    ms_per_step = float(elapsed.item()) / args.steps

    # ---- timed region 2: end to end through the public API with host buffers ----
    for _ in range(2):
        vega.chi2_batch(theta)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        c_host = vega.chi2_batch(theta)
    torch.cuda.synchronize()
    e2e_s = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(e2e_s, op=dist.ReduceOp.MAX)
    e2e_ms = float(e2e_s.item()) * 1e3 / args.steps

    # ---- extra, not the headline: VG_OPT_FOLD_METALS on (metal terms whose xi stage is static
    # served from vectors M_c X_m built at finalize instead of the per-vector matrix product) ----
    from vega_b200 import _lib as vlib
    n_folded = sum(engine.n_folded(ci) for ci in range(len(engine.infos)))
    engine.set_option(vlib.OPT_FOLD_METALS, 1)
    chi2_ref = chi2_t.clone()
    for _ in range(3):
        step_device()
    barrier()
    fold_diff = float((chi2_t - chi2_ref).abs().max().item())
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record()
    for _ in range(args.steps):
        step_device()
    f1.record()
    barrier()
    fold_elapsed = torch.tensor([f0.elapsed_time(f1)], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(fold_elapsed, op=dist.ReduceOp.MAX)
    fold_ms = float(fold_elapsed.item()) / args.steps
    engine.set_option(vlib.OPT_FOLD_METALS, 0)

    if rank == 0:
        # the reference's scalar call (what iminuit / PolyChord make): one parameter dict -> float,
        # host in, host out, a batch of one through the same plan (reported, not the headline)
        names = list(engine.sampled)
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
        stage_ms = {s: ms_stage[i] / args.steps for i, s in enumerate(STAGES)}
        gemm_ms = sum(stage_ms[s] for s in flops)
        achieved = sum(flops.values()) / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
        n_eval = sum(engine.n_eval)
        n_moment = sum(engine.n_moment(ci) for ci in range(len(engine.infos)))
        n_own = sum(sum(1 for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct)
                    for info in engine.infos)
        n_grid_eval = n_own - n_moment
        pk_instr = 0.0          # summed over the evaluations that go through the per-point pass
        for ci, info in enumerate(engine.infos):
            own = [d for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct]
            if engine.n_moment(ci) >= len(own):
                continue
            twinned = {d.pk_twin for d in own if d.pk_twin >= 0}
            part = 0.0
            for e, d in enumerate(info['evals']):
                if any(d is o for o in own):
                    hcd = d.hcd_model and (d.hcd_on[0] or d.hcd_on[1])
                    part += 12.0 if (d.pk_twin >= 0 or e in twinned) else (18.5 if hcd else 16.5)
            pk_instr += part * (len(own) - engine.n_moment(ci)) / len(own)
        grid_bytes = 6512000.0 * n_grid_eval * B
        kernels = {s: {'ms_per_step': stage_ms[s], 'launch_groups_per_step': cnt_stage[i] / args.steps}
                   for i, s in enumerate(STAGES)}
        for s in flops:
            if stage_ms[s] > 0:
                kernels[s]['tflops'] = flops[s] / (stage_ms[s] * 1e-3) / 1e12
        if stage_ms['pk_ell'] > 0:
            kernels['pk_ell']['grid_points_per_s'] = 814000.0 * n_grid_eval * B / (stage_ms['pk_ell'] * 1e-3)
            kernels['pk_ell']['equivalent_grid_GBps'] = grid_bytes / (stage_ms['pk_ell'] * 1e-3) / 1e9
        line = {
            'metric': 'chi2 evals/sec', 'value': world * B / (ms_per_step * 1e-3), 'unit': 'evals/s',
            'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3), 'ms_per_step': ms_per_step,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64',
            'data': 'synthetic',
            'config': {'workload': WORKLOAD, 'batch_per_gpu': B, 'n_theta': len(engine.sampled),
                       'evaluations_per_chi2': n_eval,
                       'evaluations_through_pk_ell': n_grid_eval,
                       'evaluations_in_moment_form': n_moment,
                       'evaluations_sharing_tables': n_eval - n_grid_eval - n_moment,
                       'evaluations_with_static_xi_stage': sum(engine.n_static_xi(ci)
                                                               for ci in range(len(engine.infos))),
                       'l2': 'inputs larger than L2: ~1.6 GB of static operands + ~2 GB of workspace '
                             'streamed per step'},
            'clocks': clock_info,
            'e2e': {'value': world * B / (e2e_ms * 1e-3), 'unit': 'evals/s', 'ms_per_step': e2e_ms,
                    'h2d_bytes_per_step': int(theta.nbytes), 'd2h_bytes_per_step': int(B * 12),
                    'chi2_checksum': float(np.sum(c_host[c_host < 1e99]))},
            'gpu_launches': int(launches),
            'roofline': {'bound': 'tensor', 'kernel': 'dgemm_tn_kernel (DMMA FP64: Hankel, metal, distortion, C^-1)',
                         'achieved': achieved, 'peak': dgemm_peak, 'unit': 'TFLOP/s',
                         'frac': achieved / dgemm_peak if dgemm_peak else None,
                         # DRAM bytes of ONE launch of the largest GEMM (auto-correlation metal
                         # matrices, M=1024 N=2500 K=37680, 12 splits) from the committed ncu --set full
                         # capture: dram__bytes_read.sum 1.384 GB + dram__bytes_write.sum 0.231 GB
                         # (profiles/r01_ncu_full_c2_raw.csv); algorithmic operand bytes 1.083 GB, the
                         # rest is the split-K partial sums
                         'traffic': 1.615e9 if B == 1024 else None,
                         'traffic_source': 'profiles/r01_ncu_full_c2_raw.csv (dgemm_tn_kernel<0>, grid 8x20x12)',
                         'peak_source': 'cuBLAS FP64 8192^3 measured live in this run (MEASURED_PEAKS.json has no '
                                        'FP64 figure); algorithmic flops per SURVEY.md 8d',
                         'share_of_step': gemm_ms / ms_per_step},
            # FP64 instructions per grid point, vector and evaluation from the SASS main loops of
            # the fast kernels (cuobjdump): 16.5 (Kaiser x Lorentzian), 18.5 (+ HCD), 12 for each
            # member of a peak/smooth pair evaluated together; counted as FMAs against the DFMA peak
            'roofline_pk_ell': {'bound': 'fp64 alu', 'peak': dfma_peak, 'unit': 'TFLOP/s (FP64 instr x 2)',
                                'fp64_instr_per_point_vector': pk_instr,
                                'achieved': (2 * pk_instr * 814000.0 * B / (stage_ms['pk_ell'] * 1e-3) / 1e12
                                             if stage_ms['pk_ell'] > 0 else None),
                                'frac': (2 * pk_instr * 814000.0 * B / (stage_ms['pk_ell'] * 1e-3) / 1e12
                                         / dfma_peak if stage_ms['pk_ell'] > 0 and dfma_peak else None),
                                'ncu_fp64_pipe_active': [0.727, 0.675],
                                'share_of_step': stage_ms['pk_ell'] / ms_per_step},
            'kernels': kernels,
            'with_folded_metals': {'note': 'not the headline: VG_OPT_FOLD_METALS = 1, the metal terms with a '
                                           'static xi stage read from M_c X_m vectors built at finalize',
                                   'folded_metal_terms': n_folded, 'ms_per_step': fold_ms,
                                   'value': world * B / (fold_ms * 1e-3), 'unit': 'evals/s',
                                   'max_abs_chi2_difference': fold_diff},
            'scalar_api': {'call': 'VegaInterface.chi2(params) -> float, batch of one', 'ms_per_call': scalar_ms,
                           'calls': scalar_calls},
            'cpu_baseline': base,
        }
        print(json.dumps(line), file=real_stdout, flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
