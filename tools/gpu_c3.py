"""C3 (DESI-DR2-shaped, full size) on the device: chi2 pass time for the batches an iminuit-style
fit consumes (1, a gradient stencil 2P+1, a Hessian stencil 2P^2+1) and one complete fit
through the batched minimizer, next to the CPU oracle's time per evaluation."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
from helpers import make_c3_config, static_tables  # noqa: E402
from oracle.vega_oracle import Oracle  # noqa: E402
from vega_b200 import VegaInterface, synthetic  # noqa: E402
from vega_b200.drivers import BatchedMinimizer, stencil, stencil_size  # noqa: E402

out = {}
t0 = time.perf_counter()
vega = VegaInterface(make_c3_config('/tmp/c3_bench'))
synthetic.inject_dense_operands(vega, with_distortion=False, with_metals=False)
synthetic.inject_global_covariance(vega)
out['init_s'] = time.perf_counter() - t0
oracle = Oracle(vega, tables=static_tables(), fast_metals=True)
synthetic.draw_mock_data(vega, oracle.compute_model())
names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'drp_QSO',
         'sigma_velo_disp_lorentz_QSO', 'bao_amp', 'bias_eta_SiII(1260)', 'bias_eta_SiIII(1207)',
         'bias_eta_SiII(1193)', 'bias_eta_SiII(1190)', 'bias_eta_CIV(eff)', 'bias_QSO',
         'BB-lyalya_lyalya-0 add post r,mu (1,2)']
names = [n for n in names if n in vega.params]
P = len(names)
base = np.array([vega.params[n] for n in names])
steps = 0.02 * np.abs(base) + 1e-4
t0 = time.perf_counter()
ref = oracle.chi2(dict(zip(names, base)))
out['oracle_s_per_eval'] = time.perf_counter() - t0
engine = vega._engine_for(tuple(names))
out['n_params'] = P
out['evaluations'] = {'per_chi2': int(sum(engine.n_eval)), 'moment_form': int(sum(engine.n_moment(ci) for ci in range(4))),
                      'static_xi': int(sum(engine.n_static_xi(ci) for ci in range(4)))}
for label, theta in (('single', base[None]), ('gradient_stencil', stencil(base, steps, hessian=False)),
                     ('hessian_stencil', stencil(base, steps, hessian=True))):
    tt = engine.theta_tensor(vega._reorder(engine, theta, names))
    for _ in range(3):
        chi2, status = engine.chi2_device(tt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        chi2, status = engine.chi2_device(tt)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    out[label] = {'vectors': int(theta.shape[0]), 'ms_per_pass': ms, 'evals_per_s': theta.shape[0] / ms * 1e3}
    if label == 'single':
        out[label]['chi2_minus_oracle'] = float(chi2[0].item() - ref)
    import ctypes as C
    import bench
    lib = engine.lib
    lib.vg_profile_enable(engine.handle, 1)
    lib.vg_launch_count(1)
    for _ in range(5):
        engine.chi2_device(tt)
    torch.cuda.synchronize()
    ms_stage, cnt_stage = (C.c_double * 16)(), (C.c_int64 * 16)()
    lib.vg_profile_read(engine.handle, ms_stage, cnt_stage)
    lib.vg_profile_enable(engine.handle, 0)
    out[label]['stages_ms'] = {s_: round(ms_stage[i] / 5, 4) for i, s_ in enumerate(bench.STAGES)}
    out[label]['launches'] = lib.vg_launch_count(0) / 5
    print(label, out[label], flush=True)
assert stencil_size(P, True) == out['hessian_stencil']['vectors']
# one complete fit: every Newton iteration is one Hessian-stencil batch
sample = {k: {n: vega.sample_params[k].get(n) for n in names} for k in ('limits', 'values', 'errors', 'fix')} \
    if all(n in vega.sample_params['limits'] for n in names) else None
t0 = time.perf_counter()
vega.minimize()
out['minimize'] = {'wall_s': time.perf_counter() - t0, 'sampled': list(vega.sample_params['limits']),
                   'fval': float(vega.minimizer.fmin.fval), 'valid': bool(vega.minimizer.fmin.is_valid)}
print(out['minimize'], flush=True)
Path('gpurun_out').mkdir(exist_ok=True)
json.dump(out, open('gpurun_out/c3.json', 'w'), indent=1)
