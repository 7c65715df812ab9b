"""Moment form and velocity-dispersion tables on the DESI-shaped C3 model: what finalize
chose per correlation, the per-point evaluations left, and the stencil passes with and without them
(python tools/gpu_c3_tables.py)."""
import json, sys, time
from pathlib import Path
import numpy as np
import torch
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / 'tests'))
import os
os.chdir(REPO / 'tests')
import bench
from vega_b200 import _lib, synthetic
from vega_b200.drivers import stencil

vega, names, init = bench.build_c3(0)
synthetic.draw_mock_data(vega, vega.compute_model())
engine = vega._engine_for()
out = {'init': init, 'tables': {}}
for ci, info in enumerate(engine.infos):
    hot = [d for d in info['evals'] if not d.direct and not d.frozen and d.table_owner < 0]
    out['tables'][info['item'].name] = {
        'served': engine.velo_tables(ci), 'own_tables': len(hot),
        'moment': engine.n_moment(ci),
        'kinds': [(d.out_slot, d.dyn_mask, d.hcd_model, list(d.hcd_on), d.fast_metals, d.same_tracer, list(d.idx_velo), d.uvb, d.uv_static)
                  for d in hot]}
base = np.array([vega.params[n] for n in names])
steps = 0.02 * np.abs(base) + 1e-4
def timed(tt, n=10):
    for _ in range(4): engine.chi2_device(tt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): c, s = engine.chi2_device(tt)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, c.cpu().numpy()
for label, theta in (('single', base[None]), ('gradient', stencil(base, steps)), ('hessian', stencil(base, steps, hessian=True))):
    tt = engine.theta_tensor(theta)
    ms_on, c_on = timed(tt)
    active = engine.velo_tables_active()
    engine.set_option(_lib.OPT_VELO_TABLES, 0)
    ms_off, c_off = timed(tt)
    engine.set_option(_lib.OPT_VELO_TABLES, 1)
    out[label] = {'vectors': int(theta.shape[0]), 'ms_tables': ms_on, 'ms_per_point': ms_off, 'active': active,
                  'max_rel_dchi2': float(np.max(np.abs(c_on - c_off) / np.abs(c_off)))}
# models: tables against the per-point pass on the Hessian stencil
theta = stencil(base, steps, hessian=True)
m_on, _ = vega.compute_model_batch(theta, names)
engine.set_option(_lib.OPT_VELO_TABLES, 0)
m_off, _ = vega.compute_model_batch(theta, names)
engine.set_option(_lib.OPT_VELO_TABLES, 1)
out['max_dxi_over_maxxi'] = {n: float(np.max(np.abs(m_on[n] - m_off[n])) / np.max(np.abs(m_off[n]))) for n in m_off}
print(json.dumps(out))
