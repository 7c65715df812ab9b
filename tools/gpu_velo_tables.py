"""Velocity-dispersion tables on the C2 workload: what finalize chose, and the pass time with and
without them (python tools/gpu_velo_tables.py)."""
import json, sys, time
from pathlib import Path
import numpy as np
import torch
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / 'tests'))
import os
os.chdir(REPO / 'tests')
from vega_b200 import VegaInterface, synthetic, _lib

vega = VegaInterface('configs/c2/main.ini')
synthetic.inject_dense_operands(vega)
t0 = time.time()
engine = vega._engine_for()
out = {'finalize_s': time.time() - t0,
       'tables': {engine.infos[ci]['item'].name: engine.velo_tables(ci) for ci in range(len(engine.infos))}}
theta = synthetic.draw_theta(vega, 1024)
th = torch.as_tensor(theta, device='cuda')
def timed(n=10):
    for _ in range(3): engine.chi2_device(th)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): c, s = engine.chi2_device(th)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, c.cpu().numpy()
ms_tab, c_tab = timed()
out['active'] = engine.velo_tables_active()
engine.set_option(_lib.OPT_VELO_TABLES, 0)
ms_pp, c_pp = timed()
out['ms_tables'] = ms_tab; out['ms_per_point'] = ms_pp
out['max_rel_dchi2'] = float(np.max(np.abs(c_tab - c_pp) / np.abs(c_pp)))
print(json.dumps(out))
