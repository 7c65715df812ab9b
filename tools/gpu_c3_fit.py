"""Wall time of complete fits on the DESI-shaped C3 model, repeated in one process (the first fit
pays the one-off costs: workspace growth, CUDA-graph captures)  (python tools/gpu_c3_fit.py)."""
import json, sys, os, time
from pathlib import Path
import numpy as np
import torch
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / 'tests'))
os.chdir(REPO / 'tests')
import bench
from vega_b200 import synthetic

vega, names, init = bench.build_c3(0)
synthetic.draw_mock_data(vega, vega.compute_model())
engine = vega._engine_for()
out = []
if len(sys.argv) > 1 and sys.argv[1] == 'passes':       # what bench.py's c3 record times before its fit
    from vega_b200.drivers import stencil
    base = np.array([vega.params[n] for n in names])
    steps = 0.02 * np.abs(base) + 1e-4
    for theta in (base[None], stencil(base, steps), stencil(base, steps, hessian=True)):
        tt = engine.theta_tensor(theta)
        for prof in (0, 1):
            engine.lib.vg_profile_enable(engine.handle, prof)
            for _ in range(14):
                engine.chi2_device(tt)
            torch.cuda.synchronize()
        engine.lib.vg_profile_enable(engine.handle, 0)
for rep in range(4):
    torch.cuda.synchronize()
    r0 = engine.lib.vg_plan_graph_replays(engine.handle)
    t0 = time.perf_counter()
    vega.minimize()
    torch.cuda.synchronize()
    mini = vega.minimizer
    out.append({'wall_s': time.perf_counter() - t0, 'chi2_evaluations': int(mini.n_eval), 'fval': float(mini.fmin.fval),
                'graph_replays': int(engine.lib.vg_plan_graph_replays(engine.handle) - r0)})
    print(out[-1], flush=True)
print('plan bytes after the fits', engine.device_bytes() / 1e9, 'GB')
theta = synthetic.draw_theta(vega, 1024, width=1.0, names=names)
engine.chi2_device(engine.theta_tensor(theta))
torch.cuda.synchronize()
print('plan bytes after a 1024-vector chunk', engine.device_bytes() / 1e9, 'GB')
print(json.dumps(out))
