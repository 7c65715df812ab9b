"""Iteration trace of the lock-step mock fits on the DESI-shaped model (diagnostics)."""
import sys
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / 'tests'))
import torch  # noqa: E402
import bench  # noqa: E402
from vega_b200 import synthetic  # noqa: E402
from vega_b200.drivers import MonteCarloFits  # noqa: E402

n_mocks = int(sys.argv[1]) if len(sys.argv) > 1 else 16
vega, names, _ = bench.build_c3(0)
fiducial = vega.compute_model()
synthetic.draw_mock_data(vega, fiducial)
mc = MonteCarloFits(vega, fiducial_model=fiducial)
mc.draw(n_mocks, seed=1)
for method, kw in (('newton', {}), ('quasi', {'share_hessian': True}), ('newton', {}), ('quasi', {'share_hessian': True})):
    t0 = time.perf_counter()
    res = mc.fit(max_iter=40, method=method, verbose=False, **kw)
    torch.cuda.synchronize()
    print(f'== {method} {kw}: {time.perf_counter() - t0:.2f} s, evals/mock {res["n_eval"] / n_mocks:.0f}, '
          f'converged {int(res["converged"].sum())}, iterations {res["n_iter"].tolist()}', flush=True)
    if 'n_gradient_iter' in res:
        print('   gradient iterations', res['n_gradient_iter'].tolist())
    print('   fval', np.round(res['fval'][:6], 3).tolist())
