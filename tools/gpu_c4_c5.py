"""C4 and C5 at their stated sizes, one process per GPU (run under torchrun or alone):
  C4  PolyChord-style batched log-likelihood: 1e5 parameter vectors sharded over the ranks,
      chi2 values all-gathered (sharding.ShardedChi2);
  C5  Monte Carlo mock fits: 1000 mocks fitted in lock-step, the mocks sharded over the ranks
      (drivers.MonteCarloFits.fit_sharded).
Usage: gpu_c4_c5.py [n_vectors] [n_mocks] [c2|c3]
  c2 (default)  the bench's C2-shaped workload (2 correlations, dense metal matrices)
  c3            the DESI-DR2-shaped workload of SURVEY 8d (4 correlations, 5625/11250-bin model
                grids, rp-only new_metals matrices, fast_metals, global covariance over 9540
                masked bins; mocks drawn from the joint covariance)"""
import json
import os
import sys
import time
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
import bench  # noqa: E402
from vega_b200 import synthetic  # noqa: E402
from vega_b200.drivers import MonteCarloFits  # noqa: E402
from vega_b200.sharding import ShardedChi2  # noqa: E402

world = int(os.environ.get('WORLD_SIZE', '1'))
rank = int(os.environ.get('RANK', '0'))
local = int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
n_vectors = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
n_mocks = int(sys.argv[2]) if len(sys.argv) > 2 else 1000

model = sys.argv[3] if len(sys.argv) > 3 else 'c2'
t_init = time.perf_counter()
if model == 'c3':
    from helpers import make_c3_config
    from vega_b200 import VegaInterface
    vega = VegaInterface(make_c3_config(f'/tmp/c3_c4c5_{rank}'))
    synthetic.inject_dense_operands(vega, with_distortion=False, with_metals=False)
    synthetic.inject_global_covariance(vega)
    names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'drp_QSO',
             'sigma_velo_disp_lorentz_QSO', 'bias_eta_SiII(1260)', 'bias_eta_SiIII(1207)',
             'bias_eta_SiII(1193)', 'bias_eta_SiII(1190)', 'bias_eta_CIV(eff)', 'bias_QSO']
    names = [n for n in names if n in vega.params]
    # the [sample] set of the fits: restrict / extend the configuration's to `names`
    sp = vega.sample_params
    from vega_b200 import utils as vutils
    for n in names:
        if n not in sp['limits']:
            sp['values'][n] = vega.params[n]
            sp['errors'][n] = 0.02 * abs(vega.params[n]) + 1e-4
            sp['limits'][n] = (None, None)
            sp['fix'][n] = False
    for key in ('values', 'errors', 'limits', 'fix'):
        sp[key] = {n: sp[key][n] for n in names}
else:
    vega = bench.build_state()
fiducial = vega.compute_model()
synthetic.draw_mock_data(vega, fiducial)
out = {'world': world, 'model': model, 'init_s': time.perf_counter() - t_init}

# ---- C4 ----
theta = synthetic.draw_theta(vega, n_vectors)
sharded = ShardedChi2.from_vega(vega)
sharded.chi2(theta[:4096])
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
chi2 = sharded.chi2(theta)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
dt = time.perf_counter() - t0
out['c4'] = {'vectors': n_vectors, 'wall_s': dt, 'evals_per_s': n_vectors / dt,
             'checksum': float(np.sum(chi2[chi2 < 1e99])), 'n_sentinel': int(np.sum(chi2 >= 1e99))}
if rank == 0:
    print('C4', out['c4'], flush=True)

# ---- C5 ----
mc = MonteCarloFits(vega, fiducial_model=fiducial)
mc.draw(n_mocks, seed=1)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
res = mc.fit_sharded(max_iter=30)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
dt = time.perf_counter() - t0
theta_fit = res['theta']
out['c5'] = {'mocks': n_mocks, 'parameters': len(mc.names), 'wall_s': dt, 'fits_per_s': n_mocks / dt,
             'converged': int(np.sum(res['converged'])),
             'mean_fval': float(np.mean(res['fval'])),
             'mean_theta': {n: float(v) for n, v in zip(mc.names, theta_fit.mean(axis=0))},
             'std_theta': {n: float(v) for n, v in zip(mc.names, theta_fit.std(axis=0))}}
if rank == 0:
    print('C5', {k: out['c5'][k] for k in ('mocks', 'parameters', 'wall_s', 'fits_per_s', 'converged', 'mean_fval')},
          flush=True)
    Path('gpurun_out').mkdir(exist_ok=True)
    json.dump(out, open(f'gpurun_out/c4_c5_{model}_n{world}.json', 'w'), indent=1)
if world > 1:
    dist.destroy_process_group()
