"""new_metals init: pair histograms on the device (vega_b200/metal_matrices.py) against the numpy
oracle (oracle/metal_matrices.py) on a DESI-sized stacked-weight table, and the hot path with
rp-only matrices against the dense 2-D ones on the same configuration."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
import helpers  # noqa: E402
from oracle.metal_matrices import MetalMatrixOracle  # noqa: E402
from vega_b200 import VegaInterface, fitsio, synthetic  # noqa: E402
from vega_b200.cosmo import ABSORBER_IGM  # noqa: E402
from vega_b200.metal_matrices import MetalMatrixBuilder  # noqa: E402

out = {}
n_pix = int(sys.argv[1]) if len(sys.argv) > 1 else 4000
helpers_write = helpers.write_new_metals_inputs


def big_inputs(tmp_dir, **kw):
    return helpers_write(tmp_dir, n_pix=n_pix, n_qso=200000)


helpers.write_new_metals_inputs = big_inputs
for rp_only in (True, False):
    key = 'rp_only' if rp_only else 'full_2d'
    ini = helpers.make_new_metals_config(f'/tmp/nm_bench_{key}', rp_only, rebin=1)
    t0 = time.perf_counter()
    vega = VegaInterface(ini)
    torch.cuda.synchronize()
    out[key] = {'interface_init_s': time.perf_counter() - t0}
    item = vega.corr_items['lyalya_lyalya']
    pair = item.metal_correlations[0]
    builder = MetalMatrixBuilder(item)
    builder.matrix(*pair)
    t0 = time.perf_counter()
    for _ in range(3):
        mat, _ = builder.matrix(*pair)
    out[key]['device_builder_s_per_pair'] = (time.perf_counter() - t0) / 3
    oracle = MetalMatrixOracle(item, item.cosmo, ABSORBER_IGM, fitsio.open)
    t0 = time.perf_counter()
    ref = (oracle.rp_matrix if rp_only else oracle.full_matrix)(*pair)[0]
    out[key]['numpy_oracle_s_per_pair'] = time.perf_counter() - t0
    got = mat.rp_mat if rp_only else mat
    out[key]['max_rel_diff'] = float(np.max(np.abs(got - ref)) / np.max(np.abs(ref)))
    out[key]['pairs_per_matrix'] = int(builder._side(0, pair[0]).shape[1] * builder._side(1, pair[1]).shape[1])
    # hot path
    synthetic.draw_mock_data(vega, vega.compute_model())
    engine = vega._engine_for()
    theta = synthetic.draw_theta(vega, 1024)
    tt = engine.theta_tensor(theta)
    for _ in range(3):
        engine.chi2_device(tt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        engine.chi2_device(tt)
    e1.record()
    torch.cuda.synchronize()
    out[key]['chi2_pass_ms_B1024'] = e0.elapsed_time(e1) / 5
    print(key, out[key], flush=True)
Path('gpurun_out').mkdir(exist_ok=True)
json.dump(out, open('gpurun_out/new_metals.json', 'w'), indent=1)
