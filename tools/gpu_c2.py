"""GPU run of the C2 workload: parity on a few vectors + first throughput numbers."""
import sys
import time
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
from helpers import make_c2  # noqa: E402
from vega_b200 import synthetic, _lib  # noqa: E402
import ctypes as C  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
t0 = time.time()
vega, oracle, theta = make_c2(B)
print(f'synth {time.time() - t0:.1f}s')
names = list(vega.sample_params['limits'])
t0 = time.time()
chi2 = vega.chi2_batch(theta[:4])
print(f'plan+first chi2 {time.time() - t0:.1f}s', chi2)
for b in range(2):
    pars = dict(zip(names, theta[b]))
    t0 = time.time()
    ref = oracle.chi2(pars)
    print(f'oracle chi2 {ref:.8f} ({time.time() - t0:.2f}s) gpu {chi2[b]:.8f} diff {chi2[b] - ref:.2e}')
engine = vega._engine_for()
print('device bytes', engine.device_bytes() / 1e9, 'GB')
lib = _lib.load()
for rep in range(3):
    lib.vg_launch_count(1)
    torch.cuda.synchronize()
    t0 = time.time()
    c = vega.chi2_batch(theta)
    dt = time.time() - t0
    print(f'B={B}: e2e {dt * 1e3:.1f} ms  {B / dt:.0f} evals/s  launches {lib.vg_launch_count(0)}  '
          f'chi2 mean {c.mean():.2f} bad {(c > 1e99).sum()}')
tt = engine.theta_tensor(theta)
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    engine.chi2_device(tt)
    e1.record()
    torch.cuda.synchronize()
    print(f'device-resident: {e0.elapsed_time(e1):.1f} ms  {B / e0.elapsed_time(e1) * 1e3:.0f} evals/s')

ms = (C.c_double * 16)()
cnt = (C.c_int64 * 16)()
lib.vg_profile_enable(engine.handle, 1)
engine.chi2_device(tt)
lib.vg_profile_read(engine.handle, ms, cnt)
tags = ['pk_ell', 'hankel_gemm', 'xi_gather', 'metal_gemm', 'dm_gemm', 'chi2_gemm', 'elementwise']
print({t: (round(ms[i], 3), cnt[i]) for i, t in enumerate(tags)})
