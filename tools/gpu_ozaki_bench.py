"""C2 step with the metal GEMM as DMMA (VG_OPT_OZAKI = 0) and as the split-integer tcgen05 GEMM with
S = 6, 7, 8 digits: ms per step, per-stage ms, and the chi2 difference to the DMMA path."""
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / 'tests'))
import torch  # noqa: E402
from bench import build_state, STAGES  # noqa: E402
from vega_b200 import synthetic, _lib  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
vega = build_state()
theta = synthetic.draw_theta(vega, B)
synthetic.draw_mock_data(vega, vega.compute_model())
engine = vega._engine_for()
lib = _lib.load()
tt = engine.theta_tensor(theta)
chi2 = torch.empty(B, dtype=torch.float64, device='cuda')
status = torch.empty(B, dtype=torch.int32, device='cuda')
out = {'batch': B}
ref = None
ONLY = [int(x) for x in sys.argv[2:]]
STEPS = 10 if not ONLY else 2
for slices in (ONLY or (0, -1, 6, 7, 8)):
    engine.set_option(_lib.OPT_OZAKI, slices)
    for _ in range(3):
        engine.chi2_device(tt, chi2, status)
    torch.cuda.synchronize()
    lib.vg_profile_enable(engine.handle, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(STEPS):
        engine.chi2_device(tt, chi2, status)
    e1.record()
    torch.cuda.synchronize()
    ms = (C.c_double * 16)()
    cnt = (C.c_int64 * 16)()
    lib.vg_profile_read(engine.handle, ms, cnt)
    lib.vg_profile_enable(engine.handle, 0)
    vals = chi2.cpu().numpy().copy()
    if ref is None:
        ref = vals
    out[f'S={slices}'] = {'digits': [engine.ozaki_slices(ci) for ci in range(len(engine.infos))],
                          'ms_per_step': e0.elapsed_time(e1) / STEPS, 'evals_per_s': B / (e0.elapsed_time(e1) / STEPS * 1e-3),
                          'stages_ms': {s: ms[i] / STEPS for i, s in enumerate(STAGES)},
                          'max_abs_dchi2': float(np.max(np.abs(vals - ref))), 'max_chi2': float(np.max(np.abs(ref))),
                          'status_any': int(status.max().item())}
print(json.dumps(out, indent=1))
(REPO / 'gpurun_out' / f'ozaki_bench_B{B}.json').write_text(json.dumps(out, indent=1))
