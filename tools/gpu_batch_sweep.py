"""chi2 pass time of the C2 workload as a function of the batch size (stencil batches of the
minimizer are ~35-600 vectors, sampler batches 1e3-1e5)."""
import json
import sys
import time
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
import bench  # noqa: E402
from vega_b200 import synthetic  # noqa: E402

vega = bench.build_state()
synthetic.draw_mock_data(vega, vega.compute_model())
engine = vega._engine_for()
out = {}
for B in [1, 8, 33, 64, 128, 256, 513, 1024, 2048, 4096]:
    theta = synthetic.draw_theta(vega, B)
    tt = engine.theta_tensor(theta)
    for _ in range(6):      # two ordinary passes, the graph capture, its first replays
        engine.chi2_device(tt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    e0.record()
    for _ in range(n):
        engine.chi2_device(tt)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    t0 = time.perf_counter()
    for _ in range(n):
        vega.chi2_batch(theta)
    host_ms = (time.perf_counter() - t0) * 1e3 / n
    out[B] = {'device_ms': ms, 'host_api_ms': host_ms, 'evals_per_s': B / (ms * 1e-3)}
    if B in (1, 33, 128, 513):       # per-stage events (they serialise the stages a little)
        import ctypes as C
        lib = engine.lib
        lib.vg_profile_enable(engine.handle, 1)
        lib.vg_launch_count(1)
        for _ in range(n):
            engine.chi2_device(tt)
        torch.cuda.synchronize()
        ms_stage, cnt_stage = (C.c_double * 16)(), (C.c_int64 * 16)()
        lib.vg_profile_read(engine.handle, ms_stage, cnt_stage)
        lib.vg_profile_enable(engine.handle, 0)
        out[B]['stages_ms'] = {s_: round(ms_stage[i] / n, 4) for i, s_ in enumerate(bench.STAGES)}
        out[B]['launches'] = lib.vg_launch_count(0) / n
    print(B, out[B], flush=True)
print(json.dumps(out))
Path('gpurun_out').mkdir(exist_ok=True)
json.dump(out, open('gpurun_out/batch_sweep.json', 'w'), indent=1)
