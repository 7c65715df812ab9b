"""Small C2 run for ncu: B vectors, one chi2 pass (after one warm-up pass)."""
import sys
from pathlib import Path
import ctypes as C
import torch
sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
import bench  # noqa: E402
from vega_b200 import synthetic, _lib  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
vega = bench.build_state()
synthetic.draw_mock_data(vega, vega.compute_model())
theta = synthetic.draw_theta(vega, B)
engine = vega._engine_for()
tt = engine.theta_tensor(theta)
for _ in range(2):
    chi2, status = engine.chi2_device(tt)
torch.cuda.synchronize()
# one more pass inside a cudaProfilerStart/Stop range: `ncu --profile-from-start off` sees only it
torch.cuda.profiler.start()
chi2, status = engine.chi2_device(tt)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
lib = _lib.load()
tf = C.c_double()
scratch = torch.zeros(148 * 8 * 256, dtype=torch.float64, device='cuda')
lib.vg_probe_dfma(148 * 8, 20000, C.c_void_p(scratch.data_ptr()), C.byref(tf), None)
print('chi2 mean', float(chi2.mean()), 'DFMA probe TFLOP/s', tf.value)
