"""Where the split-integer GEMMs start to pay: chi2 pass time of small batches on the C2 and C3
workloads with VG_OPT_OZAKI_MIN_BATCH = 1 (every batch on tcgen05), = 100000 (every batch on the
FP64 DMMA kernel) and = 64 (the default), with and without the lanes of VG_OPT_LANES  (python tools/gpu_small_batches.py)."""
import json, sys, os
from pathlib import Path
import numpy as np
import torch
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO)); sys.path.insert(0, str(REPO / 'tests'))
os.chdir(REPO / 'tests')
import bench
from vega_b200 import _lib, synthetic


def timed(engine, tt, n=20):
    for _ in range(6):          # two ordinary passes, the capture, the first replays
        engine.chi2_device(tt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        c, s = engine.chi2_device(tt)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, c.cpu().numpy()


out = {}
for label in ('c2', 'c3'):
    if label == 'c2':
        vega = bench.build_state()
        names = None
    else:
        vega, names, _ = bench.build_c3(0)
    synthetic.draw_mock_data(vega, vega.compute_model())
    engine = vega._engine_for()
    rec = {}
    for B in (1, 8, 16, 29, 33, 48, 63, 64, 96, 128, 211, 256, 512):
        theta = synthetic.draw_theta(vega, B, width=1.0, names=names)
        tt = engine.theta_tensor(theta)
        row = {}
        for name, thr, lanes in (('tcgen05', 1, 0), ('dmma_one_lane', 100000, 0), ('dmma', 100000, 1),
                                 ('default_one_lane', 64, 0), ('default', 64, 1)):
            engine.set_option(_lib.OPT_OZAKI_MIN_BATCH, thr)
            engine.set_option(_lib.OPT_LANES, lanes)
            row[name + '_ms'], row[name] = timed(engine, tt)
        for k in ('dmma_one_lane', 'default_one_lane', 'default'):
            row.pop(k)
        row['max_rel_dchi2'] = float(np.max(np.abs(row.pop('tcgen05') - row['dmma']) / np.abs(row.pop('dmma'))))
        rec[B] = row
        print(label, B, row, flush=True)
    out[label] = rec
    del engine, vega
    torch.cuda.empty_cache()
print(json.dumps(out))
os.chdir(REPO)
Path('gpurun_out').mkdir(exist_ok=True)
json.dump(out, open('gpurun_out/small_batches.json', 'w'), indent=1)
