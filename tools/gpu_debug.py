"""GPU debug run: model parity per correlation against the CPU oracle."""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
sys.path.insert(0, str(Path(__file__).resolve().parents[1] / 'tests'))
from helpers import make_oracle  # noqa: E402

configs = sys.argv[1:] or ['configs/configs/main.ini', 'configs/picca_bench/main.ini',
                           'configs/picca_bench/main_cross.ini', 'configs/full_configs/main.ini']
for main in configs:
    legacy = 'picca' in main
    vega, oracle = make_oracle(main, omega_de_none=legacy)
    t0 = time.time()
    ref = oracle.compute_model()
    t1 = time.time()
    got = vega.compute_model()
    t2 = time.time()
    print(f'== {main}: oracle {t1 - t0:.2f}s gpu(first, incl. plan) {t2 - t1:.2f}s')
    for name in ref:
        err = np.max(np.abs(got[name] - ref[name])) / np.max(np.abs(ref[name]))
        print(f'   {name:24s} max|d|/max|xi| = {err:.3e}  nan={np.isnan(got[name]).sum()}')
    if vega._has_data:
        c_ref, c_got = oracle.chi2(), vega.chi2()
        print(f'   chi2 oracle {c_ref:.10f} gpu {c_got:.10f} diff {c_got - c_ref:.3e}')
