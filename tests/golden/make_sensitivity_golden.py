"""Golden output of the UNMODIFIED reference's VegaInterface.compute_sensitivity
(vega/vega_interface.py:956-1071), generated HERE with the shims of make_golden.py:

    python tests/golden/make_sensitivity_golden.py

Output: tests/golden/reference_sensitivity.npz -- partials [2][2][Nm] per (correlation,
parameter), fisher [2][Nm] per (correlation, pair) and the nominal (value, error) pairs.
"""
import json
import os
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
REF = Path('/root/reference')
sys.path[:0] = [str(REPO / 'tests' / 'golden' / 'shims'), str(REF), str(REPO), str(REPO / 'tests' / 'golden')]
os.chdir(REF)

from vega import VegaInterface  # noqa: E402  (the reference)

NOMINAL = {'ap': (1.01, 0.03), 'at': (0.99, 0.04), 'bias_eta_LYA': (-0.2, 0.004),
           'beta_LYA': (1.6, 0.08), 'drp_QSO': (0.1, 0.15)}

if __name__ == '__main__':
    # save-components needs the metal biases applied inside P(k) (vega/metals.py:240-242)
    from make_golden import modified_copy
    ini = modified_copy('tests/full_configs/main.ini', Path('/tmp/vega_golden_sens'),
                        model={'fast_metal_bias': 'False'})
    vega = VegaInterface(str(ini))
    vega.compute_sensitivity(nominal=NOMINAL, frac=0.1, verbose=False)
    out = {}
    for name in vega.corr_items:
        for p, arr in vega.sensitivity['partials'][name].items():
            out[f'partials/{name}/{p}'] = arr
        for (p1, p2), arr in vega.sensitivity['fisher'][name].items():
            out[f'fisher/{name}/{p1}|{p2}'] = arr
    out['nominal'] = np.array(json.dumps(NOMINAL))
    np.savez_compressed(REPO / 'tests' / 'golden' / 'reference_sensitivity.npz', **out)
    print(len(out), 'arrays')
