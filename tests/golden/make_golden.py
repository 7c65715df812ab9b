"""Golden outputs of the UNMODIFIED reference, generated HERE (the reference cannot travel
to the GPU box).  Run from the repo root:

    python tests/golden/make_golden.py

It imports /root/reference/vega with the I/O shims of tests/golden/shims (astropy.io.fits ->
vega_b200.fitsio, mcfit -> the restated P2xi of oracle/hankel.py, stubs for iminuit, h5py,
git, matplotlib, getdist, picca) and stores, for the reference's own test configurations,
compute_model / chi2 / log_lik at the fiducial point and at a few perturbed parameter
dictionaries.  Output: tests/golden/reference_outputs.npz (+ the parameter dicts as JSON).
"""
import json
import os
import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parents[2]
REF = Path('/root/reference')
sys.path[:0] = [str(REPO / 'tests' / 'golden' / 'shims'), str(REF), str(REPO)]
os.chdir(REF)

from vega import VegaInterface  # noqa: E402  (the reference)

PERTURBED = {
    'full': [
        {'bias_eta_LYA': -0.21, 'beta_LYA': 1.5},
        {'ap': 0.98, 'at': 1.03, 'bao_amp': 0.9, 'bias_hcd': -0.04, 'beta_hcd': 0.6},
        {'drp_QSO': -0.3, 'sigma_velo_disp_lorentz_QSO': 5.0, 'L0_hcd': 8.0, 'alpha_LYA': 2.5,
         'BB-lyalya_lyalya-0 add post r,mu (1,2)': 0.003},
    ],
    'c1': [
        {'ap': 1.02, 'at': 0.97, 'bias_eta_LYA': -0.19, 'beta_LYA': 1.8},
        {'bias_hcd': -0.07, 'beta_hcd': 0.5, 'L0_hcd': 12.},
    ],
}
PERTURBED['full_fast'] = PERTURBED['full']
CONFIGS = {'full': 'tests/full_configs/main.ini', 'c1': 'tests/configs/main.ini'}


def fast_metals_copy(main_ini, tmp_dir):
    """Copy of a reference configuration with `fast_metals = True` under every [model]
    (vega/metals.py:58: the metal caches of compute_xi_metal_metal / _cross_main)."""
    import configparser
    tmp_dir.mkdir(parents=True, exist_ok=True)
    main = configparser.ConfigParser()
    main.optionxform = lambda option: option
    main.read(main_ini)
    new_paths = []
    for path in main['data sets']['ini files'].split():
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        from vega.utils import find_file
        cfg.read(find_file(os.path.expandvars(path)))
        cfg['model']['fast_metals'] = 'True'
        new_path = tmp_dir / Path(path).name
        with open(new_path, 'w') as f:
            cfg.write(f)
        new_paths.append(str(new_path))
    main['data sets']['ini files'] = ' '.join(new_paths)
    out = tmp_dir / 'main.ini'
    with open(out, 'w') as f:
        main.write(f)
    return str(out)


def main():
    out, meta = {}, {}
    CONFIGS['full_fast'] = fast_metals_copy(CONFIGS['full'], Path('/tmp/vega_golden_fast'))
    for key, ini in CONFIGS.items():
        vega = VegaInterface(ini)
        cases = [{}] + PERTURBED[key]
        meta[key] = {'ini': CONFIGS['full'] + ' + fast_metals = True' if key == 'full_fast' else ini,
                     'cases': cases, 'corr': list(vega.corr_items)}
        for i, pars in enumerate(cases):
            model = vega.compute_model(pars, run_init=True)
            for name, xi in model.items():
                out[f'{key}/{i}/xi/{name}'] = np.asarray(xi, dtype=float)
            out[f'{key}/{i}/chi2'] = np.array(vega.chi2(pars))
            out[f'{key}/{i}/log_lik'] = np.array(vega.log_lik(pars))
            print(key, i, float(out[f'{key}/{i}/chi2']), float(out[f'{key}/{i}/log_lik']))
    np.savez_compressed(REPO / 'tests' / 'golden' / 'reference_outputs.npz', **out)
    (REPO / 'tests' / 'golden' / 'reference_outputs.json').write_text(json.dumps(meta, indent=1))


if __name__ == '__main__':
    main()
