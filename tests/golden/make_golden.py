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


def modified_copy(main_ini, tmp_dir, model=None, control=None, data=None):
    """Copy of a reference configuration with extra options under every correlation's
    [model] / [data] section and under the main file's [control]."""
    import configparser
    from vega.utils import find_file
    tmp_dir.mkdir(parents=True, exist_ok=True)
    main = configparser.ConfigParser()
    main.optionxform = lambda option: option
    main.read(find_file(main_ini))
    new_paths = []
    for path in main['data sets']['ini files'].split():
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        cfg.read(find_file(os.path.expandvars(path)))
        for key, value in (model or {}).items():
            cfg['model'][key] = str(value)
        for key, value in (data or {}).items():
            cfg['data'][key] = str(value)
        new_path = tmp_dir / Path(path).name
        with open(new_path, 'w') as f:
            cfg.write(f)
        new_paths.append(str(new_path))
    main['data sets']['ini files'] = ' '.join(new_paths)
    if control:
        if 'control' not in main:
            main['control'] = {}
        for key, value in control.items():
            main['control'][key] = str(value)
    out = tmp_dir / 'main.ini'
    with open(out, 'w') as f:
        main.write(f)
    return str(out)


def fast_metals_copy(main_ini, tmp_dir):
    """`fast_metals = True` under every [model] (vega/metals.py:58: the metal caches of
    compute_xi_metal_metal / _cross_main)."""
    return modified_copy(main_ini, tmp_dir, model={'fast_metals': True})


# small-scale marginalisation (vega/data.py:96-128,762-828; vega/vega_interface.py:282-292,546-579)
MARG_CUTS = {'marginalize-below-rtmax': 16., 'marginalize-below-rpmax': 16.}
MARG_FIT = {'marginalize-all-rmin-cuts': True, 'marginalize-prior-sigma': 5.0}


def write_dmat_file(path, n_p=55, n_t=55, size=4.0):
    """Distortion-matrix file with its own (larger) grid, as tests/test_gpu_parity.py writes it."""
    from vega_b200 import fitsio, synthetic
    rp = (np.arange(n_p) + 0.5) * size
    rt = (np.arange(n_t) + 0.5) * size
    rp_grid, rt_grid = [g.ravel() for g in np.meshgrid(rp, rt, indexing='ij')]
    fitsio.write_tables(path, [
        ('DMAT', {'RPMIN': 0., 'RPMAX': n_p * size, 'RTMAX': n_t * size, 'NP': n_p, 'NT': n_t,
                  'COEFMOD': 1}, {'DM': synthetic.distortion_matrix(n_p, n_t)}),
        ('ATTRI', {}, {'RP': rp_grid, 'RT': rt_grid, 'Z': np.full(rp_grid.size, 2.33)})])
    return str(path)


def main():
    out, meta = {}, {}
    CONFIGS['full_fast'] = fast_metals_copy(CONFIGS['full'], Path('/tmp/vega_golden_fast'))
    # (the reference needs a distortion matrix from a file here: with the identity stand-in,
    # np.eye(...).dot(sparse templates) is not a sparse matrix and vega/data.py:817 fails)
    dmat = write_dmat_file(Path('/tmp/vega_golden_dmat55.fits'))
    CONFIGS['c1_marg'] = modified_copy(CONFIGS['c1'], Path('/tmp/vega_golden_marg'), model=MARG_CUTS,
                                       data={'distortion-file': dmat})
    CONFIGS['c1_marg_fit'] = modified_copy(CONFIGS['c1'], Path('/tmp/vega_golden_marg_fit'), model=MARG_FIT,
                                           control={'marginalize-in-fit': True},
                                           data={'distortion-file': dmat})
    PERTURBED['c1_marg'] = PERTURBED['c1_marg_fit'] = PERTURBED['c1']
    # new_metals (vega/metals.py:83-108,370-752): matrices from stacked weights, rp-only and
    # full 2-D form, on the reference's own auto + cross options; picca.constants is the
    # restatement in vega_b200/cosmo.py (tests/golden/shims/picca/constants.py)
    sys.path.insert(0, str(REPO / 'tests'))
    from helpers import make_new_metals_config
    base = str(REPO / 'tests' / 'configs' / 'full_configs' / 'main.ini')
    CONFIGS['new_metals_rp'] = make_new_metals_config('/tmp/vega_golden_nm_rp', True, base=base)
    CONFIGS['new_metals_2d'] = make_new_metals_config('/tmp/vega_golden_nm_2d', False, base=base)
    # new-bias-evolution (vega/correlation_func.py:58,238-274) on the same configuration
    CONFIGS['new_bias_evol'] = make_new_metals_config('/tmp/vega_golden_nbe', True, base=base,
                                                      new_bias_evolution=True)
    PERTURBED['new_metals_rp'] = PERTURBED['new_metals_2d'] = PERTURBED['full'][:2]
    PERTURBED['new_bias_evol'] = [PERTURBED['full'][0], PERTURBED['full'][2]]
    # Model.compute_direct through the direct_pk argument (vega/model.py:189-208)
    CONFIGS['full_direct'] = CONFIGS['full']
    PERTURBED['full_direct'] = PERTURBED['full']
    # option variants of the auto + cross configuration (tests/helpers.py: VARIANTS)
    from helpers import VARIANTS, make_variant_config, variant_cases
    for vkey, spec in VARIANTS.items():
        CONFIGS[vkey] = make_variant_config(f'/tmp/vega_golden_{vkey}', vkey, base=base)
        PERTURBED[vkey] = variant_cases(vkey, ('lyalya_lyalya', 'lyalya_qso'))
    # blinded data + parameter-level blinding (vega/vega_interface.py:389-421,853-886).  The one
    # thing replaced in the reference is WHERE get_blinding finds its offsets: its table of NERSC
    # files ships as None (vega/utils.py:331-341), so the reader below takes them from the
    # synthetic npz of helpers.make_blind_config with the reference's own per-parameter rule (:369)
    from helpers import make_blind_config, BLIND_CASES
    CONFIGS['blind'], blind_file = make_blind_config('/tmp/vega_golden_blind', base=base)
    PERTURBED['blind'] = BLIND_CASES
    import vega.utils as ref_utils

    def get_blinding_from_file(blind_pars, blinding_strat):
        assert blinding_strat is not None, 'Blinding failed, do not run!!!'
        with np.load(blind_file) as file:
            return {par: float(file[par]) for par in blind_pars}
    ref_utils.get_blinding = get_blinding_from_file
    notes = {'new_metals_rp': 'helpers.make_new_metals_config(rp_only=True): full_configs auto + cross, '
                              'new_metals = True, rp_only_metal_mats = True, seeded weight tables',
             'new_metals_2d': 'same with rp_only_metal_mats = False',
             'new_bias_evol': 'new_metals_rp + new-bias-evolution = True in [model] and [metals]',
             'full_direct': CONFIGS['full'] + ' with direct_pk = pk_full (1 + 0.05 sin(3 ln k))',
             'full_fast': CONFIGS['full'] + ' + fast_metals = True',
             'c1_marg': f"{CONFIGS['c1']} + [model] {MARG_CUTS} + distortion-file (55x55 bins)",
             'c1_marg_fit': f"{CONFIGS['c1']} + [model] {MARG_FIT} + [control] marginalize-in-fit "
                            "+ distortion-file (55x55 bins, synthetic.distortion_matrix)"}
    for vkey in VARIANTS:
        notes[vkey] = f'helpers.make_variant_config({vkey!r}): full_configs auto + cross with ' \
                      f'{ {k: v for k, v in VARIANTS[vkey].items() if k != "cases"} }'
    notes['blind'] = ('helpers.make_blind_config: auto + cross on data copies with BLINDING = desi_dr3 and a '
                      'DA_BLIND column, growth_rate sampled and blinded (offset file value 1.06), smooth '
                      'component scaled by ap_full / at_full (pinned to 1 by the blinding)')
    only = sys.argv[1:]
    if only:      # regenerate a subset, keep the rest of the stored outputs
        old = np.load(REPO / 'tests' / 'golden' / 'reference_outputs.npz')
        out.update({k: old[k] for k in old.files if k.split('/')[0] not in only})
        meta.update({k: v for k, v in json.loads(
            (REPO / 'tests' / 'golden' / 'reference_outputs.json').read_text()).items() if k not in only})
    for key, ini in CONFIGS.items():
        if only and key not in only:
            continue
        vega = VegaInterface(ini)
        if key == 'blind':
            assert vega._blind and vega._rnsps == {'growth_rate': 1.06}, (vega._blind, vega._rnsps)
        cases = [{}] + PERTURBED[key]
        meta[key] = {'ini': notes.get(key, ini), 'cases': cases, 'corr': list(vega.corr_items)}
        direct = None
        if key.endswith('_direct'):
            k = vega.fiducial['k']
            direct = vega.fiducial['pk_full'] * (1 + 0.05 * np.sin(3 * np.log(k)))
        for i, pars in enumerate(cases):
            model = vega.compute_model(pars, run_init=True, direct_pk=direct)
            for name, xi in model.items():
                out[f'{key}/{i}/xi/{name}'] = np.asarray(xi, dtype=float)
            if key in VARIANTS and VARIANTS[key].get('pk_only'):
                print(key, i, {name: xi.shape for name, xi in model.items()})
                continue
            out[f'{key}/{i}/chi2'] = np.array(vega.chi2(pars, direct_pk=direct))
            out[f'{key}/{i}/log_lik'] = np.array(vega.log_lik(pars, direct_pk=direct))
            if key.startswith('c1_marg'):
                chi2, coeff = vega.chi2(pars, return_marg_coeff=True)
                for name, c in coeff.items():
                    out[f'{key}/{i}/marg_coeff/{name}'] = np.asarray(c, dtype=float)
                model = vega.compute_model(pars, run_init=False, marg_coeff=coeff)
                for name, xi in model.items():
                    out[f'{key}/{i}/xi_marg/{name}'] = np.asarray(xi, dtype=float)
            print(key, i, float(out[f'{key}/{i}/chi2']), float(out[f'{key}/{i}/log_lik']))
        if key.startswith('new_metals') or key == 'new_bias_evol':
            # the matrices themselves: rp-only in full, the 2-D ones through three probe vectors
            # (a (Np Nt)^2 matrix is 50 MB), and the effective coordinate grids
            meta[key]['pairs'] = {}
            for name in vega.corr_items:
                metals = vega.models[name].metals
                pairs = list(vega.corr_items[name].metal_correlations)
                meta[key]['pairs'][name] = [list(p) for p in pairs]
                for c, pair in enumerate(pairs):
                    mat = metals.rp_metal_dmats[pair]
                    xi_c = metals.Xi_metal[pair]
                    out[f'{key}/coords/{name}/{c}'] = np.stack([xi_c._r, xi_c._mu, xi_c._z])
                    if not key.endswith('_2d'):
                        out[f'{key}/mat/{name}/{c}'] = np.asarray(mat, dtype=float)
                    else:
                        n = mat.shape[1]
                        probes = np.stack([np.cos(0.013 * (j + 1) * np.arange(n)) for j in range(3)], axis=1)
                        out[f'{key}/mat_probe/{name}/{c}'] = np.asarray(mat @ probes, dtype=float)
    np.savez_compressed(REPO / 'tests' / 'golden' / 'reference_outputs.npz', **out)
    (REPO / 'tests' / 'golden' / 'reference_outputs.json').write_text(json.dumps(meta, indent=1))


if __name__ == '__main__':
    main()
