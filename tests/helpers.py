"""Shared test scaffolding: build the init-time state and the CPU oracle for a config."""
import gzip
import io
from functools import lru_cache
from pathlib import Path

import numpy as np

from vega_b200 import VegaInterface
from vega_b200 import utils

REPO = Path(__file__).resolve().parents[1]


@lru_cache(maxsize=None)
def static_tables():
    """Numeric tables some model options need (Fvoigt profile, DESI instrument table)."""
    with gzip.open(utils.find_file('fvoigt_models/Fvoigt_exp.txt.gz'), 'rt') as f:
        fvoigt = np.loadtxt(io.StringIO(f.read()))
    csv = np.genfromtxt(utils.find_file(
        'instrumental_systematics/desi-instrument-syst-for-forest-auto-correlation.csv'),
        delimiter=',', names=True)
    return {'fvoigt': fvoigt, 'desi_inst': (np.asarray(csv['RT']), np.asarray(csv['XI']))}


def set_fast_metals(vega, on=True):
    """`fast_metals = True` under every [model] section (vega/metals.py:58), set before the
    oracle / the device plan read it."""
    for item in vega.corr_items.values():
        item.config['model']['fast_metals'] = str(bool(on))


def make_oracle(main_ini, omega_de_none=False, fast_metals=False, **kw):
    from oracle.vega_oracle import Oracle
    vega = VegaInterface(main_ini)
    if fast_metals:
        set_fast_metals(vega)
    if omega_de_none:
        vega.fiducial['Omega_de'] = None
    return vega, Oracle(vega, tables=static_tables(), **kw)


def make_c2(n_theta, width=3.0, main_ini='configs/c2/main.ini', fast_metals=False):
    """C2 workload: dense synthetic operands injected BEFORE the oracle is built (the oracle
    captures the metal matrices at construction), mock data drawn around the oracle model."""
    from oracle.vega_oracle import Oracle
    from vega_b200 import synthetic
    vega = VegaInterface(main_ini)
    if fast_metals:
        set_fast_metals(vega)
    synthetic.inject_dense_operands(vega)
    oracle = Oracle(vega, tables=static_tables())
    synthetic.draw_mock_data(vega, oracle.compute_model())
    return vega, oracle, synthetic.draw_theta(vega, n_theta, width=width)


# -- configurations derived from the reference's own ones (mirrors tests/golden/make_golden.py) --
MARG_CUTS = {'marginalize-below-rtmax': 16., 'marginalize-below-rpmax': 16.}
MARG_FIT = {'marginalize-all-rmin-cuts': True, 'marginalize-prior-sigma': 5.0}


def write_dmat_file(path, n_p=55, n_t=55, size=4.0, rp_min=0.):
    """Distortion-matrix file with its own, larger grid (vega/data.py:441-473)."""
    from vega_b200 import fitsio, synthetic
    rp = rp_min + (np.arange(n_p) + 0.5) * size
    rt = (np.arange(n_t) + 0.5) * size
    rp_grid, rt_grid = [g.ravel() for g in np.meshgrid(rp, rt, indexing='ij')]
    fitsio.write_tables(path, [
        ('DMAT', {'RPMIN': rp_min, 'RPMAX': rp_min + n_p * size, 'RTMAX': n_t * size, 'NP': n_p, 'NT': n_t,
                  'COEFMOD': 1}, {'DM': synthetic.distortion_matrix(n_p, n_t)}),
        ('ATTRI', {}, {'RP': rp_grid, 'RT': rt_grid, 'Z': np.full(rp_grid.size, 2.33)})])
    return str(path)


def modified_copy(main_ini, tmp_dir, model=None, control=None, data=None):
    """Copy of a configuration with extra [model] / [data] options in every correlation file and
    extra [control] options in the main file; returns the new main path."""
    import configparser
    tmp_dir = Path(tmp_dir)
    tmp_dir.mkdir(parents=True, exist_ok=True)

    def read(path):
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        cfg.read(utils.find_file(path))
        return cfg

    main = read(main_ini)
    new_paths = []
    for path in main['data sets']['ini files'].split():
        cfg = read(path)
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


def make_marg_case(tmp_path, key):
    """The two marginalisation cases of the reference goldens: 'c1_marg' (templates folded into the
    covariance) and 'c1_marg_fit' (marginalize-in-fit), both on the 55x55 distortion file."""
    from oracle.vega_oracle import Oracle
    dmat = write_dmat_file(Path(tmp_path) / 'dmat55.fits')
    if key == 'c1_marg':
        main = modified_copy('configs/configs/main.ini', Path(tmp_path) / key, model=MARG_CUTS,
                             data={'distortion-file': dmat})
    else:
        main = modified_copy('configs/configs/main.ini', Path(tmp_path) / key, model=MARG_FIT,
                             control={'marginalize-in-fit': True}, data={'distortion-file': dmat})
    vega = VegaInterface(main)
    return vega, Oracle(vega, tables=static_tables())


# ---------------------------------------------------------------------------------------------
# new_metals (vega/metals.py:83-108, 370-752): configurations with stacked-weight tables
# ---------------------------------------------------------------------------------------------
NEW_METALS_ALPHA = {'LYA': 2.9, 'LYB': 2.9, 'SiII(1190)': 1.0, 'SiII(1193)': 1.0, 'SiIII(1207)': 1.0,
                    'SiII(1260)': 1.0, 'CIV(eff)': 1.0}


def write_new_metals_inputs(tmp_dir, n_pix=520, n_qso=4000, seed=7):
    """Stacked forest weights (LOGLAM, WEIGHT) and an object catalogue (Z), seeded."""
    from vega_b200 import fitsio
    tmp_dir = Path(tmp_dir)
    tmp_dir.mkdir(parents=True, exist_ok=True)
    rng = np.random.default_rng(seed)
    step = 3e-4 if n_pix == 520 else 0.156 / n_pix             # 3600 .. ~5160 A whatever the sampling
    loglam = np.log10(3600.) + step * np.arange(n_pix)
    weight = (1 + 0.3 * np.sin(np.linspace(0, 7, n_pix))) * np.exp(-0.5 * ((loglam - 3.63) / 0.05)**2) \
        * (1 + 0.05 * rng.standard_normal(n_pix))
    forest = tmp_dir / 'forest_weights.fits'
    fitsio.write_tables(forest, [('STACK', {}, {'LOGLAM': loglam, 'WEIGHT': np.abs(weight)})])
    catalogue = tmp_dir / 'objects.fits'
    fitsio.write_tables(catalogue, [('CAT', {}, {'Z': 2.0 + 1.2 * rng.beta(2., 3., n_qso)})])
    return str(forest), str(catalogue)


def copy_data_with_cosmology(src, dst, omega_m=0.315):
    """Copy of a correlation-function file whose header carries the fiducial cosmology
    (vega/data.py:361-366 reads OMEGAM / OMEGAK / OMEGAR / WL)."""
    from vega_b200 import fitsio
    hdul = fitsio.open(utils.find_file(src))
    tables = []
    for i, hdu in enumerate(hdul[1:]):
        header = {k: hdu.header[k] for k in ('RPMIN', 'RPMAX', 'RTMAX', 'NP', 'NT', 'BLINDING')
                  if k in hdu.header}
        if i == 0:
            header.update({'OMEGAM': omega_m, 'OMEGAK': 0., 'OMEGAR': 7.9e-5, 'WL': -1.})
        tables.append((hdu.header.get('EXTNAME', f'HDU{i + 1}'), header,
                       {name: np.asarray(hdu.data[name]) for name in hdu.columns.names}))
    fitsio.write_tables(dst, tables)
    return str(dst)


def make_new_metals_config(tmp_dir, rp_only, base='configs/full_configs/main.ini',
                           corrs=('lyalya_lyalya', 'lyalya_qso'), rebin=2, new_bias_evolution=False,
                           model=None, dmat_bins=None):
    """The reference's own auto + cross configuration switched to new_metals: no metal file,
    matrices from the weight tables.  Returns the main ini path."""
    import configparser
    tmp_dir = Path(tmp_dir)
    forest, catalogue = write_new_metals_inputs(tmp_dir)

    def read(path):
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        cfg.read(utils.find_file(path))
        return cfg

    main = read(base)
    new_paths = []
    for path in main['data sets']['ini files'].split():
        if Path(path).stem not in corrs:
            continue
        cfg = read(path)
        cfg['model']['new_metals'] = 'True'
        cfg['model']['rp_only_metal_mats'] = str(bool(rp_only))
        if new_bias_evolution:       # vega/correlation_func.py:58,238-274 (needs the cosmology header)
            cfg['model']['new-bias-evolution'] = 'True'
            cfg['metals']['new-bias-evolution'] = 'True'
        data_file = tmp_dir / (Path(path).stem + '_data.fits')
        cfg['data']['filename'] = copy_data_with_cosmology(cfg['data']['filename'], data_file)
        cfg['data']['weights-tracer1'] = forest
        if cfg['data'].get('tracer2-type', 'continuous') == 'discrete':
            cfg['data']['weights-tracer2'] = catalogue
        if cfg['data'].get('tracer1-type', 'continuous') == 'discrete':
            cfg['data']['weights-tracer1'] = catalogue
            cfg['data']['weights-tracer2'] = forest
        for k, v in (model or {}).items():
            cfg['model'][k] = str(v)
        if dmat_bins is not None:    # distortion-matrix file with its own, larger model grid
            cross = cfg['data'].get('tracer2-type', 'continuous') == 'discrete'
            n_rp = 2 * dmat_bins if cross else dmat_bins
            dm_path = tmp_dir / f'dmat_{n_rp}x{dmat_bins}.fits'
            if not dm_path.exists():
                write_dmat_file(dm_path, n_rp, dmat_bins, 4.0, -4.0 * dmat_bins if cross else 0.)
            cfg['data']['distortion-file'] = str(dm_path)
        cfg['metals'].pop('filename', None)
        cfg['metal-matrix'] = {'rebin_factor': str(rebin), 'z_bins_objects': '300'}
        for name, alpha in NEW_METALS_ALPHA.items():
            cfg['metal-matrix'][f'alpha_{name}'] = str(alpha)
        new_path = tmp_dir / Path(path).name
        with open(new_path, 'w') as f:
            cfg.write(f)
        new_paths.append(str(new_path))
    main['data sets']['ini files'] = ' '.join(new_paths)
    main['output']['filename'] = str(tmp_dir / 'out')
    out = tmp_dir / 'main.ini'
    with open(out, 'w') as f:
        main.write(f)
    return str(out)


def oracle_metal_builder(corr_item, data):
    """The CPU oracle's new_metals matrices (oracle/metal_matrices.py) behind the Data hook."""
    from oracle.metal_matrices import MetalMatrixOracle
    from vega_b200 import fitsio
    from vega_b200.coordinates import Coordinates
    from vega_b200.cosmo import ABSORBER_IGM
    from vega_b200.metal_matrices import RpOnlyMatrix
    oracle = MetalMatrixOracle(corr_item, corr_item.cosmo, ABSORBER_IGM, fitsio.open)
    rp_only = corr_item.config['model'].getboolean('rp_only_metal_mats', False)
    for pair in corr_item.metal_correlations:
        mat, rp, rt, z = oracle.rp_matrix(*pair) if rp_only else oracle.full_matrix(*pair)
        data.metal_mats[pair] = RpOnlyMatrix(mat, corr_item.model_coordinates.rt_nbins) if rp_only else mat
        data.metal_coordinates[pair] = Coordinates.init_from_grids(corr_item.model_coordinates, rp, rt, z)


def make_c3_config(tmp_dir, dmat_bins=75):
    """C3: the DESI-DR2-shaped workload -- the reference's four correlations
    (lyaxlya, lyaxlyb, lyaxqso, lybxqso), data on 2500/5000 bins, distorted AND undistorted model
    on the distortion matrix's own grid (75 x 75 = 5625 / 150 x 75 = 11250 bins out to 300 Mpc/h),
    new_metals rp-only matrices on that grid, fast_metals, post-distortion broadband."""
    return make_new_metals_config(
        tmp_dir, True, corrs=('lyalya_lyalya', 'lyalya_lyalyb', 'lyalya_qso', 'lyalyb_qso'),
        model={'fast_metals': True}, dmat_bins=dmat_bins)


# ---------------------------------------------------------------------------------------------
# Option variants of the reference's auto + cross configuration.  tests/golden/make_golden.py runs
# the UNMODIFIED reference on each of them (-> reference_outputs.npz); the oracle and the CUDA path
# are compared with those outputs in tests/test_reference_golden.py.  Each one turns on model
# options no other configuration of tests/ reaches.
# ---------------------------------------------------------------------------------------------
_AUTO, _CROSS = 'lyalya_lyalya', 'lyalya_qso'
_METALS = ('SiII(1190)', 'SiII(1193)', 'SiIII(1207)', 'SiII(1260)', 'CIV(eff)')


def _separate_auto_bias_params():
    """bias_{M1}_{M2} for every metal x metal pair of the auto-correlation with M1 != M2
    (vega/metals.py:297-308), distinct values."""
    out, k = {}, 0
    for i, m1 in enumerate(_METALS):
        for m2 in _METALS[i + 1:]:
            out[f'bias_{m1}_{m2}'] = 0.6 + 0.07 * k
            k += 1
    return out


VARIANTS = {
    # vega/power_spectrum.py:224-261 (HeII branch), :190-194 + vega/utils.py:423-441 (k^n damping),
    # :281-288 (per-correlation HCD bias / beta)
    'v_pk': dict(
        sections={'model': {'HeII-reionization': True, 'pk-damping-scale': 0.9, 'pk-damping-power': 3},
                  'metals': {'pk-damping-scale': 1.3}},
        main={'parameters': {'lambda_HeII': 180., 'bias_gamma_e': 0.07, 'bias_hcd_LYAxLYA': -0.035,
                             'beta_hcd_LYAxQSO': 0.55}},
        cases=[{'bias_gamma_e': 0.11, 'lambda_HeII': 250., 'bias_prim': -0.5},
               {'bias_hcd_LYAxLYA': -0.06, 'beta_hcd_LYAxQSO': 0.8, 'bias_hcd': -0.03, 'beta_LYA': 1.4}]),
    # vega/correlation_func.py:597-686 (UV shot noise), :553-595 + vega/model.py:133-135 (DESI
    # instrumental systematics, auto-correlation only)
    'v_xi': dict(
        sections={'model': {'UVB-shotnoise': True}},
        per_corr={_AUTO: {'model': {'desi-instrumental-systematics': True}}},
        main={'parameters': {'uv_shotnoise_amp': 0.4, 'desi_inst_sys_amp': 2.5e-4}},
        cases=[{'uv_shotnoise_amp': 0.9, 'lambda_uv': 220., 'bias_gamma': 0.15},
               {'desi_inst_sys_amp': 5e-4, 'ap': 0.97, 'at': 1.04, 'drp_QSO': -0.4}]),
    # the same additive terms on the rescaled coordinates (vega/correlation_func.py:470-475,681-684)
    'v_xi_rescaled': dict(
        sections={'model': {'UVB-shotnoise': True, 'rescale-coords-systematics': True}},
        main={'parameters': {'uv_shotnoise_amp': 0.4},
              'cosmo-fit type': {'smooth-scaling': True, 'full-shape-alpha': True}},
        extra_main_params={'ap_full': 1.02, 'at_full': 0.99},
        cases=[{'ap_full': 0.96, 'at_full': 1.03, 'uv_shotnoise_amp': 0.7, 'qso_rad_strength': 0.9},
               {'drp_QSO': 1.1, 'qso_rad_asymmetry': 0.3, 'qso_rad_lifetime': 40.}]),
    # vega/broadband_poly.py:98-116: multiplicative terms before and after the distortion matrix,
    # next to additive ones in both positions and coordinate systems
    'v_bb': dict(
        replace={'broadband': {'bb1': 'mul pre r,mu 0:1:1 0:2:2', 'bb2': 'mul post rp,rt 0:1:1 0:1:1',
                               'bb3': 'add pre rp,rt 0:1:1 0:1:1', 'bb4': 'add post r,mu 0:2:1 0:6:2',
                               'bb5': 'mul pre r,mu 1:2:1 0:0:1'}},
        bb_values=True,
        cases=[{'BB-{corr}-0 mul pre r,mu (1,2)': 0.04, 'BB-{corr}-1 mul post rp,rt (1,1)': -0.03,
                'BB-{corr}-4 mul pre r,mu (2,0)': 0.02},
               {'BB-{corr}-2 add pre rp,rt (0,1)': 2e-4, 'BB-{corr}-3 add post r,mu (1,2)': 3e-3,
                'BB-{corr}-1 mul post rp,rt (0,0)': 0.05, 'bao_amp': 0.8}]),
    # vega/scale_parameters.py:38-230: phi_alpha with a rescaled smooth component
    'v_phi_alpha': dict(
        main={'cosmo-fit type': {'cosmo fit func': 'phi_alpha', 'smooth-scaling': True}},
        extra_main_params={'phi': 1.02, 'alpha': 0.99, 'phi_smooth': 0.97, 'alpha_smooth': 1.01},
        cases=[{'phi': 0.95, 'alpha': 1.03}, {'phi_smooth': 1.04, 'alpha_smooth': 0.96, 'bao_amp': 1.1}]),
    # phi_alpha in a split full-shape fit with one smooth alpha per correlation
    'v_phi_full': dict(
        main={'cosmo-fit type': {'cosmo fit func': 'phi_alpha', 'full-shape': True, 'two-alpha-smooth': True}},
        extra_main_params={'phi_full': 1.03, 'alpha': 0.98, 'alpha_smooth_LYAxLYA': 1.02,
                           'alpha_smooth_LYAxQSO': 0.97},
        cases=[{'phi_full': 0.96, 'alpha': 1.02}, {'alpha_smooth_LYAxLYA': 0.95, 'alpha_smooth_LYAxQSO': 1.05}]),
    # aiso_epsilon, full-shape with its own alpha pair, metals rescaled too (metal-scaling)
    'v_aiso': dict(
        main={'cosmo-fit type': {'cosmo fit func': 'aiso_epsilon', 'full-shape': True, 'full-shape-alpha': True,
                                 'metal-scaling': True}},
        extra_main_params={'aiso': 1.0, 'epsilon': 0.0, 'aiso_full': 1.01, 'epsilon_full': 0.012},
        cases=[{'aiso_full': 0.97, 'epsilon_full': -0.02}, {'aiso_full': 1.04, 'epsilon_full': 0.03,
                                                              'bias_eta_SiIII(1207)': -0.006}]),
    # aiso_epsilon on the peak only (the BAO branch of get_ap_at)
    'v_aiso_bao': dict(
        main={'cosmo-fit type': {'cosmo fit func': 'aiso_epsilon'}},
        extra_main_params={'aiso': 1.01, 'epsilon': 0.015},
        cases=[{'aiso': 0.96, 'epsilon': -0.03}]),
    # vega/metals.py:285-313: one beta for all metals, extra bias per metal x metal pair
    'v_metal_flags': dict(
        sections={'model': {'single-metal-beta': True, 'separate-metal-auto-biases': True}},
        extra_main_params={'beta_metals': 0.42, **_separate_auto_bias_params()},
        cases=[{'beta_metals': 0.7, 'bias_SiII(1190)_SiIII(1207)': 1.4},
               {'bias_eta_SiII(1260)': -0.004, 'bias_SiII(1193)_CIV(eff)': 0.2, 'beta_LYA': 1.5}]),
    # vega/model.py:119-131,178-186: the metal terms split into peak and smooth like the core
    'v_metal_decomp': dict(
        sections={'model': {'no-metal-decomp': False}},
        cases=[{'bao_amp': 0.7, 'bias_eta_SiIII(1207)': -0.006, 'ap': 1.03, 'at': 0.97},
               {'sigmaNL_par': 8., 'beta_SiII(1260)': 0.8, 'bias_eta_LYA': -0.19}]),
    # vega/model.py:106-107: compute_model returns P_ell(k) (no chi2 exists for this output)
    'v_model_pk': dict(
        main={'control': {'model_pk': True}}, pk_only=True,
        cases=[{'bao_amp': 0.8, 'bias_eta_LYA': -0.19, 'sigmaNL_per': 4.},
               {'beta_LYA': 1.4, 'bias_hcd': -0.03, 'L0_hcd': 8.}]),
}


def make_variant_config(tmp_dir, key, base='configs/full_configs/main.ini', corrs=(_AUTO, _CROSS)):
    """Writes the configuration files of VARIANTS[key] under tmp_dir; returns the main ini path."""
    import configparser
    spec = VARIANTS[key]
    tmp_dir = Path(tmp_dir)
    tmp_dir.mkdir(parents=True, exist_ok=True)

    def read(path):
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        cfg.read(utils.find_file(path))
        return cfg

    main = read(base)
    new_paths = []
    for path in main['data sets']['ini files'].split():
        stem = Path(path).stem
        if stem not in corrs:
            continue
        cfg = read(path)
        for section, options in spec.get('replace', {}).items():
            cfg[section] = {k: str(v) for k, v in options.items()}
        layers = [spec.get('sections', {}), spec.get('per_corr', {}).get(stem, {})]
        for layer in layers:
            for section, options in layer.items():
                for k, v in options.items():
                    cfg[section][k] = str(v)
        if spec.get('bb_values'):      # a value for every coefficient of the new broadband terms
            for name, value in broadband_parameters(cfg['broadband'], cfg['data']['name']).items():
                main['parameters'][name] = repr(value)
        new_path = tmp_dir / Path(path).name
        with open(new_path, 'w') as f:
            cfg.write(f)
        new_paths.append(str(new_path))
    main['data sets']['ini files'] = ' '.join(new_paths)
    for section, options in spec.get('main', {}).items():
        if section not in main:
            main[section] = {}
        for k, v in options.items():
            main[section][k] = str(v)
    for k, v in spec.get('extra_main_params', {}).items():
        main['parameters'][k] = repr(v)
    main['output']['filename'] = str(tmp_dir / 'out')
    out = tmp_dir / 'main.ini'
    with open(out, 'w') as f:
        main.write(f)
    return str(out)


def broadband_parameters(bb_section, corr_name):
    """Seeded values for every coefficient of a [broadband] section
    (names: vega/broadband_poly.py:62,182)."""
    rng = np.random.default_rng(11)
    out = {}
    for i, spec in enumerate(bb_section.values()):
        w = spec.split()
        lo1, hi1, s1 = (int(x) for x in w[3].split(':'))
        lo2, hi2, s2 = (int(x) for x in w[4].split(':'))
        scale = 0.02 if w[0] == 'mul' else 5e-4
        for a in range(lo1, hi1 + 1, s1):
            for b in range(lo2, hi2 + 1, s2):
                out[f'BB-{corr_name}-{i} {w[0]} {w[1]} {w[2]} ({a},{b})'] = float(
                    np.round(scale * rng.uniform(-1, 1), 6))
    return out


def variant_cases(key, corr_names):
    """Perturbed parameter dictionaries of a variant ('{corr}' in a name = every correlation)."""
    cases = []
    for case in VARIANTS[key]['cases']:
        out = {}
        for name, value in case.items():
            if '{corr}' in name:
                for corr in corr_names:
                    out[name.format(corr=corr)] = value
            else:
                out[name] = value
        cases.append(out)
    return cases


# ---------------------------------------------------------------------------------------------
# Blinded data (vega/data.py:305-340, vega/vega_interface.py:389-421,853-886, vega/utils.py:320-393)
# ---------------------------------------------------------------------------------------------
BLINDING_VALUES = {'growth_rate': 1.06}      # offset pi - exp(v^2) = +0.0657


def copy_data_blinded(src, dst, strategy='desi_dr3'):
    """Copy of a correlation file with a BLINDING card and a DA_BLIND column (DA shifted by a
    smooth, deterministic amount so that reading the wrong column cannot go unnoticed)."""
    from vega_b200 import fitsio
    hdul = fitsio.open(utils.find_file(src))
    tables = []
    for i, hdu in enumerate(hdul[1:]):
        header = {k: hdu.header[k] for k in ('RPMIN', 'RPMAX', 'RTMAX', 'NP', 'NT') if k in hdu.header}
        cols = {name: np.asarray(hdu.data[name]) for name in hdu.columns.names}
        if i == 0:
            header['BLINDING'] = strategy
            da = np.asarray(cols['DA'], dtype=float)
            cols['DA_BLIND'] = da + 1e-5 * np.cos(0.01 * np.arange(da.size))
        tables.append((hdu.header.get('EXTNAME', f'HDU{i + 1}'), header, cols))
    fitsio.write_tables(dst, tables)
    return str(dst)


def make_blind_config(tmp_dir, base='configs/full_configs/main.ini', corrs=(_AUTO, _CROSS)):
    """Auto + cross on blinded copies of the data files: growth_rate sampled (hence blinded),
    smooth component rescaled by ap_full / at_full, which the blinding pins to 1.  Returns
    (main ini path, blinding-offset file)."""
    import configparser
    tmp_dir = Path(tmp_dir)
    tmp_dir.mkdir(parents=True, exist_ok=True)

    def read(path):
        cfg = configparser.ConfigParser()
        cfg.optionxform = lambda option: option
        cfg.read(utils.find_file(path))
        return cfg

    main = read(base)
    new_paths = []
    for path in main['data sets']['ini files'].split():
        if Path(path).stem not in corrs:
            continue
        cfg = read(path)
        cfg['data']['filename'] = copy_data_blinded(cfg['data']['filename'],
                                                    tmp_dir / (Path(path).stem + '_blind.fits'))
        new_path = tmp_dir / Path(path).name
        with open(new_path, 'w') as f:
            cfg.write(f)
        new_paths.append(str(new_path))
    main['data sets']['ini files'] = ' '.join(new_paths)
    main['cosmo-fit type']['smooth-scaling'] = 'True'
    main['cosmo-fit type']['full-shape-alpha'] = 'True'
    main['parameters']['ap_full'] = '1.03'
    main['parameters']['at_full'] = '0.98'
    main['sample']['growth_rate'] = '0.5 1.5 0.96 0.1'
    if 'control' not in main:
        main['control'] = {}
    main['control']['use_template_growth_rate'] = 'False'
    main['output']['filename'] = str(tmp_dir / 'out')
    out = tmp_dir / 'main.ini'
    with open(out, 'w') as f:
        main.write(f)
    offsets = tmp_dir / 'blinding_offsets.npz'
    np.savez(offsets, **BLINDING_VALUES)
    return str(out), str(offsets)


BLIND_CASES = [{'growth_rate': 1.05, 'beta_LYA': 1.5}, {'ap_full': 1.2, 'at_full': 0.8, 'ap': 1.01},
               {'bias_eta_LYA': -0.19, 'growth_rate': 0.9}]


def make_blind_vega(tmp_dir):
    """The B200 interface on the blinded configuration, offsets from the synthetic file."""
    main, offsets = make_blind_config(tmp_dir)
    saved = dict(utils.BLINDING_FILES)
    utils.BLINDING_FILES['desi_dr3'] = {'full-shape': offsets, 'bao': None}
    try:
        return VegaInterface(main)
    finally:
        utils.BLINDING_FILES.clear()
        utils.BLINDING_FILES.update(saved)
