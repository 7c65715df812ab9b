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
