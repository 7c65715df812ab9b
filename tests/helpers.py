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


def write_dmat_file(path, n_p=55, n_t=55, size=4.0):
    """Distortion-matrix file with its own, larger grid (vega/data.py:441-473)."""
    from vega_b200 import fitsio, synthetic
    rp = (np.arange(n_p) + 0.5) * size
    rt = (np.arange(n_t) + 0.5) * size
    rp_grid, rt_grid = [g.ravel() for g in np.meshgrid(rp, rt, indexing='ij')]
    fitsio.write_tables(path, [
        ('DMAT', {'RPMIN': 0., 'RPMAX': n_p * size, 'RTMAX': n_t * size, 'NP': n_p, 'NT': n_t,
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
