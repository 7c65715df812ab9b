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
