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


def make_oracle(main_ini, omega_de_none=False, **kw):
    from oracle.vega_oracle import Oracle
    vega = VegaInterface(main_ini)
    if omega_de_none:
        vega.fiducial['Omega_de'] = None
    return vega, Oracle(vega, tables=static_tables(), **kw)
