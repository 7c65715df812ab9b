"""Outputs of the UNMODIFIED reference (tests/golden/reference_outputs.npz, generated in
the build container by tests/golden/make_golden.py) against the CPU oracle (no GPU) and
against the CUDA path (gpu)."""
import json
from pathlib import Path

import numpy as np
import pytest

from helpers import make_oracle

GOLD = Path(__file__).resolve().parent / 'golden'
INI = {'full': 'configs/full_configs/main.ini', 'c1': 'configs/configs/main.ini',
       'full_fast': 'configs/full_configs/main.ini'}


def load():
    return np.load(GOLD / 'reference_outputs.npz'), json.loads((GOLD / 'reference_outputs.json').read_text())


def check(key, compute_model, chi2, log_lik, gold, meta):
    for i, pars in enumerate(meta[key]['cases']):
        model = compute_model(pars)
        for name in meta[key]['corr']:
            ref = gold[f'{key}/{i}/xi/{name}']
            err = np.max(np.abs(model[name] - ref)) / np.max(np.abs(ref))
            assert err <= 1e-10, (key, i, name, err)
        assert abs(chi2(pars) - float(gold[f'{key}/{i}/chi2'])) <= 1e-6
        assert abs(log_lik(pars) - float(gold[f'{key}/{i}/log_lik'])) <= 1e-6


@pytest.mark.parametrize('key', ['c1', 'full', 'full_fast'])
def test_oracle_matches_reference_outputs(key):
    gold, meta = load()
    vega, oracle = make_oracle(INI[key], fast_metals=key == 'full_fast')
    check(key, oracle.compute_model, oracle.chi2, oracle.log_lik, gold, meta)


@pytest.mark.gpu
@pytest.mark.parametrize('key', ['c1', 'full', 'full_fast'])
def test_cuda_path_matches_reference_outputs(key):
    """'full_fast': the reference's fast_metals caches (vega/metals.py:144-207) -- metal x metal
    terms frozen at the plan parameters, metal x main terms sharing one xi per cache key."""
    gold, meta = load()
    vega, oracle = make_oracle(INI[key], fast_metals=key == 'full_fast')
    check(key, vega.compute_model, vega.chi2, vega.log_lik, gold, meta)
