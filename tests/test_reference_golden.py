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
        ref_chi2 = float(gold[f'{key}/{i}/chi2'])
        assert abs(chi2(pars) - ref_chi2) <= min(1e-6, 1e-7 * abs(ref_chi2) + 1e-14)
        assert abs(log_lik(pars) - float(gold[f'{key}/{i}/log_lik'])) <= 1e-6


def check_marg(key, interface, gold, meta):
    """Template amplitudes and the model with the fitted templates added
    (vega/vega_interface.py:243-246,282-292,546-579)."""
    for i, pars in enumerate(meta[key]['cases']):
        value, coeff = interface.chi2(pars, return_marg_coeff=True)
        ref_chi2 = float(gold[f'{key}/{i}/chi2'])
        assert abs(value - ref_chi2) <= 1e-7 * abs(ref_chi2) + 1e-14
        model = interface.compute_model(pars, marg_coeff=coeff)
        for name in meta[key]['corr']:
            ref = gold[f'{key}/{i}/marg_coeff/{name}']
            assert np.max(np.abs(coeff[name] - ref)) <= 1e-8 * np.max(np.abs(ref)) + 1e-16, (key, i)
            ref = gold[f'{key}/{i}/xi_marg/{name}']
            assert np.max(np.abs(model[name] - ref)) <= 1e-10 * np.max(np.abs(ref)), (key, i)


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


@pytest.mark.parametrize('key', ['c1_marg', 'c1_marg_fit'])
def test_oracle_matches_reference_marginalisation(key, tmp_path):
    from helpers import make_marg_case
    gold, meta = load()
    vega, oracle = make_marg_case(tmp_path, key)
    check(key, oracle.compute_model, oracle.chi2, oracle.log_lik, gold, meta)
    for i, pars in enumerate(meta[key]['cases']):
        value, coeff = oracle.chi2(pars, return_marg_coeff=True)
        for name in meta[key]['corr']:
            ref = gold[f'{key}/{i}/marg_coeff/{name}']
            assert np.max(np.abs(coeff[name] - ref)) <= 1e-8 * np.max(np.abs(ref)) + 1e-16


@pytest.mark.gpu
@pytest.mark.parametrize('key', ['c1_marg', 'c1_marg_fit'])
def test_cuda_path_matches_reference_marginalisation(key, tmp_path):
    """Small-scale marginalisation (vega/data.py:96-128,762-828): default mode changes only the
    covariance; marginalize-in-fit reaches the device as the folded matrix Q^T C^-1 Q."""
    from helpers import make_marg_case
    gold, meta = load()
    vega, oracle = make_marg_case(tmp_path, key)
    check(key, vega.compute_model, vega.chi2, vega.log_lik, gold, meta)
    check_marg(key, vega, gold, meta)
