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
        if f'{key}/{i}/chi2' not in gold:      # model_pk: multipoles only
            continue
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


def direct_pk_of(vega):
    """The isotropic P(k) tests/golden/make_golden.py passed as direct_pk."""
    k = vega.fiducial['k']
    return vega.fiducial['pk_full'] * (1 + 0.05 * np.sin(3 * np.log(k)))


def test_oracle_matches_reference_direct_pk():
    """Model.compute_direct (vega/model.py:189-208) through compute_model / chi2 / log_lik."""
    gold, meta = load()
    vega, oracle = make_oracle(INI['full'])
    pk = direct_pk_of(vega)
    check('full_direct', lambda p: oracle.compute_model(p, direct_pk=pk),
          lambda p: oracle.chi2(p, direct_pk=pk), lambda p: oracle.log_lik(p, direct_pk=pk), gold, meta)


@pytest.mark.gpu
def test_cuda_path_matches_reference_direct_pk():
    gold, meta = load()
    vega, oracle = make_oracle(INI['full'])
    pk = direct_pk_of(vega)
    check('full_direct', lambda p: vega.compute_model(p, direct_pk=pk),
          lambda p: vega.chi2(p, direct_pk=pk), lambda p: vega.log_lik(p, direct_pk=pk), gold, meta)
    # the option is scoped to the call: the usual decomposition is back afterwards
    check('full', vega.compute_model, vega.chi2, vega.log_lik, gold, meta)
    # batched entry point
    names = vega.theta_names
    theta = np.array([[vega.params[n] for n in names]] * 3)
    theta[1:, 0] *= [1.05, 0.9]
    got = vega.chi2_batch(theta, direct_pk=pk)
    for row, value in zip(theta, got):
        ref = oracle.chi2(dict(zip(names, row)), direct_pk=pk)
        assert abs(value - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4)


def check_new_metals_matrices(key, vega, gold, meta, rtol):
    """Metal matrices and effective coordinates of a new_metals plan against the reference's
    (vega/metals.py:502-752): the rp-only matrices entry by entry, the 2-D ones through the
    three probe vectors of make_golden.py."""
    for name, pairs in meta[key]['pairs'].items():
        data = vega.data[name]
        for c, pair in enumerate(pairs):
            coords, mat = data.get_metal(tuple(pair))
            ref = gold[f'{key}/coords/{name}/{c}']
            got = np.stack([coords.r_grid, coords.mu_grid, coords.z_grid])
            assert np.max(np.abs(got - ref)) <= rtol * np.max(np.abs(ref)), (name, pair)
            if not key.endswith('_2d'):
                ref = gold[f'{key}/mat/{name}/{c}']
                assert np.max(np.abs(mat.rp_mat - ref)) <= rtol * np.max(np.abs(ref)), (name, pair)
            else:
                ref = gold[f'{key}/mat_probe/{name}/{c}']
                n = mat.shape[1]
                probes = np.stack([np.cos(0.013 * (j + 1) * np.arange(n)) for j in range(3)], axis=1)
                assert np.max(np.abs(mat @ probes - ref)) <= rtol * np.max(np.abs(ref)), (name, pair)


def new_metals_config(key, tmp_path):
    from helpers import make_new_metals_config
    return make_new_metals_config(tmp_path, not key.endswith('_2d'),
                                  new_bias_evolution=key == 'new_bias_evol')


@pytest.mark.parametrize('key', ['new_metals_rp', 'new_metals_2d', 'new_bias_evol'])
def test_oracle_matches_reference_new_metals(key, tmp_path):
    """oracle/metal_matrices.py against the unmodified reference's compute_metal_rp_dmat /
    compute_metal_dmat, and the model built on those matrices; 'new_bias_evol' adds
    new-bias-evolution (vega/correlation_func.py:58,238-274) in [model] and [metals]."""
    from helpers import oracle_metal_builder, static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import VegaInterface
    gold, meta = load()
    vega = VegaInterface(new_metals_config(key, tmp_path), metal_builder=oracle_metal_builder)
    check_new_metals_matrices(key, vega, gold, meta, 1e-12)
    oracle = Oracle(vega, tables=static_tables())
    check(key, oracle.compute_model, oracle.chi2, oracle.log_lik, gold, meta)


@pytest.mark.gpu
@pytest.mark.parametrize('key', ['new_metals_rp', 'new_metals_2d', 'new_bias_evol'])
def test_cuda_path_matches_reference_new_metals(key, tmp_path):
    """Product path: pair histograms on the device (vg_metal_pair_histograms), the matrices
    applied by the hot path."""
    from vega_b200 import VegaInterface
    gold, meta = load()
    vega = VegaInterface(new_metals_config(key, tmp_path))
    check_new_metals_matrices(key, vega, gold, meta, 1e-11)
    check(key, vega.compute_model, vega.chi2, vega.log_lik, gold, meta)


def variant_keys():
    from helpers import VARIANTS
    return list(VARIANTS)


@pytest.mark.parametrize('key', variant_keys())
def test_oracle_matches_reference_option_variants(key, tmp_path):
    """Model options no other configuration reaches (tests/helpers.py: VARIANTS), each against the
    unmodified reference: HeII, k^n damping, per-correlation HCD bias, UV shot noise, DESI
    instrumental systematics, rescale-coords-systematics, multiplicative broadband (pre and post),
    phi_alpha / aiso_epsilon with smooth-scaling / full-shape / two-alpha-smooth / metal-scaling,
    single-metal-beta, separate-metal-auto-biases."""
    from helpers import make_variant_config, static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import VegaInterface
    gold, meta = load()
    vega = VegaInterface(make_variant_config(tmp_path, key))
    oracle = Oracle(vega, tables=static_tables())
    check(key, oracle.compute_model, oracle.chi2, oracle.log_lik, gold, meta)


@pytest.mark.gpu
@pytest.mark.parametrize('key', variant_keys())
def test_cuda_path_matches_reference_option_variants(key, tmp_path):
    """The same variants through the C ABI: scalar calls against the reference outputs, then the
    perturbed cases as ONE batch (sampled columns = every perturbed parameter) against the same."""
    from helpers import make_variant_config
    from vega_b200 import VegaInterface
    gold, meta = load()
    vega = VegaInterface(make_variant_config(tmp_path, key))
    check(key, vega.compute_model, vega.chi2, vega.log_lik, gold, meta)
    if f'{key}/0/chi2' not in gold:
        return
    cases = meta[key]['cases']
    names = sorted({n for case in cases for n in case})
    theta = np.array([[case.get(n, vega.params[n]) for n in names] for case in cases])
    models, status = vega.compute_model_batch(theta, names)
    assert not status.any()
    chi2 = vega.chi2_batch(theta, names)
    for i in range(len(cases)):
        for name in meta[key]['corr']:
            ref = gold[f'{key}/{i}/xi/{name}']
            assert np.max(np.abs(models[name][i] - ref)) <= 1e-10 * np.max(np.abs(ref)), (key, i, name)
        assert abs(chi2[i] - float(gold[f'{key}/{i}/chi2'])) <= 1e-6


def test_oracle_matches_reference_blinding(tmp_path):
    """Blinded data: DA_BLIND column, growth_rate + (pi - exp(v^2)), ap_full / at_full pinned to 1
    (vega/data.py:305-340, vega/vega_interface.py:389-421, vega/utils.py:375-393)."""
    from helpers import make_blind_vega, static_tables
    from oracle.vega_oracle import Oracle
    gold, meta = load()
    vega = make_blind_vega(tmp_path)
    assert vega._blind and vega._rnsps == {'growth_rate': 1.06}
    local = vega._get_lcl_prms({'ap_full': 1.2})
    assert local['ap_full'] == 1. and local['at_full'] == 1.
    assert abs(local['growth_rate'] - (vega.params['growth_rate'] + np.pi - np.exp(1.06**2))) < 1e-15
    oracle = Oracle(vega, tables=static_tables())
    check('blind', oracle.compute_model, oracle.chi2, oracle.log_lik, gold, meta)


def test_blinding_guards(tmp_path):
    """vega/vega_interface.py:853-886 and vega/utils.py:320-349: a BLIND_FIXED parameter may not be
    sampled, bias_QSO and beta_QSO may not both be sampled, and without an offsets file for the
    strategy the blinded parameters cannot be sampled at all."""
    import configparser
    from helpers import make_blind_config
    from vega_b200 import VegaInterface, utils
    main, offsets = make_blind_config(tmp_path)
    with pytest.raises(ValueError, match='Unknown blinding version'):
        VegaInterface(main)                      # no file registered for desi_dr3 (as shipped upstream)
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda option: option
    cfg.read(main)
    cfg['sample']['ap_full'] = '0.5 1.5 1.0 0.1'
    with open(main, 'w') as f:
        cfg.write(f)
    utils.BLINDING_FILES['desi_dr3'] = {'full-shape': offsets, 'bao': None}
    try:
        with pytest.raises(ValueError, match='must be fixed'):
            VegaInterface(main)
    finally:
        del utils.BLINDING_FILES['desi_dr3']


@pytest.mark.gpu
def test_cuda_path_matches_reference_blinding(tmp_path):
    """The blinding transform on the device (vg_plan_set_theta_transform): scalar calls, then a
    batch whose columns include the blinded growth_rate and the pinned ap_full / at_full."""
    from helpers import make_blind_vega
    gold, meta = load()
    vega = make_blind_vega(tmp_path)
    check('blind', vega.compute_model, vega.chi2, vega.log_lik, gold, meta)
    cases = meta['blind']['cases']
    names = sorted({n for case in cases for n in case})
    theta = np.array([[case.get(n, vega.params[n]) for n in names] for case in cases])
    models, status = vega.compute_model_batch(theta, names)
    assert not status.any()
    chi2 = vega.chi2_batch(theta, names)
    for i in range(len(cases)):
        for name in meta['blind']['corr']:
            ref = gold[f'blind/{i}/xi/{name}']
            assert np.max(np.abs(models[name][i] - ref)) <= 1e-10 * np.max(np.abs(ref)), (i, name)
        assert abs(chi2[i] - float(gold[f'blind/{i}/chi2'])) <= 1e-6


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


@pytest.mark.gpu
def test_cuda_path_matches_reference_sensitivity():
    """VegaInterface.compute_sensitivity (vega/vega_interface.py:956-1071) against the unmodified
    reference's output (tests/golden/make_sensitivity_golden.py): partial derivatives of the
    distorted / undistorted peak and smooth components and the per-bin Fisher information.
    The finite difference divides the 1e-10 model tolerance by a step of 0.2-1 % of the
    parameter, hence 1e-6 * max|partial| here."""
    from vega_b200 import VegaInterface
    gold = np.load(GOLD / 'reference_sensitivity.npz')
    nominal = {k: tuple(v) for k, v in json.loads(str(gold['nominal'])).items()}
    vega = VegaInterface(INI['full'])
    sens = vega.compute_sensitivity(nominal=nominal, frac=0.1, verbose=False)
    assert list(sens['nominal']) == list(nominal)
    n_checked = 0
    for key in gold.files:
        if key == 'nominal':
            continue
        kind, name, par = key.split('/')
        ref = gold[key]
        got = sens[kind][name][par if kind == 'partials' else tuple(par.split('|'))]
        assert got.shape == ref.shape, key
        assert np.array_equal(np.isnan(got), np.isnan(ref)), key
        if kind == 'partials':
            for d in range(2):
                for c in range(2):
                    scale = np.max(np.abs(ref[d, c]))
                    assert np.max(np.abs(got[d, c] - ref[d, c])) <= 1e-6 * scale, (key, d, c)
        else:
            scale = np.nanmax(np.abs(ref), axis=1, keepdims=True)
            ok = ~np.isnan(ref)
            assert np.all((np.abs(got - ref) <= 1e-5 * scale + 0 * ref)[ok]), key
        n_checked += 1
    assert n_checked == 80
