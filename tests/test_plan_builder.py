"""Host-side plan logic (no GPU): descriptor flags / parameter columns produced from the
reference's ini options, the (ap, at) rule selection, static-vs-sampled factor split."""
import configparser

import pytest

from vega_b200 import VegaInterface, _lib
from vega_b200.plan import PlanBuilder, resolve_scale


def builder(main, sampled=None):
    vega = VegaInterface(main)
    return vega, PlanBuilder(vega, sampled=sampled)


def test_full_config_descriptors():
    vega, pb = builder('configs/full_configs/main.ini')
    info = pb.describe_corr('lyalya_lyalya')
    assert len(info['evals']) == 2 + 15 + 1
    peak, smooth = info['evals'][:2]
    # last: the direct_pk evaluation (vega/model.py:189-208) = the smooth options on P(k) slot 3
    direct = info['evals'][-1]
    assert direct.direct == 1 and direct.pk_lin_sel == 3 and direct.out_slot == 1
    assert direct.static_mask == smooth.static_mask and direct.scale_form == smooth.scale_form
    assert sum(d.direct for d in info['evals']) == 1
    assert peak.is_peak == 1 and peak.pk_lin_sel == 0 and peak.out_slot == 0
    assert smooth.is_peak == 0 and smooth.pk_lin_sel == 1 and smooth.out_slot == 1
    # bias_eta_LYA and beta_LYA are sampled: Kaiser is per vector, every other factor static
    assert peak.dyn_mask == 0
    assert peak.static_mask == _lib.F_ARINYO | _lib.F_GK | _lib.F_PEAKNL | _lib.F_FSGAUSS
    assert smooth.static_mask == _lib.F_ARINYO | _lib.F_GK | _lib.F_FSGAUSS
    assert peak.hcd_model == 1 and peak.hcd_grid >= 0 and peak.uvb == 1
    assert peak.scale_form == 1 and smooth.scale_form == 0      # BAO fit: only the peak is rescaled
    assert peak.idx_scale[0] == pb.col('ap') and peak.idx_scale[1] == pb.col('at')
    # metals: no HCD (their PowerSpectrum is built from the [metals] section), bias weights
    metal = info['evals'][2]
    assert metal.hcd_model == 0 and metal.fast_metals == 1 and metal.weight_bias == 1
    assert metal.scale_form == 0 and metal.pk_lin_sel == 2
    mixed = [d for d, pair in zip(info['evals'][2:], info['item'].metal_correlations) if pair[0] != pair[1]]
    assert all(d.weight_const == 2.0 for d in mixed)            # vega/metals.py:238-239


def test_static_xi_flag_follows_the_sampled_set():
    """vg_eval_desc.static_xi: set only when nothing the xi stage reads is sampled."""
    vega, pb = builder('configs/full_configs/main.ini', sampled=['bias_eta_LYA', 'beta_LYA', 'ap', 'at'])
    evals = pb.describe_corr('lyalya_lyalya')['evals']
    assert evals[0].static_xi == 0                        # the peak is rescaled by the sampled (ap, at)
    assert all(d.static_xi == 1 for d in evals[1:])       # smooth and metals: scale_form 0
    vega, pb = builder('configs/full_configs/main.ini', sampled=['alpha_LYA'])
    info = pb.describe_corr('lyalya_lyalya')
    assert info['evals'][0].static_xi == 0 and info['evals'][1].static_xi == 0
    for d, pair in zip(info['evals'][2:], info['item'].metal_correlations):
        assert d.static_xi == int('LYA' not in pair)       # only the LYA x metal terms read alpha_LYA
    vega, pb = builder('configs/full_configs/main.ini', sampled=['drp_QSO'])
    assert all(d.static_xi == 0 for d in pb.describe_corr('lyalya_qso')['evals'])
    assert all(d.static_xi == 1 for d in pb.describe_corr('lyalya_lyalya')['evals'])
    # additive xi terms keep the gather
    vega, pb = builder('configs/picca_bench/main_cross.ini', sampled=['bias_eta_LYA'])
    assert pb.describe_corr('test_7')['evals'][0].static_xi == 0


def test_peak_and_smooth_are_paired_when_their_sampled_factors_agree():
    """vg_eval_desc.pk_twin (vega/model.py:176-186: peak and smooth differ by the linear P(k) and
    the peak's non-linear broadening)."""
    vega, pb = builder('configs/full_configs/main.ini', sampled=['bias_eta_LYA', 'sigma_velo_disp_lorentz_QSO'])
    for name in vega.corr_items:
        evals = pb.describe_corr(name)['evals']
        assert evals[0].pk_twin == -1 and evals[1].pk_twin == 0
        assert all(d.pk_twin == -1 for d in evals[2:])
    # a sampled sigmaNL makes the broadening a per-vector factor of the peak only
    vega, pb = builder('configs/full_configs/main.ini', sampled=['sigmaNL_par'])
    assert pb.describe_corr('lyalya_qso')['evals'][1].pk_twin == -1


def test_sampled_parameter_makes_factor_dynamic():
    vega, pb = builder('configs/full_configs/main.ini',
                       sampled=['sigma_velo_disp_lorentz_QSO', 'L0_hcd', 'par_sigma_smooth'])
    info = pb.describe_corr('lyalya_qso')
    peak = info['evals'][0]
    assert peak.dyn_mask & _lib.F_VELO and peak.dyn_mask & _lib.F_FSGAUSS
    assert peak.hcd_grid == -1                                   # L0 sampled: F evaluated per vector
    assert peak.idx_drp == pb.col('drp_QSO')
    assert peak.radiation == 1


def test_cross_config_options():
    vega, pb = builder('configs/picca_bench/main_cross.ini')
    d7 = pb.describe_corr('test_7')['evals'][0]
    assert d7.rel == 1 and d7.asy == 1 and d7.radiation == 1 and d7.velo_kind == 2
    d = {n: pb.describe_corr(n)['evals'][0] for n in vega.corr_items}
    assert any(x.zevol_kind[0] == 1 for x in d.values())        # a Croom QSO evolution config exists
    assert pb.describe_corr('test_7')['legacy'] is True


def scale_cfg(**kw):
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda o: o
    cfg['cosmo-fit type'] = {k.replace('_', '-') if k != 'cosmo_fit_func' else 'cosmo fit func': str(v)
                             for k, v in kw.items()}
    return cfg['cosmo-fit type']


def test_resolve_scale_rules():
    """vega/scale_parameters.py:38-230"""
    cfg = scale_cfg(cosmo_fit_func='ap_at')
    assert resolve_scale(cfg, {}, True, 'LYAxLYA', False) == (1, 'ap', 'at')
    assert resolve_scale(cfg, {}, False, 'LYAxLYA', False) == (0, None, None)
    assert resolve_scale(cfg, {}, True, 'LYAxSiII', True) == (0, None, None)
    cfg = scale_cfg(cosmo_fit_func='phi_alpha', smooth_scaling=True)
    assert resolve_scale(cfg, {}, True, 'x', False) == (3, 'phi', 'alpha')
    assert resolve_scale(cfg, {}, False, 'x', False) == (3, 'phi_smooth', 'alpha_smooth')
    cfg = scale_cfg(cosmo_fit_func='phi_alpha', full_shape=True)
    assert resolve_scale(cfg, {}, False, 'x', False) == (3, 'phi_full', 'alpha_smooth')
    assert resolve_scale(cfg, {}, True, 'x', False) == (3, 'phi_full', 'alpha')
    cfg = scale_cfg(cosmo_fit_func='phi_alpha', smooth_scaling=True, two_alpha_smooth=True)
    assert resolve_scale(cfg, {}, False, 'LYAxQSO', False) == (3, 'phi_smooth', 'alpha_smooth_LYAxQSO')
    cfg = scale_cfg(cosmo_fit_func='ap_at', full_shape=True)
    with pytest.raises(ValueError):
        resolve_scale(cfg, {}, False, 'x', False)
    cfg = scale_cfg(cosmo_fit_func='aiso_epsilon', full_shape=True, full_shape_alpha=True)
    assert resolve_scale(cfg, {}, False, 'x', False) == (2, 'aiso_full', 'epsilon_full')


def test_fast_metals_descriptors():
    """fast_metals = True (vega/metals.py:144-207): metal x metal terms are frozen, metal x main
    terms with the same cache key (betas + every parameter, not the pair) read the first
    pair's knot tables on the first pair's coordinates."""
    vega, pb = builder('configs/full_configs/main.ini')
    item = vega.corr_items['lyalya_lyalya']
    item.config['model']['fast_metals'] = 'True'
    info = pb.describe_corr('lyalya_lyalya')
    main = [item.tracer1['name'], item.tracer2['name']]
    evals = info['evals'][2:]
    pairs = item.metal_correlations
    with_main = [n1 in main or n2 in main for n1, n2 in pairs]
    assert all(d.frozen == (0 if wm else 1) for d, wm in zip(evals, with_main))
    owners = [d for d, wm in zip(evals, with_main) if wm and d.table_owner < 0]
    aliases = [d for d, wm in zip(evals, with_main) if wm and d.table_owner >= 0]
    # every metal beta of the test configuration is the same fixed value: one owner
    assert len(owners) == 1 and len(aliases) == sum(with_main) - 1
    owner_id = 2 + with_main.index(True)
    assert all(d.table_owner == owner_id and d.coord_set == owners[0].coord_set for d in aliases)
    # the aliases keep their own bias columns (the weight) and metal slot
    assert len({d.out_slot for d in evals}) == len(evals)
    assert len({tuple(d.idx_bias) + tuple(d.idx_bias_eta) for d in aliases + owners}) == len(aliases) + 1
    # growth rate pinned to the fiducial value for the metal evaluations
    assert owners[0].idx_growth_rate == pb.col('growth_rate@fiducial')
    assert info['evals'][0].idx_growth_rate == pb.col('growth_rate')


def test_unsupported_options_fail_loudly():
    vega, pb = builder('configs/full_configs/main.ini')
    model = vega.corr_items['lyalya_lyalya'].config['model']
    model['fht_extrap'] = 'True'
    with pytest.raises(NotImplementedError):
        pb.describe_corr('lyalya_lyalya')
    model['fht_extrap'] = 'False'
    # the reference's metal x metal cache is keyed on the pair alone: with split metals it would
    # return the peak term for the smooth one, so the combination is refused
    model['no-metal-decomp'], model['fast_metals'] = 'False', 'True'
    with pytest.raises(NotImplementedError):
        pb.describe_corr('lyalya_lyalya')


def test_split_metals_emit_a_peak_and_a_smooth_evaluation_per_pair():
    """no-metal-decomp = False (vega/model.py:119-131,178-186)."""
    vega, pb = builder('configs/full_configs/main.ini')
    item = vega.corr_items['lyalya_lyalya']
    item.config['model']['no-metal-decomp'] = 'False'
    info = pb.describe_corr('lyalya_lyalya')
    n_pairs = len(item.metal_correlations)
    metals = [d for d in info['evals'] if d.out_slot >= 2]
    assert len(metals) == 2 * n_pairs == len(info['metal_mats'])
    for c in range(n_pairs):
        peak, smooth = metals[2 * c], metals[2 * c + 1]
        assert (peak.is_peak, peak.pk_lin_sel, peak.idx_weight_extra2) == (1, 0, pb.col('bao_amp'))
        assert (smooth.is_peak, smooth.pk_lin_sel, smooth.idx_weight_extra2) == (0, 1, -1)
        assert peak.coord_set == smooth.coord_set and peak.out_slot + 1 == smooth.out_slot
        assert info['metal_mats'][2 * c] is info['metal_mats'][2 * c + 1]


def test_small_covariances_keep_the_reference_host_inverse():
    """Below DEVICE_LINALG_MIN rows the masked inverse is numpy's LU inverse, bit-identical to
    the reference's (vega/utils.py:271-299), whatever the machine."""
    import numpy as np
    from vega_b200 import utils
    rng = np.random.default_rng(0)
    a = rng.standard_normal((60, 60))
    cov = a @ a.T + 60 * np.eye(60)
    mask = rng.random(60) > 0.3
    masked = cov[:, mask][mask, :]
    assert np.array_equal(utils.compute_masked_invcov(cov, mask), np.linalg.inv(masked))
    assert utils.compute_log_cov_det(cov, mask) == np.linalg.slogdet(masked)[1]
    assert not utils._factor_cache


def test_uv_bias_term_is_static_unless_a_uv_parameter_is_sampled():
    """vg_eval_desc::uv_static: the scale-dependent UV bias (vega/power_spectrum.py:224-261) is a static
    function of k -- and the evaluation eligible for the moment form -- exactly when none of
    bias_prim / lambda_uv / bias_gamma is a batch column."""
    vega, pb = builder('configs/full_configs/main.ini')
    peak = pb.describe_corr('lyalya_lyalya')['evals'][0]
    assert peak.uvb == 1 and peak.uv_on[0] == 1 and peak.uv_static == 1
    cross = pb.describe_corr('lyalya_qso')['evals'][0]
    assert cross.uvb == 1 and list(cross.uv_on) == [1, 0] and cross.uv_static == 1
    vega, pb = builder('configs/full_configs/main.ini', sampled=['bias_eta_LYA', 'beta_LYA', 'bias_gamma'])
    assert pb.describe_corr('lyalya_lyalya')['evals'][0].uv_static == 0
    # metal evaluations are built from the [metals] section: no UV term at all
    metal = pb.describe_corr('lyalya_lyalya')['evals'][2]
    assert metal.uvb == 0 and metal.uv_static == 0


def test_theta_limits_follow_the_sample_section():
    """The box handed to vg_plan_set_theta_limits is the [sample] limits (default_values.txt for
    `name = True` entries): what the velocity-dispersion tables are built over."""
    vega, pb = builder('configs/c2/main.ini')
    limits = vega.sample_params['limits']
    assert limits['sigma_velo_disp_lorentz_QSO'] == (0.0, 15.0)
    assert 'sigma_velo_disp_lorentz_QSO' in pb.sampled
    peak = pb.describe_corr('qsoxlya')['evals'][0]
    assert peak.dyn_mask == _lib.F_VELO and max(peak.idx_velo) == pb.col('sigma_velo_disp_lorentz_QSO')
