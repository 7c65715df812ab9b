"""Known answers of the reference's tests/test_pk.py (means of the full P(k,mu) model,
Kaiser sum, fast-metals mean) on the CPU oracle and, through the stage-0 tap of the C ABI
(materialising kernel K1), on the CUDA path."""
import configparser

import numpy as np
import pytest

from oracle import vega_oracle as vo
from vega_b200 import VegaInterface, fitsio, utils

AUTO_PARAMS = {'bias_LYA': -0.12, 'beta_LYA': 1.6, 'bias_gamma': 0.1125, 'bias_prim': -0.66,
               'lambda_uv': 300, 'bias_hcd': -0.05, 'beta_hcd': 0.5, 'L0_hcd': 10,
               'sigmaNL_par': 6.36984, 'sigmaNL_per': 3.24, 'par_sigma_smooth': 2,
               'per_sigma_smooth': 2.5, 'dnl_arinyo_q1': 0.8558, 'dnl_arinyo_kv': 1.11454,
               'dnl_arinyo_av': 0.5378, 'dnl_arinyo_bv': 1.607, 'dnl_arinyo_kp': 19.47}
CROSS_PARAMS = {'bias_LYA': -0.12, 'beta_LYA': 1.6, 'bias_QSO': 3.7, 'beta_QSO': 0.26,
                'bias_gamma': 0.1125, 'bias_prim': -0.66, 'lambda_uv': 300, 'bias_hcd': -0.05,
                'beta_hcd': 0.5, 'L0_hcd': 10, 'sigmaNL_par': 6.36984, 'sigmaNL_per': 3.24,
                'par_sigma_smooth': 2, 'per_sigma_smooth': 2.5, 'sigma_velo_disp_lorentz_QSO': 7.2}
AUTO_MODEL = {'model-hcd': 'Rogers', 'UVB-fluctuations': 'True', 'fvoigt_model': 'exp',
              'small scale nl': 'arinyo', 'fullshape smoothing': 'gauss'}
CROSS_MODEL = {'model-hcd': 'Rogers', 'UVB-fluctuations': 'True', 'fvoigt_model': 'exp',
               'fullshape smoothing': 'gauss', 'velocity dispersion': 'lorentz'}
LYA = {'name': 'LYA', 'type': 'continuous'}
QSO = {'name': 'QSO', 'type': 'discrete'}


def fiducial():
    hdul = fitsio.open(utils.find_file('PlanckDR16/PlanckDR16.fits'))
    return {'z_eff': 2.25, 'k': np.array(hdul[1].data['K'], float),
            'pk_full': np.array(hdul[1].data['PK'], float),
            'pk_smooth': np.array(hdul[1].data['PKSB'], float),
            'z_fiducial': hdul[1].header['ZREF']}


def section(options):
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda o: o
    cfg['model'] = {'bin_size_rp': '4', 'bin_size_rt': '4', **options}
    return cfg['model']


def test_oracle_pk_known_answers():
    """reference tests/test_pk.py:172-231,262-266"""
    fid = fiducial()
    g = vo.PkGrid(section({'model binning': 'False'}), fid, LYA, LYA, 'lyaxlya')
    pars = {'bias_LYA': -0.12, 'beta_LYA': 1.6, 'peak': False}
    pk = vo.pk_model(g, fid['pk_smooth'], pars)
    assert pk.shape == (1000, 814)
    assert np.sum(pk[:, :1] / fid['pk_smooth'][0]) == pytest.approx(37.13279)
    g = vo.PkGrid(section({}), fid, LYA, LYA, 'lyaxlya')
    assert np.sum(vo.binning_kernel(g, pars)) == pytest.approx(450783.949889)
    assert np.mean(vo.pk_model(g, fid['pk_smooth'], pars, fast_metals=True)) == pytest.approx(1228.9847366)

    g = vo.PkGrid(section(AUTO_MODEL), fid, LYA, LYA, 'lyaxlya')
    assert np.mean(vo.pk_model(g, fid['pk_full'] - fid['pk_smooth'], {**AUTO_PARAMS, 'peak': True})) \
        == pytest.approx(2.8794436016)
    assert np.mean(vo.pk_model(g, fid['pk_smooth'], {**AUTO_PARAMS, 'peak': False})) \
        == pytest.approx(19.67878957)
    g = vo.PkGrid(section(CROSS_MODEL), fid, LYA, QSO, 'lyaxqso')
    assert np.mean(vo.pk_model(g, fid['pk_full'] - fid['pk_smooth'], {**CROSS_PARAMS, 'peak': True})) \
        == pytest.approx(-2.9406788865)
    assert np.mean(vo.pk_model(g, fid['pk_smooth'], {**CROSS_PARAMS, 'peak': False})) \
        == pytest.approx(-401.0937936)


def write_config(tmp_path, name, tracers, model, params, data_file):
    corr = configparser.ConfigParser()
    corr.optionxform = lambda o: o
    corr['data'] = {'name': name, 'tracer1': tracers[0]['name'], 'tracer2': tracers[1]['name'],
                    'tracer1-type': tracers[0]['type'], 'tracer2-type': tracers[1]['type'],
                    'filename': data_file}
    corr['cuts'] = {'r-min': '10', 'r-max': '180'}
    corr['model'] = {'z evol LYA': 'bias_vs_z_std', 'z evol QSO': 'bias_vs_z_std', **model}
    corr['parameters'] = {f'par binsize {name}': '4', f'per binsize {name}': '4'}
    with open(tmp_path / 'corr.ini', 'w') as f:
        corr.write(f)
    main = configparser.ConfigParser()
    main.optionxform = lambda o: o
    main['data sets'] = {'zeff': '2.25', 'ini files': str(tmp_path / 'corr.ini')}
    main['cosmo-fit type'] = {'cosmo fit func': 'ap_at'}
    main['fiducial'] = {'filename': 'PlanckDR16/PlanckDR16.fits'}
    main['sample'] = {}
    main['parameters'] = {'ap': '1', 'at': '1', 'bao_amp': '1', 'growth_rate': '0.970386',
                          'alpha_LYA': '2.9', 'alpha_QSO': '1.44', 'drp_QSO': '0.',
                          **{k: repr(float(v)) for k, v in params.items()}}
    with open(tmp_path / 'main.ini', 'w') as f:
        main.write(f)
    return tmp_path / 'main.ini'


@pytest.mark.gpu
@pytest.mark.parametrize('kind', ['auto', 'cross'])
@pytest.mark.parametrize('sampled', [[], ['sigmaNL_par', 'par_sigma_smooth', 'dnl_arinyo_q1', 'L0_hcd',
                                          'sigma_velo_disp_lorentz_QSO']])
def test_cuda_pk_known_answers(tmp_path, kind, sampled):
    """Same numbers from the materialising kernel, with the factors folded into the static
    grid (nothing sampled) and evaluated per vector (their parameters sampled)."""
    from vega_b200.plan import PlanBuilder
    if kind == 'auto':
        main = write_config(tmp_path, 'lyaxlya', (LYA, LYA), AUTO_MODEL, AUTO_PARAMS, 'data/dr16_simple_auto.fits')
        want = (2.8794436016, 19.67878957)
    else:
        main = write_config(tmp_path, 'lyaxqso', (LYA, QSO), CROSS_MODEL, CROSS_PARAMS, 'data/dr16_simple_cross.fits')
        want = (-2.9406788865, -401.0937936)
    vega = VegaInterface(str(main))
    sampled = [s for s in sampled if s in vega.params]
    engine = PlanBuilder(vega, sampled=sampled).build()
    theta = np.array([[vega.params[s] for s in sampled]]) if sampled else np.zeros((1, 1))
    for eval_index, value in enumerate(want):
        grid, status = engine.stage(0, 0, theta, eval_index=eval_index)
        assert status[0] == 0
        assert np.mean(grid) == pytest.approx(value)
