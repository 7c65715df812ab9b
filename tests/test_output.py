"""Result files in the reference's FITS layout (vega/output.py:13-350), written with the package's
own FITS writer and read back."""
import types

import numpy as np
import pytest

from helpers import make_oracle
from vega_b200 import fitsio


def test_write_results_round_trip(tmp_path):
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    vega.main_config['output']['filename'] = str(tmp_path / 'result')
    model = oracle.compute_model()
    names = ['bias_eta_LYA', 'beta_LYA', 'BB-lyalya_lyalya-0 add post r,mu (1,2)']
    cov = np.array([[4e-4, 1e-5, 0.], [1e-5, 9e-2, 0.], [0., 0., 1e-6]])
    minimizer = types.SimpleNamespace(
        values={n: vega.params[n] for n in names}, errors=dict(zip(names, np.sqrt(np.diag(cov)))),
        covariance=cov, fmin=types.SimpleNamespace(fval=9541.25, is_valid=True))
    stats = {name: {'masked_size': int(vega.data[name].data_mask.sum()), 'chisq': 1.5, 'reduced_chisq': 1.0,
                    'p_value': 0.5, 'bestfit_marg_coeff': None} for name in vega.corr_items}
    scan = [{'ap': ap, 'bias_eta_LYA': -0.2 + 0.01 * i, 'fval': 10. + i} for i, ap in enumerate((0.98, 1.0, 1.02))]
    vega._analysis = types.SimpleNamespace(grids={'ap': np.linspace(0.98, 1.02, 3)})
    path = vega.output.write_results(model, vega.params, minimizer, stats, scan, vega.models)
    assert path.suffix == '.fits'
    with pytest.raises(OSError):                       # overwrite = False by default
        vega.output.write_results(model, vega.params, minimizer, stats, scan, vega.models)

    hdul = fitsio.open(path)
    by_name = {h.header.get('EXTNAME'): h for h in hdul[1:]}
    assert list(by_name) == ['MODEL_' + n for n in vega.corr_items] + ['BESTFIT', 'SCAN']
    for name in vega.corr_items:
        hdu = by_name['MODEL_' + name]
        assert np.array_equal(hdu.data[name + '_MODEL'], model[name])
        assert np.array_equal(hdu.data[name + '_MASK'], vega.data[name].data_mask)
        assert np.array_equal(hdu.data[name + '_MODEL_MASK'], vega.data[name].model_mask)
        assert np.array_equal(hdu.data[name + '_DATA'], vega.data[name].data_vec, equal_nan=True)
        assert np.array_equal(hdu.data[name + '_RP'], vega.corr_items[name].dist_model_coordinates.rp_grid)
        assert hdu.header['bias_eta_LYA'] == vega.params['bias_eta_LYA']
        assert hdu.header['BB-lyalya_lyalya-0 add post r,mu (1,2)'] == \
            vega.params['BB-lyalya_lyalya-0 add post r,mu (1,2)']
        assert hdu.header['chisq'] == 1.5 and hdu.header['masked_size'] == stats[name]['masked_size']
    best = by_name['BESTFIT']
    assert list(best.data['names']) == names
    assert np.array_equal(best.data['covariance'], cov)
    assert np.allclose(best.data['errors'], np.sqrt(np.diag(cov)))
    assert best.header['FVAL'] == 9541.25 and best.header['VALID'] is True
    table = by_name['SCAN']
    assert np.array_equal(table.data['ap'][:3], [0.98, 1.0, 1.02]) and np.array_equal(table.data['fval'][:3], [10., 11., 12.])
    assert [str(n) for n in table.data['names'][:3]] == ['ap', 'bias_eta_LYA', 'fval']
    assert table.header['ap_min'] == 0.98 and table.header['ap_num_bins'] == 3


def test_scan_table_with_more_names_than_grid_points(tmp_path):
    """astropy pads the shorter columns of BinTableHDU.from_columns (vega/output.py:326): every
    parameter name is kept when the scan has fewer grid points than sampled parameters."""
    import configparser
    from vega_b200.output import Output
    cfg = configparser.ConfigParser()
    cfg['output'] = {'filename': str(tmp_path / 'scan'), 'overwrite': 'True'}
    out = Output(cfg['output'], data={}, corr_items={})
    names = ['ap', 'at', 'bias_eta_LYA', 'beta_LYA', 'fval']
    scan = [{n: float(i + j) for j, n in enumerate(names)} for i in range(2)]
    extname, header, columns = out._scan_table(scan)
    assert [str(n) for n in columns['names']] == names
    assert np.array_equal(columns['ap'], [0., 1., 0., 0., 0.]) and np.array_equal(columns['fval'][:2], [4., 5.])


def test_write_monte_carlo_round_trip(tmp_path):
    """Output.write_monte_carlo (vega/output.py:442-520): Bestfit / FitInfo / Mocks tables, the
    per-rank file name of bin/run_vega_mc_fits_mpi.py:155-159."""
    import configparser
    import types
    from vega_b200 import fitsio
    from vega_b200.output import Output
    cfg = configparser.ConfigParser()
    cfg['output'] = {'filename': str(tmp_path / 'run' / 'out'), 'overwrite': 'False'}
    rng = np.random.default_rng(0)
    n_mocks, names = 5, ['ap', 'at', 'bias_eta_LYA']
    analysis = types.SimpleNamespace(
        has_monte_carlo=True,
        mc_bestfits={n: rng.standard_normal((n_mocks, 2)) for n in names},
        mc_covariances=[rng.standard_normal((3, 3)) for _ in range(n_mocks)],
        mc_chisq=list(rng.uniform(9000, 10000, n_mocks)), mc_valid_minima=[True, True, False, True, True],
        mc_valid_hesse=[True] * n_mocks, mc_failed_mask=[False] * n_mocks,
        mc_mocks={'lyaxlya': [rng.standard_normal(7) for _ in range(n_mocks)],
                  'qsoxlya': [rng.standard_normal(11) for _ in range(n_mocks)]})
    out = Output(cfg['output'], data={}, corr_items={}, analysis=analysis)
    path = out.write_monte_carlo(cpu_id=3)
    assert path == tmp_path / 'run' / 'monte_carlo' / 'monte_carlo_3.fits'
    with pytest.raises(OSError):
        out.write_monte_carlo(cpu_id=3)
    hdul = fitsio.open(path)
    by_name = {h.header.get('EXTNAME'): h for h in hdul[1:]}
    assert list(by_name) == ['Bestfit', 'FitInfo', 'Mocks']
    best = by_name['Bestfit'].data
    assert [str(n) for n in best['names']] == names
    for i, n in enumerate(names):
        assert np.array_equal(best['values'][i], analysis.mc_bestfits[n][:, 0])
        assert np.array_equal(best['errors'][i], analysis.mc_bestfits[n][:, 1])
    cov = np.array(analysis.mc_covariances).reshape(n_mocks * 3, 3).T
    assert np.array_equal(np.asarray(best['covariance']).reshape(3, -1), cov)
    info = by_name['FitInfo'].data
    assert np.array_equal(info['chisq'], analysis.mc_chisq)
    assert list(info['valid_minima']) == analysis.mc_valid_minima
    mocks = by_name['Mocks'].data
    assert np.array_equal(mocks['qsoxlya'], np.array(analysis.mc_mocks['qsoxlya']))
    analysis.has_monte_carlo = False
    with pytest.raises(AssertionError):
        out.write_monte_carlo()
