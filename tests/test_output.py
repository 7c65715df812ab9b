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
    assert np.array_equal(table.data['ap'], [0.98, 1.0, 1.02]) and np.array_equal(table.data['fval'], [10., 11., 12.])
    assert table.header['ap_min'] == 0.98 and table.header['ap_num_bins'] == 3
