"""Pins the CPU oracle against the reference's own golden vectors / known answers
(reference tests/test_vega.py:10-44)."""
import numpy as np
import pytest

from vega_b200 import fitsio, utils
from helpers import make_oracle


@pytest.mark.parametrize('kind,main', [('auto', 'configs/picca_bench/main.ini'),
                                       ('cross', 'configs/picca_bench/main_cross.ini')])
def test_picca_golden_vectors(kind, main):
    vega, oracle = make_oracle(main, omega_de_none=True)
    xi = oracle.compute_model()
    hdul = fitsio.open(utils.find_file('data/picca_bench_data.fits'))
    hdu = hdul[1] if kind == 'auto' else hdul[2]
    for i in (0, 1, 2, 4, 5, 6, 7):
        gold = np.array(hdu.data[f'{kind}_test_{i}'])
        assert np.allclose(xi[f'test_{i}'], gold), (kind, i)


def test_full_config_log_lik():
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    # reference tests/test_vega.py:13-14 pins this with math.isclose (rel 1e-9)
    assert np.isclose(oracle.log_lik(), -8766.997108462287, rtol=1e-9, atol=0)
