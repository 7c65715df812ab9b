"""Velocity-dispersion tables (csrc/vg_moments.cuh, VG_OPT_VELO_TABLES): evaluations whose only sampled
non-polynomial P(k,mu) factor is one velocity dispersion (vega/power_spectrum.py:588-636) read their
knot tables from a Chebyshev series in sigma_v built at finalize.  The per-point pass is the parity
reference (option off), the oracle the arbiter."""
import numpy as np
import pytest

from helpers import make_c2

pytestmark = pytest.mark.gpu
SIGMA = 'sigma_velo_disp_lorentz_QSO'


def _rel(a, b):
    return np.max(np.abs(a - b)) / np.max(np.abs(b))


@pytest.fixture(scope='module')
def c2():
    vega, oracle, theta = make_c2(160, width=3.0)
    names = list(vega.sample_params['limits'])
    lo, hi = vega.sample_params['limits'][SIGMA]
    # the dispersion over its WHOLE [sample] range, both ends included
    theta[:, names.index(SIGMA)] = np.linspace(lo, hi, theta.shape[0])
    return vega, oracle, theta, names


def test_tables_are_built_and_checked_at_finalize(c2):
    vega, _, _, _ = c2
    engine = vega._engine_for()
    served = [engine.velo_tables(ci) for ci in range(len(engine.infos))]
    # the cross-correlation's peak, smooth and shared QSO x metal table read the sampled Lorentzian
    assert sum(n for n, _, _ in served) == 3
    for n, orders, err in served:
        if n:
            assert 8 <= orders <= 64 and err <= 2e-12


def test_tables_match_the_per_point_pass_and_the_oracle(c2):
    from vega_b200 import _lib
    vega, oracle, theta, names = c2
    engine = vega._engine_for()
    chi2_tab = vega.chi2_batch(theta)
    assert engine.velo_tables_active()
    models_tab, _ = vega.compute_model_batch(theta)
    assert engine.velo_tables_active()
    engine.set_option(_lib.OPT_VELO_TABLES, 0)
    chi2_pp = vega.chi2_batch(theta)
    assert not engine.velo_tables_active()
    models_pp, _ = vega.compute_model_batch(theta)
    engine.set_option(_lib.OPT_VELO_TABLES, 1)
    for name in models_pp:
        assert _rel(models_tab[name], models_pp[name]) <= 1e-11, name
    assert np.max(np.abs(chi2_tab - chi2_pp) / np.abs(chi2_pp)) <= 1e-10
    for b in (0, 1, 77, theta.shape[0] - 1):           # sigma = lo, just above, mid-range, hi
        ref = oracle.compute_model(dict(zip(names, theta[b])))
        for name in ref:
            assert _rel(models_tab[name][b], ref[name]) <= 1e-10, (b, name)
        c_ref = oracle.chi2(dict(zip(names, theta[b])))
        assert abs(chi2_tab[b] - c_ref) <= 1e-10 * abs(c_ref), (b, chi2_tab[b], c_ref)


def test_vectors_outside_the_range_keep_the_per_point_pass_and_small_batches_use_the_tables(c2):
    from vega_b200 import _lib
    vega, _, theta, names = c2
    engine = vega._engine_for()
    out = theta.copy()
    out[5, names.index(SIGMA)] = vega.sample_params['limits'][SIGMA][1] + 0.5
    chi2 = vega.chi2_batch(out)
    assert not engine.velo_tables_active()
    # batches of a few vectors (a gradient stencil, the scalar API; CUDA-graph replay from the third call)
    small = [vega.chi2_batch(theta[60:89]) for _ in range(4)]
    assert engine.velo_tables_active()
    one = vega.chi2(dict(zip(names, theta[70])))
    engine.set_option(_lib.OPT_VELO_TABLES, 0)
    np.testing.assert_array_equal(chi2, vega.chi2_batch(out))
    ref = vega.chi2_batch(theta[60:89])
    for got in small:
        np.testing.assert_allclose(got, ref, rtol=1e-11)
    np.testing.assert_allclose(one, ref[10], rtol=1e-11)
    engine.set_option(_lib.OPT_VELO_TABLES, 1)
    # device-resident theta takes the same decision (one range kernel per call)
    import torch
    th = torch.as_tensor(theta, device='cuda')
    chi2_dev, _ = engine.chi2_device(th)
    assert engine.velo_tables_active()
    np.testing.assert_allclose(chi2_dev.cpu().numpy(), vega.chi2_batch(theta), rtol=1e-13)
    chi2_dev, _ = engine.chi2_device(torch.as_tensor(out, device='cuda'))
    assert not engine.velo_tables_active()
