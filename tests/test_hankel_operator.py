"""Host-side Hankel + spline operators (vega_b200/hankel.py) against the CPU oracle's
FFT-based transforms and scipy's splines.  No GPU needed: the device applies exactly these
matrices, so this pins the operator half of kernel K2b."""
import numpy as np
import pytest
from scipy.interpolate import interp1d, splev, splrep

from oracle import hankel as oh
from vega_b200 import fitsio, hankel, utils


def template():
    hdul = fitsio.open(utils.find_file('PlanckDR16/PlanckDR16.fits'))
    return np.array(hdul[1].data['K'], float), np.array(hdul[1].data['PK'], float)


def spline_from_operator(xk, y, m, x):
    """Numpy twin of spline_eval() in csrc/vg_xi.cuh."""
    j = np.clip(np.searchsorted(xk, x, side='right') - 1, 0, xk.size - 2)
    h = xk[j + 1] - xk[j]
    t = (x - xk[j]) / h
    a = 1 - t
    return a * y[j] + t * y[j + 1] + ((a**3 - a) * m[j] + (t**3 - t) * m[j + 1]) * h * h / 6


@pytest.mark.parametrize('ell', [0, 2, 4, 6])
def test_mcfit_operator_matches_fft(ell):
    k, pk = template()
    pk_ell = pk * np.exp(-(k * 3.)**2) * (1 + 0.3 * ell)
    r_ref, xi_ref = oh.P2xi(k, l=ell, lowring=True)(pk_ell)
    r, H = hankel.fftlog_mcfit_operator(k, ell)
    assert np.allclose(r, r_ref, rtol=1e-14, atol=0)
    xi = H @ pk_ell
    scale = np.max(np.abs(xi_ref[(r > 1) & (r < 400)]))
    sel = (r > 1) & (r < 400)
    assert np.max(np.abs(xi - xi_ref)[sel]) < 1e-12 * scale


@pytest.mark.parametrize('ell', [0, 2, 4, 6])
def test_operator_spline_matches_interp1d(ell):
    k, pk = template()
    pk_ell = pk * np.exp(-(k * 3.)**2)
    xk, op = hankel.build_operator_set(k, (0, 2, 4, 6), legacy=False)
    i = ell // 2
    y, m = op[i, 0] @ pk_ell, op[i, 1] @ pk_ell
    r = np.exp(np.linspace(np.log(2.), np.log(380.), 4000))
    r_ref, xi_ref = oh.P2xi(k, l=ell, lowring=True)(pk_ell)
    ref = interp1d(np.log(r_ref), xi_ref, kind='cubic')(np.log(r))
    got = spline_from_operator(xk[i], y, m, np.log(r))
    assert np.max(np.abs(got - ref)) < 1e-11 * np.max(np.abs(ref))


@pytest.mark.parametrize('ell', [0, 2, 4, 6])
def test_legacy_operator_matches_fft_and_splrep(ell):
    k, pk = template()
    mu = ((np.arange(1000) + 0.5) / 1000)[:, None]
    pk2d = pk * np.exp(-(k * 3.)**2) * (1 + 1.5 * mu**2)**2
    r = np.exp(np.linspace(np.log(2.), np.log(380.), 3000))
    ref = oh.legacy_multipoles(r, k, pk2d, [0, 2, 4, 6], mu, 1e-3)[ell // 2]
    from scipy.special import legendre
    pk_ell = np.sum(1e-3 * legendre(ell)(mu) * pk2d, axis=0) * (2 * ell + 1)
    xk, op = hankel.build_operator_set(k, (0, 2, 4, 6), legacy=True)
    i = ell // 2
    got = spline_from_operator(xk[i], op[i, 0] @ pk_ell, op[i, 1] @ pk_ell, np.log(r))
    assert np.max(np.abs(got - ref)) < 1e-11 * np.max(np.abs(ref))


def test_static_tables_rel_asy():
    k, pk = template()
    r = np.exp(np.linspace(np.log(2.), np.log(380.), 500))
    mu = ((np.arange(1000) + 0.5) / 1000)[:, None]
    for ells, tform, n_pow in (([1, 3], 'rel', 1.), ([0, 2], 'asy', 2.)):
        ref = oh.legacy_multipoles(r, k, pk, ells, mu, 1e-3, tform=tform)
        for ell in ells:
            tab = hankel.build_static_table(k, pk, ell, n_pow)
            got = spline_from_operator(tab[0], tab[1], tab[2], np.log(r))
            assert np.max(np.abs(got - ref[ell // 2])) < 1e-11 * np.max(np.abs(ref[ell // 2]))


@pytest.mark.parametrize('lowring', [True, False])
def test_p2xi_analytic_gaussian_pair(lowring):
    """P(k) = exp(-k^2 a^2 / 2)  <->  xi_0(r) = (2 pi)^-3/2 a^-3 exp(-r^2 / 2 a^2), with and
    without the low-ringing shift.  With lowring off the output grid is y = 1 / (x e^Delta)
    reversed, i.e. ln(xy) = 0 as in mcfit (xy = 1), not Delta."""
    k, _ = template()
    a = 30.
    pk = np.exp(-0.5 * (k * a)**2)
    for make in (lambda: oh.P2xi(k, l=0, lowring=lowring)(pk),
                 lambda: (lambda r, H: (r, H @ pk))(*hankel.fftlog_mcfit_operator(k, 0, lowring=lowring))):
        r, xi = make()
        exact = (2 * np.pi)**-1.5 / a**3 * np.exp(-0.5 * (r / a)**2)
        sel = (r > 1.) & (r < 150.)
        assert np.max(np.abs(xi - exact)[sel]) < 2e-6 * exact.max()
    if not lowring:
        delta = np.log(k[-1] / k[0]) / (k.size - 1)
        assert np.allclose(r, np.exp(-delta) / k[::-1], rtol=1e-14, atol=0)


@pytest.mark.parametrize('ell', [2, 4])
def test_p2xi_lowring_off_agrees_with_lowring_on(ell):
    """Both grids sample the same continuous transform (interpolated onto common radii)."""
    k, pk = template()
    pk_ell = pk * np.exp(-(k * 3.)**2)
    r_on, xi_on = oh.P2xi(k, l=ell, lowring=True)(pk_ell)
    r_off, xi_off = oh.P2xi(k, l=ell, lowring=False)(pk_ell)
    r = np.exp(np.linspace(np.log(5.), np.log(300.), 400))
    on = interp1d(np.log(r_on), xi_on, kind='cubic')(np.log(r))
    off = interp1d(np.log(r_off), xi_off, kind='cubic')(np.log(r))
    assert np.max(np.abs(on - off)) < 2e-5 * np.max(np.abs(on))
