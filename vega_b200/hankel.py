"""Init-time construction of the Hankel + spline operators the device applies as one
FP64 tensor-core GEMM over the batch (kernel K2b).

The reference turns each multipole P_l(k) into xi_l on a log-r grid with an FFTLog and
then reads it off with a global cubic spline in log r:
  * default path: ``mcfit.P2xi(k, l=ell, lowring=True)`` (vega/pktoxi.py:53,141) followed
    by ``interp1d(log r, xi, kind='cubic')`` (vega/pktoxi.py:144-152);
  * ``old_fftlog = True``: the in-tree ``PktoXi.Pk2Mp`` (vega/pktoxi.py:230-279) followed
    by ``splrep(log r - dr/2, xi, k=3, s=0)`` / ``splev`` (vega/pktoxi.py:274-277).
Both steps are LINEAR in P_l(k) for a fixed k grid, so here they are folded, once per
plan, into dense operators: ``xi_knots = H P_l`` (FFTLog as a matrix) and
``m_knots = D H P_l`` (second derivatives of the not-a-knot interpolating cubic spline,
which is the spline both interp1d(kind='cubic') and splrep(k=3, s=0) construct).
The device then evaluates the spline piecewise from (xi_knots, m_knots).

mcfit is an unpinned third-party dependency whose source is not part of the reference
tree; ``fftlog_mcfit_operator`` follows its published algorithm.
"""
import numpy as np
from scipy.special import loggamma


def _mellin_sph_bessel(ell, z):
    """Mellin transform of sqrt(2/pi) * j_ell at complex z."""
    return np.exp(np.log(2.) * (z - 1.5) + loggamma(0.5 * (ell + z)) - loggamma(0.5 * (3. + ell - z)))


def fftlog_mcfit_operator(k, ell, q=1.5, lowring=True):
    """Matrix form of ``r, xi = mcfit.P2xi(k, l=ell, lowring=lowring)(pk_ell, extrap=False)``.

    Returns (r[n], H[n, n]) with xi = H @ pk_ell.
    """
    k = np.asarray(k, dtype=float)
    nin = k.size
    delta = np.log(k[-1] / k[0]) / (nin - 1)
    n_fft = int(2 ** np.ceil(np.log2(2 * nin)))
    if lowring:
        lnxy = delta / np.pi * np.angle(_mellin_sph_bessel(ell, q + 1j * np.pi / delta))
    else:
        lnxy = 0.0          # mcfit: lnxy = log(xy), xy = 1, when the low-ringing shift is off
    r = np.exp(lnxy - delta) / k[::-1]
    m = np.arange(n_fft // 2 + 1)
    u = _mellin_sph_bessel(ell, q + 2j * np.pi * m / (n_fft * delta))
    u = u * np.exp(-2j * np.pi * lnxy * m / (n_fft * delta))

    pad = n_fft - nin
    pad_left, pad_right = pad // 2, pad - pad // 2
    prefac = k**3 / (2 * np.pi)**1.5 * k**(-q)
    postfac = (-1.)**(ell // 2) * r**(-q)

    f = np.zeros((nin, n_fft))
    f[np.arange(nin), pad_left + np.arange(nin)] = prefac
    g = np.fft.hfft(np.fft.rfft(f, axis=1) * u, n=n_fft, axis=1) / n_fft
    g = g[:, pad_right:pad_right + nin] * postfac
    return r, np.ascontiguousarray(g.T)


def fftlog_legacy_operator(k, ell, n_pow=2., multipole_prefactor=True):
    """Matrix form of the in-tree Hamilton FFTLog of vega/pktoxi.py:230-279.

    Returns (knots[n] = log r - dr/2, H[n, n]) with xi = H @ pk_ell; the last sample is
    forced to zero as in the reference (:273).  ``multipole_prefactor`` applies the
    (-1)^(ell//2) / (2 pi^2) of the standard transform (:259-261); the relativistic /
    asymmetry transforms (tform 'rel' n_pow=1, 'asy' n_pow=2) take the linear P(k) as is.
    """
    k = np.asarray(k, dtype=float)
    n = k.size
    k0 = k[0]
    span = np.log(k.max() / k0)
    emm = n * np.fft.fftfreq(n)
    r = np.exp(-emm * span / n)
    dr = abs(np.log(r[1] / r[0]))
    order = np.argsort(r)
    r = r[order]

    mu = ell + 0.5
    qq = 2 - n_pow - 0.5
    x = qq + 2 * np.pi * 1j * emm / span
    um = (k0 * 1.)**(-2 * np.pi * 1j * emm / span) * 2**x * np.exp(
        loggamma((mu + 1 + x) / 2) - loggamma((mu + 1 - x) / 2))
    um[0] = np.real(um[0])

    pre = k**n_pow * np.sqrt(np.pi / 2)
    if multipole_prefactor:
        pre = pre * (-1)**(ell // 2) / 2 / np.pi**2
    a = np.fft.fft(np.diag(pre), axis=1) * um
    xi = np.real(np.fft.ifft(a, axis=1)[:, order]) / r**(3 - n_pow)
    xi[:, -1] = 0.
    return np.log(r) - dr / 2, np.ascontiguousarray(xi.T)


def notaknot_second_derivative_operator(x):
    """D[n, n] with m = D @ y the knot second derivatives of the interpolating cubic spline
    with not-a-knot end conditions through (x_i, y_i)."""
    x = np.asarray(x, dtype=float)
    n = x.size
    h = np.diff(x)
    T = np.zeros((n, n))
    R = np.zeros((n, n))
    for i in range(1, n - 1):
        T[i, i - 1], T[i, i], T[i, i + 1] = h[i - 1], 2 * (h[i - 1] + h[i]), h[i]
        R[i, i - 1], R[i, i], R[i, i + 1] = 6 / h[i - 1], -6 / h[i - 1] - 6 / h[i], 6 / h[i]
    T[0, 0], T[0, 1], T[0, 2] = h[1], -(h[0] + h[1]), h[0]
    T[-1, -3], T[-1, -2], T[-1, -1] = h[-1], -(h[-2] + h[-1]), h[-2]
    return np.linalg.solve(T, R)


_CACHE = {}


def build_operator_set(k, ells, legacy, lowring=True):
    """(knots[n_ell, n], op[n_ell, 2, n, nk]) for vg_plan_add_hankel."""
    key = (k.tobytes(), tuple(ells), bool(legacy), bool(lowring))
    if key not in _CACHE:
        knots, ops = [], []
        for ell in ells:
            if legacy:
                xk, H = fftlog_legacy_operator(k, ell)
            else:
                r, H = fftlog_mcfit_operator(k, ell, lowring=lowring)
                xk = np.log(r)
            D = notaknot_second_derivative_operator(xk)
            knots.append(xk)
            ops.append(np.stack([H, D @ H]))
        _CACHE[key] = (np.ascontiguousarray(np.stack(knots)), np.ascontiguousarray(np.stack(ops)))
    return _CACHE[key]


def build_static_table(k, pk, ell, n_pow):
    """[3, n] spline table (knots, values, second derivatives) of the relativistic /
    asymmetry multipole of the linear P(k) (vega/pktoxi.py:321-382)."""
    xk, H = fftlog_legacy_operator(k, ell, n_pow=n_pow, multipole_prefactor=False)
    y = H @ pk
    m = notaknot_second_derivative_operator(xk) @ y
    return np.ascontiguousarray(np.stack([xk, y, m]))
