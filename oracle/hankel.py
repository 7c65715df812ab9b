"""ORACLE (test infrastructure, never shipped or timed as the product).

CPU restatement of the two Hankel transforms on Vega's hot path.

1. ``P2xi``: the third-party ``mcfit.P2xi`` FFTLog that the reference
   constructs at vega/pktoxi.py:53 and calls at vega/pktoxi.py:141,188.  mcfit
   is an *unpinned* PyPI dependency (pyproject.toml:36, ``"mcfit"``), its source
   is not under /root/reference and it is not installed in this image, so this
   follows the published algorithm (Hamilton 2000 FFTLog with a power-law bias
   q=1.5, low-ringing shift, zero padding to N = 2^ceil(log2(2 Nin))) as laid
   out in SURVEY.md section 8(c).  PARITY UNPINNED at this boundary: the
   reference holds no golden xi vector for the mcfit path (only the weak
   log_lik known answer of tests/test_vega.py:14); the restatement is validated
   against analytic transform pairs instead (tests/test_oracle_hankel.py).

2. ``legacy_multipoles``: the in-tree FFTLog ``PktoXi.Pk2Mp``
   (vega/pktoxi.py:230-279), the path the picca golden vectors pin
   (tests/test_vega.py:21-44).
"""
import numpy as np
from scipy.special import loggamma
from scipy.interpolate import splrep, splev


def mellin_sph_bessel(ell):
    """Mellin transform of sqrt(2/pi) j_ell, as a function of complex z."""
    def kernel(z):
        return np.exp(np.log(2.) * (z - 1.5) + loggamma(0.5 * (ell + z))
                      - loggamma(0.5 * (3. + ell - z)))
    return kernel


class P2xi:
    """xi_ell(r) = i^ell int k^2 dk / (2 pi^2) P(k) j_ell(kr) on a log grid.

    Same constructor / call shape as the object built at vega/pktoxi.py:53:
    ``P2xi(k, l=ell, lowring=True)`` then ``r, xi = obj(pk_ell, extrap=False)``.
    """

    def __init__(self, k, l=0, q=1.5, N=2j, lowring=True):  # noqa: E741
        self.x = np.asarray(k, dtype=float)
        self.ell = l
        self.q = q
        nin = self.x.size
        self.nin = nin
        self.delta = np.log(self.x[-1] / self.x[0]) / (nin - 1)
        if isinstance(N, complex):
            self.N = int(2 ** np.ceil(np.log2(nin * N.imag)))
        else:
            self.N = int(N)
        assert self.N >= nin

        kernel = mellin_sph_bessel(l)
        if lowring:
            lnxy = self.delta / np.pi * np.angle(kernel(q + 1j * np.pi / self.delta))
        else:
            lnxy = 0.0      # mcfit: lnxy = log(xy) with the default xy = 1 when lowring is off
        self.lnxy = lnxy
        self.y = np.exp(lnxy - self.delta) / self.x[::-1]

        m = np.arange(self.N // 2 + 1)
        u = kernel(q + 2j * np.pi * m / (self.N * self.delta))
        u = u * np.exp(-2j * np.pi * lnxy * m / (self.N * self.delta))
        self.u = u

        npad = self.N - nin
        self.pad_left = npad // 2
        self.pad_right = npad - self.pad_left
        self.prefac = self.x**3 / (2 * np.pi)**1.5 * self.x**(-q)
        # i^ell for even ell is real
        assert l % 2 == 0
        self.postfac = (-1.)**(l // 2) * self.y**(-q)

    def __call__(self, F, extrap=False):
        assert not extrap, 'only zero padding is restated (fht_extrap=False default)'
        f = np.zeros(self.N)
        f[self.pad_left:self.pad_left + self.nin] = self.prefac * F
        fk = np.fft.rfft(f)
        g = np.fft.hfft(fk * self.u, n=self.N) / self.N
        # output padding is mirrored with respect to the input padding
        g = g[self.pad_right:self.pad_right + self.nin]
        return self.y, self.postfac * g


def legacy_grid(k):
    """Output r grid of the legacy transform (vega/pktoxi.py:238-247)."""
    k0 = k[0]
    span = np.log(k.max() / k0)
    n = len(k)
    emm = n * np.fft.fftfreq(n)
    r = np.exp(-emm * span / n)
    dr = abs(np.log(r[1] / r[0]))
    order = np.argsort(r)
    return k0, span, emm, r[order], dr, order


def legacy_xi_samples(k, pk_ell, ell, n_pow=2.):
    """FFTLog samples xi_ell(r_j) of vega/pktoxi.py:262-275 for one multipole.

    ``pk_ell`` is the already prefactored multipole: for the standard transform
    (2 ell + 1) sum_mu dmu L_ell P times (-1)^(ell//2) / (2 pi^2); for 'rel'
    (n_pow=1) and 'asy' (n_pow=2) the isotropic linear P(k) itself.
    """
    k0, span, emm, r, dr, order = legacy_grid(k)
    mu = ell + 0.5
    q = 2 - n_pow - 0.5
    x = q + 2 * np.pi * 1j * emm / span
    lg1 = loggamma((mu + 1 + x) / 2)
    lg2 = loggamma((mu + 1 - x) / 2)
    um = (k0 * 1.)**(-2 * np.pi * 1j * emm / span) * 2**x * np.exp(lg1 - lg2)
    um[0] = np.real(um[0])
    an = np.fft.fft(pk_ell * k**n_pow * np.sqrt(np.pi / 2)) * um
    xi = np.fft.ifft(an)[order]
    xi = xi / r**(3 - n_pow)
    xi[-1] = 0
    return r, dr, np.real(xi)


def legacy_multipoles(ar, k, pk, ell_vals, muk, dmuk, tform=None):
    """Restatement of PktoXi.Pk2Mp (vega/pktoxi.py:230-279).

    Returns an array [len(ell_vals), len(ar)] filled at row ell//2.
    """
    from scipy.special import legendre
    out = np.zeros([len(ell_vals), len(ar)])
    for ell in ell_vals:
        if tform == 'rel':
            pk_ell, n_pow = pk, 1.
        elif tform == 'asy':
            pk_ell, n_pow = pk, 2.
        else:
            pk_ell = np.sum(dmuk * legendre(ell)(muk) * pk, axis=0) * (2 * ell + 1)
            pk_ell = pk_ell * (-1)**(ell // 2) / 2 / np.pi**2
            n_pow = 2.
        r, dr, xi = legacy_xi_samples(k, pk_ell, ell, n_pow)
        tck = splrep(np.log(r) - dr / 2, xi, k=3, s=0)
        out[ell // 2, :] = splev(np.log(ar), tck)
    return out
