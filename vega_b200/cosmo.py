"""Fiducial cosmology and rest wavelengths used by the new_metals matrices
(vega/correlation_item.py:138-151, vega/metals.py:443-470,523-541).

The reference takes both from ``picca.constants`` (``Cosmo`` and ``ABSORBER_IGM``), an
un-vendored dependency that is absent from /root/reference and from this image.  PARITY
UNPINNED for this module: it restates picca's published algorithm -- H(z) of a
w-CDM model in h units, the comoving distance as a trapezoid sum on a fixed grid of 10 000
redshifts up to z = 10, linear interpolation in between -- and the rest wavelengths in
Angstrom of the transitions vega's configurations name.  Everything built on top of it
(vega/metals.py:372-752) is pinned against the unmodified reference run with this module
standing in for picca.constants (tests/golden/make_golden.py).
"""
import numpy as np

SPEED_LIGHT = 299792.458   # km / s

# rest-frame wavelengths [Angstrom] of the absorbers that appear in vega's ini files
ABSORBER_IGM = {
    'LYA': 1215.67,
    'LYB': 1025.72,
    'LY3': 972.537,
    'LY4': 949.7431,
    'LY5': 937.8035,
    'SiII(1260)': 1260.4221,
    'SiIII(1207)': 1206.500,
    'SiII(1193)': 1193.2897,
    'SiII(1190)': 1190.4158,
    'SiII(1304)': 1304.3702,
    'SiII(1527)': 1526.70698,
    'SiIV(1394)': 1393.76018,
    'SiIV(1403)': 1402.77291,
    'CII(1335)': 1334.5323,
    'CII(1037)': 1036.7909,
    'CII(1036)': 1036.3367,
    'CIV(1548)': 1548.2049,
    'CIV(1551)': 1550.77845,
    'CIV(eff)': 1549.06,
    'NV(1239)': 1238.821,
    'NV(1243)': 1242.804,
    'OI(1302)': 1302.1685,
    'OVI(1032)': 1031.912,
    'OVI(1038)': 1037.613,
    'MgII(2796)': 2796.3511,
    'MgII(2804)': 2803.5324,
}


class Cosmo:
    """Distances of a w-CDM model in Mpc/h: get_r_comov, get_dist_m, get_dist_hubble, get_hubble."""

    def __init__(self, Om, Ok=0., Or=0., wl=-1., verbose=False, nbins=10000, zmax=10.):
        self.Om, self.Ok, self.Or, self.wl = Om, Ok, Or, wl
        Ol = 1. - Ok - Om - Or
        H0 = 100.
        dz = zmax / nbins
        z = np.arange(nbins) * dz
        hubble = H0 * np.sqrt(Ol * (1. + z)**(3. * (1. + wl)) + Ok * (1. + z)**2
                              + Om * (1. + z)**3 + Or * (1. + z)**4)
        inv = SPEED_LIGHT / hubble
        r_comov = np.concatenate([[0.], np.cumsum((inv[:-1] + inv[1:]) / 2. * dz)])
        if Ok == 0.:
            dist_m = r_comov
        elif Ok < 0.:
            scale = H0 * np.sqrt(-Ok) / SPEED_LIGHT
            dist_m = np.sin(scale * r_comov) / scale
        else:
            scale = H0 * np.sqrt(Ok) / SPEED_LIGHT
            dist_m = np.sinh(scale * r_comov) / scale
        self._z, self._hubble, self._r_comov, self._dist_m = z, hubble, r_comov, dist_m

    def _interp(self, z, table):
        z = np.asarray(z, dtype=float)
        if np.any(z < self._z[0]) or np.any(z > self._z[-1]):
            raise ValueError('redshift outside the tabulated range [0, {:.3f}]'.format(self._z[-1]))
        return np.interp(z, self._z, table)

    def get_r_comov(self, z):
        return self._interp(z, self._r_comov)

    def get_dist_m(self, z):
        return self._interp(z, self._dist_m)

    def get_hubble(self, z):
        return self._interp(z, self._hubble)

    def get_dist_hubble(self, z):
        return self._interp(z, SPEED_LIGHT / self._hubble)
