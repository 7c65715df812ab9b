"""new_metals: metal matrices built from the stacked forest weights / the object catalogue
(vega/metals.py:83-108, 370-752), host side.

The O(N1 x N2) part -- every histogram the reference takes over the table of pixel pairs --
runs on the device through `vg_metal_pair_histograms` (csrc/vg_metalmat.cuh); what is left
here is reading the weight tables, the comoving distances, the normalisations and the small
transverse matrix.  There is no CPU pair loop in this module: without the CUDA library the
call fails (`_lib.load()`).

Two forms, as in the reference:
  * `rp_only_metal_mats = True`  -> `RpOnlyMatrix`: one (Np x Np) matrix applied to every rt
    column of xi (vega/metals.py:354-358).  The device keeps it in that form (Np^2 instead of
    (Np Nt)^2 numbers per metal correlation; csrc/vg_post.cuh `metal_rp_kernel`).
  * otherwise -> the dense (Np Nt)^2 product of the rp and rt matrices (:614-622), applied by
    the metal GEMM like a matrix read from a metal file.
"""
import ctypes as C

import numpy as np

from . import _lib, fitsio
from .coordinates import Coordinates
from .cosmo import ABSORBER_IGM


def _dptr(a):
    return a.ctypes.data_as(C.c_void_p)


class RpOnlyMatrix:
    """rp-only metal matrix: xi(rp, rt) -> sum_rp' A[rp, rp'] xi(rp', rt)."""

    def __init__(self, rp_mat, rt_nbins):
        self.rp_mat = np.ascontiguousarray(rp_mat, dtype=np.float64)
        self.rt_nbins = int(rt_nbins)
        n = self.rp_mat.shape[0] * self.rt_nbins
        self.shape = (n, n)

    def dot(self, xi):
        return (self.rp_mat @ np.asarray(xi).reshape(self.rp_mat.shape[1], self.rt_nbins)).ravel()


class MetalMatrixBuilder:
    """Per correlation: `matrix(abs1, abs2)` -> (matrix, Coordinates) for a metal pair."""

    def __init__(self, corr_item, wavelengths=None):
        if corr_item.cosmo is None:
            raise ValueError('new_metals needs the fiducial cosmology (OMEGAM in the data file header).')
        self.lib = _lib.load()
        self.item = corr_item
        self.cosmo = corr_item.cosmo
        self.wavelengths = wavelengths or ABSORBER_IGM
        self.cfg = corr_item.config['metal-matrix']
        self.coords = corr_item.model_coordinates
        self.rp_only = corr_item.config['model'].getboolean('rp_only_metal_mats', False)
        self.zmin = corr_item.config['data'].getfloat('zmin', 0.0)
        self.zmax = corr_item.config['data'].getfloat('zmax', 10.0)
        self.main = [corr_item.tracer1, corr_item.tracer2]
        self.signed = int(any(t['type'] == 'discrete' for t in self.main))
        self._tables = {}

    # -- weight tables (vega/metals.py:389-449) ------------------------------------------------
    def _forest(self, tracer):
        key = ('forest', tracer['weights-path'])
        if key not in self._tables:
            with fitsio.open(tracer['weights-path']) as hdul:
                wave = 10**np.asarray(hdul[1].data['LOGLAM'], dtype=float)
                weights = np.asarray(hdul[1].data['WEIGHT'], dtype=float)
            factor = self.cfg.getint('rebin_factor', None)
            if factor is not None:
                groups = wave.size // factor
                wave = wave[:groups * factor].reshape(groups, factor).mean(-1)
                weights = weights[:groups * factor].reshape(groups, factor).mean(-1)
            self._tables[key] = (wave, weights)
        return self._tables[key]

    def _objects(self, tracer):
        key = ('objects', tracer['weights-path'])
        if key not in self._tables:
            with fitsio.open(tracer['weights-path']) as hdul:
                z_cat = np.asarray(hdul[1].data['Z'], dtype=float)
            z_ref = self.cfg.getfloat('z_ref_objects', 2.25)
            z_evol = self.cfg.getfloat('z_evol_objects', 1.44)
            w_cat = ((1. + z_cat) / (1. + z_ref))**(z_evol - 1.)
            w, edges = np.histogram(z_cat, bins=self.cfg.getint('z_bins_objects', 1000), weights=w_cat)
            wz, _ = np.histogram(z_cat, bins=edges, weights=w_cat * z_cat)
            filled = w > 0
            self._tables[key] = (wz[filled] / w[filled], w[filled])
        return self._tables[key]

    def _side(self, which, true_abs):
        """[5][n]: r(true z), r(assumed z), true z, assumed z, weight (vega/metals.py:521-541)."""
        tracer = self.main[which]
        if tracer['type'] == 'continuous':
            wave, weights = self._forest(tracer)
            z_true = wave / self.wavelengths[true_abs] - 1.
            z_assumed = wave / self.wavelengths[tracer['name']] - 1.
            a_true = self.cfg.getfloat(f'alpha_{true_abs}')
            a_main = self.cfg.getfloat('alpha_{}'.format(tracer['name']), 2.9)
            weights = weights * (1 + z_true)**(a_true + a_main - 2)          # :480-500
        else:
            z_true, weights = self._objects(tracer)
            z_assumed = z_true
        if np.any(z_true < 0) or np.any(z_assumed < 0):
            raise ValueError('Attempting to compute distance to a negative redshift')
        return np.ascontiguousarray(np.stack([
            self.cosmo.get_r_comov(z_true), self.cosmo.get_r_comov(z_assumed), z_true, z_assumed,
            weights * np.ones_like(z_true)]), dtype=np.float64)

    # -- the device pass ---------------------------------------------------------------------
    def _histograms(self, abs1, abs2, with_ratio):
        c = self.coords
        s1, s2 = self._side(0, abs1), self._side(1, abs2)
        n1, n2 = s1.shape[1], s2.shape[1]
        nb = c.rp_nbins
        rp_edges = np.linspace(c.rp_min, c.rp_max, nb + 1)
        ratio_edges, nr = None, 0
        if with_ratio:
            nr = 4 * (c.rt_nbins + 1)
            lim = np.zeros(2)
            _lib.check(self.lib.vg_metal_pair_ratio_range(_dptr(s1), n1, _dptr(s2), n2, _dptr(lim)))
            lo, hi = lim
            if lo == hi:                      # numpy's rule for a degenerate range
                lo, hi = lo - 0.5, hi + 0.5
            ratio_edges = np.linspace(lo, hi, nr + 1)
        H, per_true = np.zeros((nb, nb)), np.zeros(nb)
        per_assumed, ratio = np.zeros((3, nb)), np.zeros(max(nr, 1))
        _lib.check(self.lib.vg_metal_pair_histograms(
            _dptr(s1), n1, _dptr(s2), n2, self.signed, self.zmin, self.zmax, _dptr(rp_edges), nb,
            _dptr(ratio_edges) if nr else None, nr, _dptr(H), _dptr(per_true), _dptr(per_assumed),
            _dptr(ratio)))
        w = per_assumed[0]
        safe = w + (w == 0)
        return H, per_true, per_assumed[1] / safe, per_assumed[2] / safe, ratio_edges, ratio[:nr]

    def matrix(self, abs1, abs2):
        c = self.coords
        nt = c.rt_nbins
        H, per_true, rp_eff, z_eff, ratio_edges, ratio_w = self._histograms(abs1, abs2, not self.rp_only)
        if self.rp_only:                                                   # :698-751
            mat = RpOnlyMatrix(H * ((per_true > 0) / (per_true + (per_true == 0)))[None, :], nt)
            rt_eff = np.arange(c.rt_binsize / 2, c.rt_max, c.rt_binsize)
        else:                                                              # :556-644
            norm = H.sum(axis=0)
            along = H / (norm + (norm == 0))
            mat = np.einsum('ij,kl->ikjl', along, self._transverse(ratio_edges, ratio_w)).reshape(
                c.rp_nbins * nt, c.rp_nbins * nt)
            lo, hi = np.arange(nt) * c.rt_max / nt, (1 + np.arange(nt)) * c.rt_max / nt
            rt_eff = (2 * (hi**3 - lo**3)) / (3 * (hi**2 - lo**2))
        grids = Coordinates.init_from_grids(
            c, np.repeat(rp_eff, nt), np.tile(rt_eff, c.rp_nbins), np.repeat(z_eff, nt))
        return mat, grids

    def _transverse(self, ratio_edges, ratio_w):
        """rt migration matrix from the distance-ratio histogram (vega/metals.py:585-610)."""
        c = self.coords
        nt = c.rt_nbins
        rt_edges = np.linspace(0, c.rt_max, nt + 1)
        ratios = (ratio_edges[1:] + ratio_edges[:-1]) / 2
        centres = (rt_edges[:-1] + rt_edges[1:]) / 2
        half = c.rt_binsize / 2
        oversample = 7
        offsets = np.linspace(-half, half * (1 - 2 / oversample), oversample)
        across = np.zeros((nt, nt))
        for i, rt in enumerate(centres):
            sub = rt + offsets
            across[:, i], _ = np.histogram(np.outer(ratios, sub).ravel(), bins=rt_edges,
                                           weights=np.outer(ratio_w, sub).ravel())
        norm = across.sum(axis=0)
        return across / (norm + (norm == 0))


def build_new_metals(corr_item, data, wavelengths=None):
    """Fill data.metal_coordinates / data.metal_mats for every metal correlation of a
    new_metals configuration (what Metals.__init__ does at vega/metals.py:94-108)."""
    builder = MetalMatrixBuilder(corr_item, wavelengths)
    for pair in corr_item.metal_correlations:
        mat, grids = builder.matrix(*pair)
        data.metal_mats[pair] = mat
        data.metal_coordinates[pair] = grids
