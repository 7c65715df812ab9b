"""Coordinate grids of one correlation (static per-bin tables uploaded once).

Same attribute surface as the reference's vega/coordinates.py:4-217 (rp/rt/r/mu/z
grids, regular-grid centres, bin sizes, scale-cut masks) because Data, the
broadband design tables and the plan builder read those names.
"""
import numpy as np


class Coordinates:
    def __init__(self, rp_min, rp_max, rt_max, rp_nbins, rt_nbins,
                 rp_grid=None, rt_grid=None, z_grid=None, z_eff=None,
                 r_grid=None, mu_grid=None):
        self.rp_min, self.rp_max, self.rt_max = rp_min, rp_max, rt_max
        self.rp_nbins, self.rt_nbins = rp_nbins, rt_nbins
        self.rp_binsize = (rp_max - rp_min) / rp_nbins
        self.rt_binsize = rt_max / rt_nbins

        # Bin centres of the regular grid, rp-major (vega/coordinates.py:43-48)
        rp_centres = np.arange(rp_min + self.rp_binsize / 2, rp_max, self.rp_binsize)
        rt_centres = np.arange(self.rt_binsize / 2, rt_max, self.rt_binsize)
        rt_mesh, rp_mesh = np.meshgrid(rt_centres, rp_centres)
        self.rp_regular_grid = rp_mesh.flatten()
        self.rt_regular_grid = rt_mesh.flatten()
        self.r_regular_grid = np.sqrt(self.rp_regular_grid**2 + self.rt_regular_grid**2)
        self.mu_regular_grid = self._safe_ratio(self.rp_regular_grid, self.r_regular_grid)

        self.rp_grid = self.rp_regular_grid if rp_grid is None else np.asarray(rp_grid, float)
        self.rt_grid = self.rt_regular_grid if rt_grid is None else np.asarray(rt_grid, float)
        if r_grid is None:
            self.r_grid = np.sqrt(self.rp_grid**2 + self.rt_grid**2)
        else:
            self.r_grid = np.asarray(r_grid, float)
        if mu_grid is None:
            self.mu_grid = self._safe_ratio(self.rp_grid, self.r_grid)
        else:
            self.mu_grid = np.asarray(mu_grid, float)

        if z_grid is None and z_eff is None:
            self.z_grid = None
        else:
            self.z_grid = z_eff if z_grid is None else np.asarray(z_grid, float)

    @staticmethod
    def _safe_ratio(num, den):
        out = np.zeros_like(den)
        w = den > 0.
        out[w] = num[w] / den[w]
        return out

    @classmethod
    def init_from_grids(cls, other, rp_grid, rt_grid, z_grid):
        return cls(other.rp_min, other.rp_max, other.rt_max, other.rp_nbins, other.rt_nbins,
                   rp_grid=rp_grid, rt_grid=rt_grid, z_grid=z_grid)

    @classmethod
    def init_from_r_mu_grids(cls, r_grid, mu_grid, z_eff=None):
        if len(r_grid) != len(mu_grid):
            raise ValueError(
                'r_grid and mu_grid must either be on a meshgrid or have the same size')
        rp_grid = r_grid * mu_grid
        rt_grid = r_grid * np.sqrt(1 - mu_grid**2)
        return cls(rp_min=rp_grid.min(), rp_max=rp_grid.max(), rt_max=rt_grid.max(),
                   rp_nbins=len(r_grid), rt_nbins=len(r_grid), rp_grid=rp_grid, rt_grid=rt_grid,
                   r_grid=r_grid, mu_grid=mu_grid, z_eff=z_eff)

    def get_mask_to_other(self, other):
        assert self.rp_binsize == other.rp_binsize
        assert self.rt_binsize == other.rt_binsize
        mask = (self.rp_grid >= other.rp_min) & (self.rp_grid <= other.rp_max)
        return mask & (self.rt_grid <= other.rt_max)

    def get_mask_scale_cuts(self, cuts_config, small_scale_mask=False):
        """Strict inequalities on the REGULAR grid centres (vega/coordinates.py:146-182)."""
        def cut(key, default):
            return cuts_config.getfloat(key, default)

        mask = (self.rp_regular_grid > cut('rp-min', 0.)) \
            & (self.rt_regular_grid > cut('rt-min', 0.))
        mask &= self.r_regular_grid > cut('r-min', 10.)
        if small_scale_mask:
            return mask
        mask &= (self.rp_regular_grid < cut('rp-max', 300.)) \
            & (self.rt_regular_grid < cut('rt-max', 300.))
        mask &= self.r_regular_grid < cut('r-max', 180.)
        mask &= (self.mu_regular_grid > cut('mu-min', -1.)) \
            & (self.mu_regular_grid < cut('mu-max', +1.))
        return mask

    def get_mask_marginalization_scales(self, cuts_config, marginalization_cuts):
        mask = np.ones_like(self.rp_regular_grid, dtype=bool)
        if 'rtmax' in marginalization_cuts:
            mask &= self.rt_regular_grid < marginalization_cuts['rtmax']
        if 'rtmin' in marginalization_cuts:
            mask &= self.rt_regular_grid > marginalization_cuts['rtmin']
        if 'rpmax' in marginalization_cuts:
            mask &= np.abs(self.rp_regular_grid) < marginalization_cuts['rpmax']
        if 'rpmin' in marginalization_cuts:
            mask &= np.abs(self.rp_regular_grid) > marginalization_cuts['rpmin']
        if 'all-rmin' in marginalization_cuts:
            mask = ~self.get_mask_scale_cuts(cuts_config, small_scale_mask=True)
        return mask
