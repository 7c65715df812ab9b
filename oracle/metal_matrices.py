"""TEST INFRASTRUCTURE ONLY -- CPU restatement (numpy) of the reference's new_metals matrices.

Follows vega/metals.py:370-752 (`rebin`, `get_forest_weights`, `get_qso_weights`,
`get_rp_pairs`, `get_forest_weight_scaling`, `compute_metal_dmat`, `compute_metal_rp_dmat`):
a metal absorber seen at observed wavelength lambda is assigned the redshift of the main
tracer's transition, so a pair of pixels at true separation rp_true is binned at
rp_assumed; the matrix is the weighted 2-D histogram of (assumed, true) separations over all
pixel pairs of the two stacked-weight tables, normalised per true bin.

Pinned against the unmodified reference run here with vega_b200/cosmo.py standing in for
picca.constants (tests/golden/make_golden.py, key `new_metals`).  Only tests/, smoke() and
the CPU-baseline leg of bench.py may import this module; the product path is
vega_b200/metal_matrices.py (CUDA pair histogram).
"""
import numpy as np


def rebin(vector, factor):
    """Mean over consecutive groups of `factor` entries, tail dropped (vega/metals.py:370-387)."""
    n = vector.size // factor
    return vector[:n * factor].reshape(n, factor).mean(axis=1)


class TracerSide:
    """Redshifts and weights of one side of the pair table for a (true absorber, main tracer)
    combination (vega/metals.py:521-541)."""

    def __init__(self, true_z, assumed_z, weights):
        self.true_z, self.assumed_z, self.weights = true_z, assumed_z, weights


class MetalMatrixOracle:
    """`fits_open(path)[1].data[column]` reads the weight tables; `cfg` is the per-correlation
    ConfigParser; `coords` the model Coordinates; `cosmo` has get_r_comov(z)."""

    def __init__(self, corr_item, cosmo, wavelengths, fits_open):
        self.item = corr_item
        self.cosmo = cosmo
        self.wavelengths = wavelengths
        self.fits_open = fits_open
        self.cfg = corr_item.config['metal-matrix']
        self.coords = corr_item.model_coordinates
        self.zmin = corr_item.config['data'].getfloat('zmin', 0.0)
        self.zmax = corr_item.config['data'].getfloat('zmax', 10.0)
        self.main = [corr_item.tracer1, corr_item.tracer2]
        self.signed = any(t['type'] == 'discrete' for t in self.main)

    # vega/metals.py:389-415
    def forest_table(self, tracer):
        assert tracer['type'] == 'continuous'
        table = self.fits_open(tracer['weights-path'])[1].data
        wave, weights = 10**np.asarray(table['LOGLAM'], float), np.asarray(table['WEIGHT'], float)
        factor = self.cfg.getint('rebin_factor', None)
        if factor is not None:
            wave, weights = rebin(wave, factor), rebin(weights, factor)
        return wave, weights

    # vega/metals.py:417-449
    def object_table(self, tracer):
        assert tracer['type'] == 'discrete'
        z_cat = np.asarray(self.fits_open(tracer['weights-path'])[1].data['Z'], float)
        z_ref = self.cfg.getfloat('z_ref_objects', 2.25)
        z_evol = self.cfg.getfloat('z_evol_objects', 1.44)
        nbins = self.cfg.getint('z_bins_objects', 1000)
        w_cat = ((1. + z_cat) / (1. + z_ref))**(z_evol - 1.)
        sum_w, edges = np.histogram(z_cat, bins=nbins, weights=w_cat)
        sum_wz, _ = np.histogram(z_cat, bins=edges, weights=w_cat * z_cat)
        keep = sum_w > 0
        return sum_wz[keep] / sum_w[keep], sum_w[keep]

    # vega/metals.py:480-500 and :521-541
    def side(self, which, true_abs):
        tracer = self.main[which]
        if tracer['type'] == 'continuous':
            wave, weights = self.forest_table(tracer)
            true_z = wave / self.wavelengths[true_abs] - 1.
            assumed_z = wave / self.wavelengths[tracer['name']] - 1.
            alpha_true = self.cfg.getfloat(f'alpha_{true_abs}')
            alpha_main = self.cfg.getfloat('alpha_{}'.format(tracer['name']), 2.9)
            weights = weights * (1 + true_z)**(alpha_true + alpha_main - 2)
            return TracerSide(true_z, assumed_z, weights)
        z, weights = self.object_table(tracer)
        return TracerSide(z, z, weights * 1.)

    # vega/metals.py:451-478
    def separations(self, z1, z2):
        if np.any(z1 < 0) or np.any(z2 < 0):
            raise ValueError('Attempting to compute distance to a negative redshift')
        r1, r2 = self.cosmo.get_r_comov(z1), self.cosmo.get_r_comov(z2)
        rp = (r1[:, None] - r2[None, :]).ravel()
        if not self.signed:
            rp = np.abs(rp)
        return rp, ((r1[:, None] + r2[None, :]) / 2).ravel()

    def pair_table(self, abs1, abs2):
        s1, s2 = self.side(0, abs1), self.side(1, abs2)
        rp_true, dist_true = self.separations(s1.true_z, s2.true_z)
        rp_assumed, dist_assumed = self.separations(s1.assumed_z, s2.assumed_z)
        weights = (s1.weights[:, None] * s2.weights[None, :]).ravel()
        z_pair = (s1.assumed_z[:, None] + s2.assumed_z[None, :]) / 2.
        weights = weights * ((z_pair >= self.zmin) & (z_pair <= self.zmax)).ravel()
        z_true_pair = ((s1.true_z[:, None] + s2.true_z[None, :]) / 2.).ravel()
        return rp_true, dist_true, rp_assumed, dist_assumed, weights, z_true_pair

    def effective_rp_z(self, rp_assumed, weights, z_true_pair, edges):
        """Weighted mean assumed rp and true redshift per assumed-rp bin (:625-637, :716-728)."""
        w, _ = np.histogram(rp_assumed, bins=edges, weights=weights)
        w_rp, _ = np.histogram(rp_assumed, bins=edges, weights=weights * rp_assumed)
        w_z, _ = np.histogram(rp_assumed, bins=edges, weights=weights * z_true_pair)
        safe = w + (w == 0)
        return w_rp / safe, w_z / safe

    # vega/metals.py:679-752
    def rp_matrix(self, abs1, abs2):
        c = self.coords
        rp_true, _, rp_assumed, _, weights, z_true_pair = self.pair_table(abs1, abs2)
        edges = np.linspace(c.rp_min, c.rp_max, c.rp_nbins + 1)
        mat, _, _ = np.histogram2d(rp_assumed, rp_true, bins=(edges, edges), weights=weights)
        per_true, _ = np.histogram(rp_true, bins=edges, weights=weights)
        mat = mat * ((per_true > 0) / (per_true + (per_true == 0)))[None, :]
        rp_eff, z_eff = self.effective_rp_z(rp_assumed, weights, z_true_pair, edges)
        rt_centres = np.arange(c.rt_binsize / 2, c.rt_max, c.rt_binsize)
        # flat index = rt_index + rt_nbins * rp_index
        return (mat, np.repeat(rp_eff, c.rt_nbins), np.tile(rt_centres, c.rp_nbins),
                np.repeat(z_eff, c.rt_nbins))

    # vega/metals.py:502-677
    def full_matrix(self, abs1, abs2):
        c = self.coords
        rp_true, dist_true, rp_assumed, dist_assumed, weights, z_true_pair = self.pair_table(abs1, abs2)
        rp_edges = np.linspace(c.rp_min, c.rp_max, c.rp_nbins + 1)
        along, _, _ = np.histogram2d(rp_assumed, rp_true, bins=(rp_edges, rp_edges), weights=weights)
        norm = along.sum(axis=0)
        along = along / (norm + (norm == 0))

        # transverse: histogram of the distance ratio, weighted by solid angle, restricted to
        # pairs that are truly close along the line of sight (:566-584)
        rt_edges = np.linspace(0, c.rt_max, c.rt_nbins + 1)
        ratio_w, ratio_edges = np.histogram(
            dist_assumed / dist_true, bins=4 * rt_edges.size,
            weights=weights / dist_true**2 * (np.abs(rp_true) < 20.))
        ratios = (ratio_edges[1:] + ratio_edges[:-1]) / 2
        centres = (rt_edges[:-1] + rt_edges[1:]) / 2
        half = c.rt_binsize / 2
        oversample = 7
        offsets = np.linspace(-half, half * (1 - 2 / oversample), oversample)
        across = np.zeros((c.rt_nbins, c.rt_nbins))
        for i, rt in enumerate(centres):          # (:598-606)
            sub = rt + offsets
            across[:, i], _ = np.histogram((ratios[:, None] * sub[None, :]).ravel(), bins=rt_edges,
                                           weights=(ratio_w[:, None] * sub[None, :]).ravel())
        norm = across.sum(axis=0)
        across = across / (norm + (norm == 0))

        n = c.rp_nbins * c.rt_nbins
        mat = np.einsum('ij,kl->ikjl', along, across).reshape(n, n)      # (:614-622)
        rp_eff, z_eff = self.effective_rp_z(rp_assumed, weights, z_true_pair, rp_edges)
        lo = np.arange(c.rt_nbins) * c.rt_max / c.rt_nbins
        hi = (1 + np.arange(c.rt_nbins)) * c.rt_max / c.rt_nbins
        rt_eff = (2 * (hi**3 - lo**3)) / (3 * (hi**2 - lo**2))           # solid-angle mean (:639-644)
        return (mat, np.repeat(rp_eff, c.rt_nbins), np.tile(rt_eff, c.rp_nbins),
                np.repeat(z_eff, c.rt_nbins))
