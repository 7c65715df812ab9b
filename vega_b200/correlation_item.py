"""Per-correlation configuration holder (init-time plumbing).

Follows the attribute surface of the reference's vega/correlation_item.py:8-140
that the data reader, the plan builder and the oracle consume: tracer names and
types, the parsed ini sections, coordinate grids, the metal-correlation list.
plus the small-scale marginalisation templates (vega/correlation_item.py:169-275).
"""
import numpy as np


class CorrelationItem:
    cosmo = None
    model_coordinates = None
    dist_model_coordinates = None
    data_coordinates = None
    low_mem_mode = False

    def __init__(self, config, model_pk=False):
        self.config = config
        self.model_pk = model_pk
        data_cfg = config['data']
        model_cfg = config['model']
        self.name = data_cfg.get('name')
        self.tracer1 = {'name': data_cfg.get('tracer1'), 'type': data_cfg.get('tracer1-type')}
        self.tracer2 = {'name': data_cfg.get('tracer2', self.tracer1['name']),
                        'type': data_cfg.get('tracer2-type', self.tracer1['type'])}

        self.cov_rescale = data_cfg.getfloat('cov_rescale', None)
        self.has_distortion = data_cfg.getboolean('distortion', True)
        self.has_data = data_cfg.getboolean('has_datafile', True)
        if 'filename' not in data_cfg:
            self.has_data = False
        self.test_flag = data_cfg.getboolean('test', False)

        self.new_metals = model_cfg.getboolean('new_metals', False)
        if self.new_metals:      # vega/correlation_item.py:44-49
            self.tracer1['weights-path'] = data_cfg.get('weights-tracer1')
            self.tracer2['weights-path'] = data_cfg.get('weights-tracer2', None)
            if self.tracer2['weights-path'] is None:
                self.tracer2['weights-path'] = self.tracer1['weights-path']

        self.marginalize_small_scales = {}
        for key, name in (('marginalize-below-rtmax', 'rtmax'), ('marginalize-above-rtmin', 'rtmin'),
                          ('marginalize-below-rpmax', 'rpmax'), ('marginalize-above-rpmin', 'rpmin')):
            value = model_cfg.getfloat(key, 0)
            if value > 0:
                self.marginalize_small_scales[name] = value
        if model_cfg.getboolean('marginalize-all-rmin-cuts', False):
            self.marginalize_small_scales['all-rmin'] = True
        self.marginalize_small_scales_prior_sigma = model_cfg.getfloat(
            'marginalize-prior-sigma', 10.0)
        self.marginalize_match_data_bins = model_cfg.getboolean(
            'marginalize-match-data-bins', False)
        self.fit_marg_scales = model_cfg.getboolean('fit-marginalized-scales', False)

        self.has_metals = False
        self.has_bb = False
        self.tracer_catalog = {}
        self.metal_correlations = []

    def init_metals(self, tracer_catalog, metal_correlations):
        """Canonical, de-duplicated metal-correlation list with the main tracers kept
        in their slot (vega/correlation_item.py:76-105)."""
        self.tracer_catalog = tracer_catalog
        self.metal_correlations = []
        for corr in metal_correlations:
            pair = tuple(sorted([corr[0], corr[1]]))
            if pair[0] == self.tracer2['name'] or pair[1] == self.tracer1['name']:
                pair = (pair[1], pair[0])
            if pair not in self.metal_correlations:
                self.metal_correlations.append(pair)
        self.has_metals = True

    def init_broadband(self, coeff_binning_model):
        self.coeff_binning_model = coeff_binning_model
        self.has_bb = True

    def init_coordinates(self, model_coordinates, dist_model_coordinates=None,
                         data_coordinates=None):
        self.model_coordinates = model_coordinates
        self.data_coordinates = model_coordinates if data_coordinates is None \
            else data_coordinates
        self.dist_model_coordinates = model_coordinates if dist_model_coordinates is None \
            else dist_model_coordinates

    def init_cosmo(self, cosmo_params):
        """vega/correlation_item.py:138-151; the cosmology object is vega_b200/cosmo.py's
        restatement of picca's (used by new_metals)."""
        from .cosmo import Cosmo
        self.cosmo_params = cosmo_params
        self.cosmo = Cosmo(Om=cosmo_params['Omega_m'], Ok=cosmo_params['Omega_k'],
                           Or=cosmo_params['Omega_r'], wl=cosmo_params['wl'])

    def check_if_blind_corr(self, blind_tracers):
        if 'all' in blind_tracers:
            return True
        return any(t in self.tracer1['name'] or t in self.tracer2['name']
                   for t in blind_tracers)

    # -- small-scale marginalisation (vega/correlation_item.py:175-275) ------------------------
    def get_undist_xi_marg_templates(self):
        """Sparse [n_model, n_templates] matrix of indicator templates on the UNdistorted model
        grid: one column per marginalised bin, or one per nearest distorted bin with
        marginalize-match-data-bins.  The prior sigma is applied by the caller
        (Data.get_dist_xi_marg_templates)."""
        from scipy.sparse import csr_array
        mc, cuts = self.model_coordinates, self.marginalize_small_scales
        n = mc.rt_regular_grid.size
        if 'all-rmin' in cuts:
            # everything the small-scale part of the scale cuts removes, evaluated on the
            # distorted grid and blown up to the finer model grid
            dist = self.dist_model_coordinates
            cb = mc.rp_nbins // dist.rp_nbins
            keep_dist = dist.get_mask_scale_cuts(self.config['cuts'], small_scale_mask=True) \
                .reshape(dist.rp_nbins, dist.rt_nbins)
            keep = np.zeros((mc.rp_nbins, mc.rt_nbins), dtype=bool)
            block = np.kron(keep_dist, np.ones((cb, cb), dtype=bool))
            keep[:block.shape[0], :block.shape[1]] = block[:mc.rp_nbins, :mc.rt_nbins]
            selected = ~keep.ravel()
            print(f'Marginalizing distortion scales with {int(selected.sum())} points based on scale cuts.')
        else:
            selected = np.ones(n, dtype=bool)
            if 'rtmax' in cuts:
                selected &= mc.rt_regular_grid < cuts['rtmax']
            if 'rtmin' in cuts:
                selected &= mc.rt_regular_grid > cuts['rtmin']
            if 'rpmax' in cuts:
                selected &= np.abs(mc.rp_regular_grid) < cuts['rpmax']
            if 'rpmin' in cuts:
                selected &= np.abs(mc.rp_regular_grid) > cuts['rpmin']
            if not selected.any():
                raise ValueError('No common indices found for small-scale marginalization templates.')
        bins = np.nonzero(selected)[0]
        if self.marginalize_match_data_bins:
            dist = self.dist_model_coordinates
            d2 = (dist.rp_grid[None, :] - mc.rp_grid[bins][:, None])**2 \
                + (dist.rt_grid[None, :] - mc.rt_grid[bins][:, None])**2
            nearest = d2.argmin(axis=1)
            columns = np.searchsorted(np.unique(nearest), nearest)
            n_templates = np.unique(nearest).size
        else:
            columns, n_templates = np.arange(bins.size), bins.size
        return csr_array((np.ones(bins.size), (bins, columns)), shape=(n, n_templates))
