"""Static inputs of one correlation: data vector, masks, covariance, distortion
and metal matrices (init-time; these arrays are uploaded to HBM once per plan).

Attribute names follow the reference's vega/data.py (``distortion_mat``,
``inv_masked_cov``, ``masked_data_vec``, ``model_mask``, ``data_mask``,
``metal_mats``, ``metal_coordinates``, ``log_cov_det``, ``data_size`` ...) because
callers (run_vega_mc scripts, Analysis, Output) and the plan builder read them.
Differences from the reference, all init-time representation choices:
  * matrices are kept DENSE float64 (the device applies them as FP64 GEMMs over the
    parameter batch; the reference's CSR mat-vec, vega/data.py:342-346, is the CPU form);
  * a missing distortion / metal matrix is recorded as ``None`` plus an identity flag
    instead of materialising ``np.eye`` (vega/data.py:77-80,683-684) -- the model is
    unchanged, the device skips the product.
"""
import numpy as np

from . import fitsio, utils
from .coordinates import Coordinates
from .utils import find_file, compute_masked_invcov, compute_log_cov_det

BLINDING_STRATEGIES = ['desi_dr3']


class Data:
    def __init__(self, corr_item, marginalize_in_fit=False, metal_builder=None):
        self.corr_item = corr_item
        self.tracer1 = corr_item.tracer1
        self.tracer2 = corr_item.tracer2
        cfg = corr_item.config
        self.use_metal_autos = cfg['model'].getboolean('use_metal_autos', True)
        self.cholesky_masked_cov = cfg['data'].getboolean('cholesky-masked-cov', True)

        self._data_vec = None
        self._masked_data_vec = None
        self._cov_mat = None
        self._distortion_mat = None
        self._inv_masked_cov = None
        self._log_cov_det = None
        self._blind = None
        self._blinding_strat = None
        self.cosmo_params = None
        self.model_coordinates = None
        self.dist_model_coordinates = None
        self.data_coordinates = None
        self.nb = None
        self.metal_mats = {}
        self.metal_coordinates = {}

        self._read_data(cfg['data'].get('filename'), cfg['cuts'],
                        cfg['data'].get('distortion-file', None),
                        cfg['data'].get('covariance-file', None),
                        cfg['data'].getfloat('cov_rescale', None))
        corr_item.init_coordinates(
            self.model_coordinates, self.dist_model_coordinates, self.data_coordinates)

        if 'metals' in cfg:
            if corr_item.new_metals:      # no metal file: vega/data.py:60-66
                tracer_catalog, metal_correlations = self._metal_tracers_and_pairs(cfg['metals'])
            else:
                tracer_catalog, metal_correlations = self._init_metals(cfg['metals'])
            corr_item.init_metals(tracer_catalog, metal_correlations)
        if 'broadband' in cfg:
            corr_item.init_broadband(self.coeff_binning_model)
        if self.cosmo_params is not None:
            corr_item.init_cosmo(self.cosmo_params)
        if 'metals' in cfg and corr_item.new_metals:
            # metal matrices and coordinates from the stacked weights (the reference does this in
            # Metals.__init__, vega/metals.py:94-108).  `metal_builder` is a hook for callers that
            # bring their own matrices; the package's builder runs the pair histograms on the device
            if metal_builder is None:
                from .metal_matrices import build_new_metals as metal_builder
            metal_builder(corr_item, self)

        # Identity stand-ins (vega/data.py:77-80): kept implicit
        self.distortion_is_identity = self._distortion_mat is None
        self.cov_is_identity = self._cov_mat is None
        if self.cov_is_identity:
            self._cov_mat = np.eye(self.full_data_size)
        self.variance = self.cov_mat.diagonal()

        self.marg_templates = None
        self.cov_marg_update = None
        self.marg_diff2coeff_matrix = None
        self.num_marg_modes = 0
        if corr_item.marginalize_small_scales:
            self._init_marginalization(marginalize_in_fit)

        self._cholesky = None
        self._scale = 1.
        self.scaled_inv_masked_cov = None
        self.scaled_log_cov_det = None
        self.mc_mock = None
        self.masked_mc_mock = None
        self.effective_data_size = self.data_size - self.num_marg_modes

    # -- small-scale marginalisation (vega/data.py:96-128,762-828) ------------------------------
    def _init_marginalization(self, marginalize_in_fit):
        """Templates for the marginalised bins pushed through the distortion matrix.  Default
        mode: their prior covariance sigma^2 T T^T (SVD-compressed) is added to the masked
        covariance, so the chi2 path is unchanged.  marginalize-in-fit: the covariance is left
        alone and `marg_diff2coeff_matrix` = (T^T C^-1 T + S)^-1 T^T C^-1 gives the best-fit
        template amplitudes from a residual (used by VegaInterface.chi2)."""
        item = self.corr_item
        self.marg_templates, self.cov_marg_update = self.get_dist_xi_marg_templates()
        ntemps = self.marg_templates.shape[1]
        self._inv_masked_cov = None
        inv_cov = self.inv_masked_cov            # of the covariance BEFORE the update
        self._inv_masked_cov = None
        if not marginalize_in_fit:
            self._cov_mat = np.array(self._cov_mat, dtype=float)
            self._cov_mat[np.ix_(self.data_mask, self.data_mask)] += self.cov_marg_update
            self.cov_is_identity = False
            self._log_cov_det = None
            self.variance = self._cov_mat.diagonal()
        else:
            self.cov_marg_update = None
        t_masked = self.marg_templates[self.model_mask, :]
        prior = None
        if not (item.fit_marg_scales and item.marginalize_match_data_bins):
            prior = np.full(ntemps, item.marginalize_small_scales_prior_sigma**-2)
        self.marg_diff2coeff_matrix = utils.diff_to_coeff_matrix(t_masked, inv_cov, prior)

    def marg_projector(self):
        """Q = I - T_masked D2C on the masked bins: with the template amplitudes fitted on the fly
        (vega/vega_interface.py:282-292) the residual entering chi2 is Q (d - m), so the
        quadratic form is (d - m)^T [Q^T C^-1 Q] (d - m) -- one modified matrix for the device."""
        if self.marg_diff2coeff_matrix is None:
            return None
        t_masked = self.marg_templates[self.model_mask, :]
        return np.eye(int(self.model_mask.sum())) - np.asarray(t_masked @ self.marg_diff2coeff_matrix)

    def get_dist_xi_marg_templates(self, factor=1e-8, return_AAT=True):
        """(templates [n_dist, n_templates] sparse, covariance update on the masked bins)."""
        from scipy.sparse import csr_array
        item = self.corr_item
        if not item.marginalize_small_scales:
            raise ValueError('Marginalization not configured')
        templates = item.get_undist_xi_marg_templates()
        if not self.distortion_is_identity:
            templates = csr_array(self._distortion_mat @ templates)
        if item.fit_marg_scales:
            self.data_mask = self.data_mask | self.data_coordinates.get_mask_marginalization_scales(
                item.config['cuts'], item.marginalize_small_scales)
            self.model_mask = self.model_mask | self.dist_model_coordinates.get_mask_marginalization_scales(
                item.config['cuts'], item.marginalize_small_scales)
            if self.data_mask.sum() != self.model_mask.sum():
                raise ValueError(
                    'Data and model masks should be the same after marginalization scale cuts.')
            self._masked_data_vec = None
        if not return_AAT:
            return templates
        t = (templates * item.marginalize_small_scales_prior_sigma)[self.model_mask, :].toarray()
        print(f'  There are {templates.shape[1]} templates. '
              'SVD of template matrix to remove degenerate modes.')
        update, self.num_marg_modes = utils.template_covariance(t, factor)
        print(f'  There are {self.num_marg_modes} remaining modes for marginalization.')
        return templates, update

    # -- properties with the reference's names ------------------------------------
    @property
    def blind(self):
        return self._blind

    @property
    def blinding_strat(self):
        return self._blinding_strat

    @property
    def data_vec(self):
        return self._data_vec

    @property
    def masked_data_vec(self):
        if self._masked_data_vec is None:
            self._masked_data_vec = self.data_vec[self.data_mask]
        return self._masked_data_vec

    @property
    def data_size(self):
        return self.masked_data_vec.size

    @property
    def cov_mat(self):
        if self._cov_mat is None:
            raise AttributeError('No covariance matrix found.')
        return self._cov_mat

    @property
    def has_cov_mat(self):
        return self._cov_mat is not None

    @property
    def has_distortion(self):
        # The reference always ends up with a matrix (identity when absent)
        return True

    @property
    def distortion_mat(self):
        """Dense (Nd, Nm) distortion matrix; identity materialised on demand."""
        if self._distortion_mat is None:
            return np.eye(self.full_data_size)
        return self._distortion_mat

    @property
    def inv_masked_cov(self):
        if self._inv_masked_cov is None:
            if self.cov_is_identity:
                self._inv_masked_cov = np.eye(int(self.data_mask.sum()))
            else:
                self._inv_masked_cov = compute_masked_invcov(self.cov_mat, self.data_mask)
        return self._inv_masked_cov

    @property
    def log_cov_det(self):
        if self._log_cov_det is None:
            self._log_cov_det = compute_log_cov_det(self.cov_mat, self.data_mask)
        return self._log_cov_det

    # -- readers ---------------------------------------------------------------------
    def _read_data(self, data_path, cuts_config, dmat_path=None, cov_path=None,
                   cov_rescale=None):
        """Correlation file -> data vector, grids, optional DM / CO columns
        (vega/data.py:285-420)."""
        hdul = fitsio.open(find_file(data_path))
        header = hdul[1].header
        columns = hdul[1].columns.names

        self._blinding_strat = header.get('BLINDING', None)
        if self._blinding_strat in ('none', 'None'):
            self._blinding_strat = None
        if self._blinding_strat in BLINDING_STRATEGIES:
            self._blind = True
            if self._blinding_strat == 'desi_dr3':
                assert 'DA_BLIND' in columns, 'Blinding failed, do not run!!!'
            if 'DA_BLIND' in columns:
                self._data_vec = hdul[1].data['DA_BLIND']
            elif 'DA' in columns:
                self._data_vec = hdul[1].data['DA']
            else:
                raise ValueError('No DA or DA_BLIND column found in data file.')
        elif self._blinding_strat is None or \
                self._blinding_strat in ['desi_m2', 'desi_y1', 'desi_y3']:
            self._blind = False
            self._data_vec = hdul[1].data['DA']
        else:
            self._blind = True
            raise ValueError(f'Unknown blinding strategy {self._blinding_strat}.')
        self._data_vec = np.ascontiguousarray(self._data_vec, dtype=float)

        if dmat_path is None:
            for name in ('DM_BLIND', 'DM'):
                if name in columns:
                    self._distortion_mat = np.ascontiguousarray(
                        hdul[1].data[name], dtype=float)
                    break

        if cov_path is not None:
            with fitsio.open(find_file(cov_path)) as cov_hdul:
                self._cov_mat = np.array(cov_hdul[1].data['CO'], dtype=float)
        elif 'CO' in columns:
            self._cov_mat = np.array(hdul[1].data['CO'], dtype=float)
        if cov_rescale is not None and self._cov_mat is not None:
            self._cov_mat *= cov_rescale

        if 'OMEGAM' in header:
            self.cosmo_params = {
                'Omega_m': header['OMEGAM'], 'Omega_k': header.get('OMEGAK', 0.),
                'Omega_r': header.get('OMEGAR', 0.), 'wl': header.get('WL', -1.)}
        if 'NB' in columns:
            self.nb = hdul[1].data['NB']

        grid_args = (header['RPMIN'], header['RPMAX'], header['RTMAX'], header['NP'], header['NT'])
        self.data_coordinates = Coordinates(
            *grid_args, rp_grid=hdul[1].data['RP'], rt_grid=hdul[1].data['RT'],
            z_grid=hdul[1].data['Z'])
        if dmat_path is None:
            if len(hdul) > 2:
                self.model_coordinates = Coordinates(
                    *grid_args, rp_grid=hdul[2].data['DMRP'], rt_grid=hdul[2].data['DMRT'],
                    z_grid=hdul[2].data['DMZ'])
            self.coeff_binning_model = 1

        self.data_mask = self.data_coordinates.get_mask_scale_cuts(cuts_config)
        if dmat_path is not None:
            self._read_dmat(dmat_path)
        if self.model_coordinates is None:
            self.model_coordinates = self.data_coordinates
        if self.dist_model_coordinates is None:
            self.dist_model_coordinates = self.model_coordinates
        self.model_mask = self.dist_model_coordinates.get_mask_scale_cuts(cuts_config)
        self.full_data_size = len(self.data_vec)

        self.r_min_cut = cuts_config.getfloat('r-min', 10.)
        self.r_max_cut = cuts_config.getfloat('r-max', 180.)
        self.mu_min_cut = cuts_config.getfloat('mu-min', -1.)
        self.mu_max_cut = cuts_config.getfloat('mu-max', +1.)

    def _read_dmat(self, dmat_path):
        """Separate distortion-matrix file carrying its own (distorted and undistorted)
        model grids (vega/data.py:441-473)."""
        hdul = fitsio.open(find_file(dmat_path))
        header = hdul[1].header
        columns = hdul[1].columns.names
        if 'DM' in columns:
            dmat = hdul[1].data['DM']
        elif 'DM_BLIND' in columns:
            dmat = hdul[1].data['DM_BLIND']
        else:
            raise ValueError('No DM or DM_BLIND column found in distortion matrix file.')
        self._distortion_mat = np.ascontiguousarray(dmat, dtype=float)

        self.coeff_binning_model = header['COEFMOD']
        coef = self.coeff_binning_model
        self.model_coordinates = Coordinates(
            header['RPMIN'], header['RPMAX'], header['RTMAX'],
            header['NP'] * coef, header['NT'] * coef,
            rp_grid=hdul[2].data['RP'], rt_grid=hdul[2].data['RT'], z_grid=hdul[2].data['Z'])
        self.dist_model_coordinates = Coordinates(
            header['RPMIN'], header['RPMAX'], header['RTMAX'], header['NP'], header['NT'])

    # -- metals ----------------------------------------------------------------------
    def _use_correlation(self, name1, name2):
        """CIV(eff) only auto-correlates; SiII autos optional (vega/data.py:632-653)."""
        if name1 == 'CIV(eff)' or name2 == 'CIV(eff)':
            return name1 == name2
        if 'SiII' in name1 and 'SiII' in name2 and not self.use_metal_autos:
            return False
        return True

    def _init_metals(self, metal_config):
        """Tracer catalogue, metal-correlation list, per-correlation coordinates and
        matrices from the metal file (vega/data.py:475-513,556-687)."""
        assert ('in tracer1' in metal_config) or ('in tracer2' in metal_config), (
            "The metals config must specify 'in tracer1' and/or 'in tracer2'")
        in1 = metal_config.get('in tracer1').split() if 'in tracer1' in metal_config else None
        in2 = metal_config.get('in tracer2').split() if 'in tracer2' in metal_config else None

        catalog = {self.tracer1['name']: self.tracer1, self.tracer2['name']: self.tracer2}
        for metal in (in1 or []) + (in2 or []):
            catalog[metal] = {'name': metal, 'type': 'continuous'}

        hdul = fitsio.open(find_file(metal_config.get('filename')))
        prefix = 'DM_'
        if hdul[1].header.get('BLINDING', 'none') != 'none':
            prefix = 'DM_BLIND_'
        file_columns = hdul[2].columns.names
        pairs = self._metal_pairs(in1, in2)
        for tracers in pairs:
            name = tracers[0] + '_' + tracers[1]
            if 'RP_' + name not in file_columns:
                name = tracers[1] + '_' + tracers[0]
            self._read_metal_correlation(hdul, tracers, name, prefix)
        return catalog, pairs

    def _metal_tracers_and_pairs(self, metal_config):
        """Tracer catalogue and metal-correlation list without a metal file
        (vega/data.py:475-554, the new_metals branch)."""
        assert ('in tracer1' in metal_config) or ('in tracer2' in metal_config), (
            "The metals config must specify 'in tracer1' and/or 'in tracer2'")
        in1 = metal_config.get('in tracer1').split() if 'in tracer1' in metal_config else None
        in2 = metal_config.get('in tracer2').split() if 'in tracer2' in metal_config else None
        catalog = {self.tracer1['name']: self.tracer1, self.tracer2['name']: self.tracer2}
        for metal in (in1 or []) + (in2 or []):
            catalog[metal] = {'name': metal, 'type': 'continuous'}
        return catalog, self._metal_pairs(in1, in2)

    def _metal_pairs(self, in1, in2):
        t1, t2 = self.tracer1['name'], self.tracer2['name']
        pairs = []
        if in2 is not None:
            pairs += [(t1, metal) for metal in in2 if self._use_correlation(t1, metal)]
        if in1 is not None:
            pairs += [(metal, t2) for metal in in1 if self._use_correlation(metal, t2)]
        if in1 is not None and in2 is not None:
            for i, metal1 in enumerate(in1):
                start = i if self.tracer1 == self.tracer2 else 0
                pairs += [(metal1, metal2) for metal2 in in2[start:]
                          if self._use_correlation(metal1, metal2)]
        return pairs

    def _read_metal_correlation(self, hdul, tracers, name, prefix):
        header = hdul[1].header
        self.metal_coordinates[tracers] = Coordinates(
            header['RPMIN'], header['RPMAX'], header['RTMAX'], header['NP'], header['NT'],
            rp_grid=hdul[2].data['RP_' + name], rt_grid=hdul[2].data['RT_' + name],
            z_grid=hdul[2].data['Z_' + name])
        dm_name = prefix + name
        if dm_name in hdul[2].columns.names:
            self.metal_mats[tracers] = np.ascontiguousarray(hdul[2].data[dm_name], dtype=float)
        elif len(hdul) > 3 and dm_name in hdul[3].columns.names:
            self.metal_mats[tracers] = np.ascontiguousarray(hdul[3].data[dm_name], dtype=float)
        elif self.corr_item.test_flag:
            self.metal_mats[tracers] = None   # identity (vega/data.py:683-684)
        else:
            raise ValueError("Cannot find correct metal matrices."
                             " Check that blinding is consistent between cf and metal files.")

    def get_metal(self, corr_hash):
        """(coordinates, matrix-or-None) of a metal correlation, either order."""
        if corr_hash in self.metal_coordinates:
            return self.metal_coordinates[corr_hash], self.metal_mats[corr_hash]
        return self.metal_coordinates[corr_hash[::-1]], self.metal_mats[corr_hash[::-1]]

    # -- Monte Carlo mocks -------------------------------------------------------------
    def _set_mc_scale(self, scale):
        """Scaled inverse covariance / log-determinant of the Monte Carlo mode; returns True when
        the scale changed (the Cholesky factor is then stale)."""
        changed = not np.isclose(scale, self._scale)
        if changed:
            self._scale = scale
            self.scaled_inv_masked_cov = self.inv_masked_cov / scale
            self.scaled_log_cov_det = self.log_cov_det + np.log(scale)
        if self.scaled_inv_masked_cov is None:
            self.scaled_inv_masked_cov = self.inv_masked_cov
        if self.scaled_log_cov_det is None:
            self.scaled_log_cov_det = self.log_cov_det
        return changed

    def _fiducial_on_data_grid(self, fiducial_model):
        """A fiducial given on the distorted model grid is cut to the data bins."""
        if fiducial_model.size == self.full_data_size:
            return fiducial_model
        if fiducial_model.size != self.dist_model_coordinates.rp_grid.size:
            raise ValueError("Could not match fiducial model to data or model size.")
        return fiducial_model[self.dist_model_coordinates.get_mask_to_other(self.data_coordinates)]

    def create_monte_carlo(self, fiducial_model, scale=None, seed=None, forecast=False):
        """One Gaussian mock around a fiducial model, mock = fiducial + L z with L the Cholesky
        factor of scale * covariance (of the masked block when `cholesky_masked_cov`) and z from
        numpy's global stream -- same draw order as vega/data.py:689-760, so a seed reproduces the
        reference's mock.  `forecast`: the fiducial itself.  Sets mc_mock / masked_mc_mock and
        bumps mc_version (the device re-uploads the data when it changes)."""
        stale = self._set_mc_scale(1 if scale is None else scale)
        if not forecast and (stale or self._cholesky is None):
            cov = self.cov_mat
            if self.cholesky_masked_cov:
                cov = cov[:, self.data_mask][self.data_mask, :]
            self._cholesky = np.linalg.cholesky(self._scale * cov)
        if seed is not None:
            np.random.seed(seed)
        fiducial = self._fiducial_on_data_grid(fiducial_model)
        if forecast:
            self.mc_mock = fiducial
        elif self.cholesky_masked_cov:
            noise = self._cholesky.dot(np.random.randn(int(self.data_mask.sum())))
            self.mc_mock = np.full(self.full_data_size, np.nan)
            self.mc_mock[self.data_mask] = fiducial[self.data_mask] + noise
        else:
            self.mc_mock = fiducial + self._cholesky.dot(np.random.randn(self.full_data_size))
        self.masked_mc_mock = self.mc_mock[self.data_mask]
        self.mc_version = getattr(self, 'mc_version', 0) + 1
        return self.mc_mock
