"""VegaInterface: the drop-in boundary of the B200 path.

Keeps the reference's public surface (vega/vega_interface.py:22-446):
``compute_model(params)``, ``chi2(params)``, ``log_lik(params)`` take a PARTIAL
parameter dict merged over ``self.params`` and return dict-of-arrays / float, so
iminuit, PolyChord adapters and the run_vega_mc scripts can drive it unchanged.
What is new is underneath: every call is a batch of one through the batched
engine (``chi2_batch(theta[B, P])``, ``log_lik_batch``, ``compute_model_batch``),
which runs hand-written sm_100a kernels through the C ABI of include/vega_b200.h.
There is no CPU fallback: without the CUDA library these methods raise.

Init-time work (ini parsing, FITS reading, masks, C^-1) stays on the host and
mirrors vega/vega_interface.py:36-206,665-954.
"""
import configparser
import copy
import os
from pathlib import Path

import numpy as np

from . import _lib, fitsio, utils
from .correlation_item import CorrelationItem
from .data import Data
from .parameters import get_default_values


def _read_ini(path):
    config = configparser.ConfigParser()
    config.optionxform = lambda option: option
    config.read(path)
    return config


class VegaInterface:
    _blind = None
    _use_global_cov = False
    global_cov = None

    def __init__(self, main_path, device=None, batch_chunk=None, metal_builder=None):
        main_path = utils.find_file(main_path)
        utils.add_search_dir(Path(main_path).resolve().parent)
        self.main_config = _read_ini(main_path)

        self.fiducial = self._read_fiducial(self.main_config['fiducial'])
        self.fiducial['z_eff'] = self.main_config['data sets'].getfloat('zeff')
        output_cfg = self.main_config['output'] if 'output' in self.main_config else {}
        self.fiducial['save-components'] = False
        if 'output' in self.main_config:
            self.fiducial['save-components'] = (
                output_cfg.getboolean('write_cf', False) or output_cfg.getboolean('write_pk', False))
        ini_files = self.main_config['data sets'].get('ini files').split()
        global_cov_file = self.main_config['data sets'].get('global-cov-file', None)

        control = self.main_config['control'] if 'control' in self.main_config else None
        self.model_pk = control.getboolean('model_pk', False) if control else False
        self.low_mem_mode = False
        self.marginalize_in_fit = control.getboolean('marginalize-in-fit', False) \
            if control else False

        self.corr_items = {}
        for path in ini_files:
            config = _read_ini(utils.find_file(os.path.expandvars(path)))
            name = config['data'].get('name')
            self.corr_items[name] = CorrelationItem(config, self.model_pk)

        self.params = self._read_parameters(self.corr_items, self.main_config['parameters'])
        self.sample_params = self._read_sample(self.main_config['sample'])

        use_template_growth_rate = control.getboolean('use_template_growth_rate', True) \
            if control else True
        if use_template_growth_rate and 'growth_rate' in self.fiducial:
            assert 'growth_rate' not in self.sample_params['limits'], (
                "use_template_growth_rate is True, but growth_rate is in the sample params.")
            self.params['growth_rate'] = self.fiducial['growth_rate']
        elif 'growth_rate' not in self.fiducial:
            if 'growth_rate' in self.params:
                self.fiducial['growth_rate'] = self.params['growth_rate']
        for key in ('par_sigma_smooth', 'per_sigma_smooth'):
            if key in self.params:
                self.fiducial[key] = self.params[key]

        self._has_data = all(item.has_data for item in self.corr_items.values())
        self.data = {name: (Data(item, self.marginalize_in_fit, metal_builder) if self._has_data else None)
                     for name, item in self.corr_items.items()}

        self._blind = False
        self._rnsps = None
        if self._has_data:
            self._init_blinding()

        self.mc_config = None
        if 'monte carlo' in self.main_config:
            self.mc_config = {'params': {p: float(v) for p, v in
                                         self.main_config['mc parameters'].items()},
                              'sample': self._read_sample(self.main_config['monte carlo'])}

        self.priors = {}
        if 'priors' in self.main_config:
            self.priors = self._init_priors(self.main_config['priors'])
            for param in self.priors:
                not_sampled = param not in self.sample_params['limits']
                if self.mc_config is not None:
                    not_sampled &= param not in self.mc_config['sample']['limits']
                if not_sampled:
                    raise ValueError(
                        f'Prior specified for a parameter that is not sampled: {param}')

        cov_scale = control.getfloat('cov_scale', None) if control else None
        if global_cov_file is not None:
            self.read_global_cov(global_cov_file, cov_scale)

        self.corr_num_marg_modes = {name: (self.data[name].num_marg_modes if self._has_data else 0)
                                    for name in self.corr_items}
        self._random_marg_coeff = None
        self.run_sampler = control.getboolean('run_sampler', False) if control else False
        self.sampler = control.get('sampler', None) if control else None
        self.monte_carlo = False
        self.minimizer = None
        # the reference's per-correlation Model objects (vega/vega_interface.py:136-139)
        from .model import Model
        self.models = {name: Model(self, name) for name in self.corr_items}
        self._analysis = None
        self._output = None
        self.plots = None

        self._device = device
        self._batch_chunk = batch_chunk
        self._engine = None
        self._engines = {}
        self._templates = None          # (pk_full, pk_smooth) when Model.compute was given other templates
        self._templates_pinned = False
        self._fast_metals = False

    # -- init-time readers (vega/vega_interface.py:665-851) ---------------------------
    @staticmethod
    def _read_fiducial(fiducial_config):
        path = utils.find_file(os.path.expandvars(fiducial_config.get('filename')))
        hdul = fitsio.open(path)
        header = hdul[1].header
        fiducial = {
            'z_fiducial': header['ZREF'], 'Omega_m': header['OM'], 'Omega_de': header['OL'],
            'k': np.ascontiguousarray(hdul[1].data['K'], dtype=float),
            'pk_full': np.ascontiguousarray(hdul[1].data['PK'], dtype=float),
            'pk_smooth': np.ascontiguousarray(hdul[1].data['PKSB'], dtype=float),
        }
        if 'F_ZREF' in header:
            fiducial['growth_rate'] = header['F_ZREF']
        return fiducial

    @staticmethod
    def _read_parameters(corr_items, parameters_config):
        params = {}
        for item in corr_items.values():
            if 'parameters' in item.config:
                for param, value in item.config.items('parameters'):
                    params[param] = float(value)
        for param, value in parameters_config.items():
            params[param] = float(value)
        return params

    def _read_sample(self, sample_config):
        """``param = min max [start [sigma]]`` or ``param = True`` with defaults from
        parameters/default_values.txt (vega/vega_interface.py:737-816)."""
        sample = {'limits': {}, 'values': {}, 'errors': {}, 'fix': {}}
        defaults = get_default_values()

        def need_default(param):
            if param not in defaults:
                raise ValueError(f'Default values not found for: {param}. Please add them to '
                                 'default_values.txt, or provide the full sampling specification.')

        for param, values in sample_config.items():
            if param not in self.params:
                print(f'Warning: You tried sampling the parameter: {param}. As this parameter '
                      'was not specified under [parameters], it will be skipped.')
                continue
            words = values.split()
            if len(words) > 1:
                lower = None if words[0] == 'None' else float(words[0])
                upper = None if words[1] == 'None' else float(words[1])
                sample['limits'][param] = (lower, upper)
            else:
                if words[0] not in ['True', 'true', 't', 'y', 'yes']:
                    continue
                need_default(param)
                sample['limits'][param] = defaults[param]['limits']
            if len(words) > 2:
                sample['values'][param] = float(words[2])
            else:
                need_default(param)
                sample['values'][param] = self.params[param]
            if len(words) > 3:
                assert len(words) == 4
                sample['errors'][param] = float(words[3])
            else:
                need_default(param)
                sample['errors'][param] = defaults[param]['error']
            sample['fix'][param] = False
        return sample

    @staticmethod
    def _init_priors(prior_config):
        priors = {}
        for param, prior in prior_config.items():
            words = prior.split()
            if len(words) != 3:
                raise ValueError('Prior configuration must have the format:'
                                 ' "<param> = gaussian <mean> <sigma>"')
            if words[0] not in ['gaussian', 'Gaussian']:
                raise ValueError('Only gaussian priors are supported.')
            priors[param] = np.array(words[1:]).astype(float)
        return priors

    def read_global_cov(self, global_cov_file, scale=None):
        """Joint covariance over all correlations (vega/vega_interface.py:888-954)."""
        with fitsio.open(utils.find_file(global_cov_file)) as hdul:
            self.global_cov = np.array(hdul[1].data['COV'], dtype=float)
        if scale is not None:
            self.global_cov *= scale
        self._use_global_cov = True
        self.full_data_mask = np.concatenate(
            [self.data[name].data_mask for name in self.corr_items])
        self.full_model_mask = np.concatenate(
            [self.data[name].model_mask for name in self.corr_items])
        self.masked_global_invcov = utils.compute_masked_invcov(
            self.global_cov, self.full_data_mask)
        self.masked_global_log_cov_det = utils.compute_log_cov_det(
            self.global_cov, self.full_data_mask)

    @staticmethod
    def _gaussian_chi2_prior(value, mean, sigma):
        return (value - mean)**2 / sigma**2

    @staticmethod
    def _gaussian_lik_prior(sigma):
        return -0.5 * np.log(2 * np.pi) - np.log(sigma)

    def _init_blinding(self):
        """Blinding at the parameter level (vega/vega_interface.py:853-886)."""
        blinding_strat = None
        for data_obj in self.data.values():
            if data_obj.blind:
                self._blind = True
                if blinding_strat is None:
                    blinding_strat = data_obj.blinding_strat
                elif blinding_strat != data_obj.blinding_strat:
                    raise ValueError('Different blinding strategies found in the data sets.')
        if not self._blind:
            return
        blind_pars = []
        for par in self.sample_params['limits'].keys():
            if par in utils.BLIND_FIXED_PARS:
                raise ValueError(f'Running on blind data, parameter {par} must be fixed.')
            if par not in utils.VEGA_BLINDED_PARS:
                continue
            tracers = utils.VEGA_BLINDED_PARS[par]
            if any(corr.check_if_blind_corr(tracers) for corr in self.corr_items.values()):
                blind_pars += [par]
        if len(blind_pars) > 0:
            self._rnsps = utils.get_blinding(blind_pars, blinding_strat)
        if ('bias_QSO' in self.sample_params['limits']) and ('beta_QSO' in self.sample_params['limits']):
            raise ValueError('Running on blind data and sampling bias_QSO and beta_QSO.')

    def _get_lcl_prms(self, params=None):
        """Local copy of the parameters with the blinding applied (vega/vega_interface.py:389-421).
        The device applies the same transform to the theta columns (vg_plan_set_theta_transform);
        this host version serves the plan's constants and callers that read the dictionary."""
        local_params = copy.deepcopy(self.params)
        if params is not None:
            local_params |= params
        assert self._blind is not None, (
            'Blinding flag is not set. Call _init_blinding() before computing the model.')
        if self._rnsps is not None:
            assert self._blind, ('Blinding offsets (_rnsps) are set but blinding flag is False. '
                                 'This is an inconsistent state.')
            local_params = utils.apply_blinding(local_params, self._rnsps)
            for par in local_params:
                if par in utils.BLIND_FIXED_PARS:
                    local_params[par] = 1.
        return local_params

    # -- device engine -----------------------------------------------------------------------
    def _engine_for(self, names=(), values=None):
        """The device plan whose batch columns are the [sample] set plus every other name in
        `names` (parameters outside the columns are constants folded into the plan's static
        grids).  The plan of the [sample] set alone is kept for good; calls that bring extra
        names -- a scalar chi2({'bao_amp': .9}), compute_sensitivity -- get a second plan with
        exactly those extra columns, so they never widen (or slow down) the first one.  A name
        absent from [parameters] is merged the way _get_lcl_prms merges it
        (vega/vega_interface.py:402-404): it joins the parameter set with the value passed."""
        from .plan import PlanBuilder
        if self._templates is not None and not self._templates_pinned:
            self.set_templates(self.fiducial['pk_full'], self.fiducial['pk_smooth'])
        base_names = list(self.sample_params['limits'].keys())
        extra = [n for n in dict.fromkeys(names) if n not in base_names]
        key = tuple(sorted(extra))
        new_values = {}
        for n in extra:
            if n not in self.params:
                if values is None or n not in values:
                    raise KeyError(f'Unknown parameter {n}: it is not listed under [parameters].')
                new_values[n] = float(values[n])
        entry = self._engines.get(key)
        if entry is not None and (entry['params'] != self.params or entry['new'] != new_values):
            entry = None
        if entry is None:
            for old_key in [k for k in self._engines if k != () and k != key]:
                del self._engines[old_key]          # one plan with extra columns at a time
            builder = PlanBuilder(self, sampled=base_names + list(key), extra_params=new_values)
            entry = {'engine': builder.build(chunk=self._batch_chunk), 'params': dict(self.params),
                     'new': new_values, 'data_token': None}
            if self._templates is not None:
                entry['engine'].set_templates(*self._templates)
            self._engines[key] = entry
        engine = entry['engine']
        if self._has_data:
            token = self._data_token()
            if entry['data_token'] is None:
                entry['data_token'] = (False,) + self._data_token(False)[1:]   # what build() uploaded
            if not self._same_token(entry['data_token'], token):
                # the reference reads data.masked_mc_mock / scaled_inv_masked_cov on EVERY chi2 call
                # (vega/vega_interface.py:296-312): any re-draw or change of scale is uploaded
                for ci, name in enumerate(self.corr_items):
                    engine.set_data(ci, self.data[name], monte_carlo=bool(self.monte_carlo))
                entry['data_token'] = token
        self._engine = engine
        return engine

    def _data_token(self, monte_carlo=None):
        """Identity of the arrays the chi2 reads (the Monte Carlo flag, and per correlation the
        data vector / inverse covariance objects of that mode)."""
        mc = bool(self.monte_carlo) if monte_carlo is None else monte_carlo
        token = [mc]
        for name in self.corr_items:
            data = self.data[name]
            if mc:
                token += [data.masked_mc_mock, data.scaled_inv_masked_cov, getattr(data, 'mc_version', 0)]
            else:
                token += [data.masked_data_vec, data.inv_masked_cov, 0]
        return tuple(token)

    @staticmethod
    def _same_token(a, b):
        return len(a) == len(b) and all((x is y) or (not isinstance(x, np.ndarray) and x == y)
                                        for x, y in zip(a, b))

    def _theta_transform(self, sampled):
        """Per-column blinding transform of the device (vg_plan_set_theta_transform)."""
        if self._rnsps is None:
            return None
        offset = np.array([utils.blinding_offset(self._rnsps[n]) if n in self._rnsps else 0.
                           for n in sampled], dtype=np.float64)
        force = np.array([n in utils.BLIND_FIXED_PARS for n in sampled], dtype=np.int32)
        return offset, force

    def set_templates(self, pk_full, pk_smooth):
        """Templates other than the fiducial file's for Model.compute(pars, pk_full, pk_smooth)
        (vega/model.py:157-187): the device re-derives what it derived from the linear P(k)
        (vg_plan_set_templates).  A no-op when they are the ones in use."""
        fid = self.fiducial
        pk_full = np.ascontiguousarray(pk_full, dtype=np.float64)
        pk_smooth = np.ascontiguousarray(pk_smooth, dtype=np.float64)
        current = self._templates if self._templates is not None else (fid['pk_full'], fid['pk_smooth'])
        if np.array_equal(pk_full, current[0]) and np.array_equal(pk_smooth, current[1]):
            return
        if pk_full.size != fid['k'].size or pk_smooth.size != pk_full.size:
            raise ValueError('Templates must be sampled on the k grid of the fiducial file.')
        is_fiducial = np.array_equal(pk_full, fid['pk_full']) and np.array_equal(pk_smooth, fid['pk_smooth'])
        self._templates = None if is_fiducial else (pk_full.copy(), pk_smooth.copy())
        for entry in self._engines.values():
            entry['engine'].set_templates(pk_full, pk_smooth)

    def templates(self, pk_full, pk_smooth):
        """Context in which the engine evaluates on the given templates (Model.compute); the
        interface's own compute_model / chi2 / log_lik always use the fiducial file's, as the
        reference does (vega/vega_interface.py:236-240)."""
        import contextlib

        @contextlib.contextmanager
        def scope():
            self.set_templates(pk_full, pk_smooth)
            self._templates_pinned = True
            try:
                yield
            finally:
                self._templates_pinned = False
        return scope()

    def _direct_mode(self, engine, direct_pk):
        """Context in which the engine evaluates Model.compute_direct on `direct_pk`
        (vega/model.py:189-208, the direct_pk argument of compute_model / chi2 / log_lik)."""
        import contextlib

        @contextlib.contextmanager
        def mode():
            if direct_pk is None:
                yield
                return
            for item in self.corr_items.values():
                if item.has_metals and not item.config['model'].getboolean('no-metal-decomp', True):
                    # vega/model.py:119-121 would evaluate the metals on direct_pk as well
                    raise NotImplementedError('direct_pk with no-metal-decomp = False is not built.')
            engine.set_direct_pk(direct_pk)
            engine.set_option(_lib.OPT_DIRECT_PK, 1)
            try:
                yield
            finally:
                engine.set_option(_lib.OPT_DIRECT_PK, 0)
        return mode()

    def _theta_row(self, engine, params=None):
        """The merge of vega/vega_interface.py:402-404 for the plan's columns; the blinding
        transform of :405-419 happens on the device."""
        own = self.params
        if params:
            return np.array([[params[n] if n in params else own[n] for n in engine.sampled]], dtype=np.float64)
        return np.array([[own[n] for n in engine.sampled]], dtype=np.float64)

    @property
    def theta_names(self):
        """Default column names of the batched entry points: the [sample] set."""
        return list(self.sample_params['limits'].keys())

    # -- the reference's scalar API (batch of one) ----------------------------------------------
    def compute_model(self, params=None, run_init=True, direct_pk=None, marg_coeff=None):
        """Model xi of every correlation for one parameter dict
        (vega/vega_interface.py:208-248).  Returns {name: ndarray(n_dist)}; with `model_pk` in
        [control], the P_ell(k) multipoles [n_ell, nk] instead (vega/model.py:106-107,186)."""
        if self.model_pk:
            bao_amp = self._get_lcl_prms(params)['bao_amp']
            return {name: bao_amp * self.models[name].pk_ells(params, 'peak')
                    + self.models[name].pk_ells(params, 'smooth') for name in self.corr_items}
        engine = self._engine_for(tuple(params) if params else (), params)
        with self._direct_mode(engine, direct_pk):
            models, status = engine.model_device(engine.theta_tensor(self._theta_row(engine, params)))
            status = int(status.cpu()[0])
        if status & _lib.STATUS_ARINYO:
            raise utils.VegaArinyoError
        if status & _lib.STATUS_BOUNDS:
            raise utils.VegaBoundsError
        model_cf = {name: xi[0].cpu().numpy() for name, xi in models.items()}
        if marg_coeff is not None:          # vega/vega_interface.py:243-246
            for name in self.data:
                if self.data[name].marg_templates is not None:
                    model_cf[name] += self.data[name].marg_templates.dot(marg_coeff[name])
        return model_cf

    def compute_marg_coeff(self, model_cf):
        """Best-fit amplitudes of the marginalisation templates for a model
        (vega/vega_interface.py:546-579)."""
        out = {}
        for name in self.corr_items:
            corr_data = self.data[name]
            vec = corr_data.masked_mc_mock if self.monte_carlo else corr_data.masked_data_vec
            diff = vec - model_cf[name][corr_data.model_mask]
            if corr_data.marg_diff2coeff_matrix is not None:
                out[name] = corr_data.marg_diff2coeff_matrix.dot(diff)
        return out

    def chi2(self, params=None, direct_pk=None, return_marg_coeff=False):
        """chi2 incl. priors for one parameter dict (vega/vega_interface.py:250-325);
        model-domain failures give 1e100."""
        assert self._has_data
        engine = self._engine_for(tuple(params) if params else (), params)
        with self._direct_mode(engine, direct_pk):
            chi2, status = engine.chi2_host(self._theta_row(engine, params))
        if status[0] & (_lib.STATUS_ARINYO | _lib.STATUS_BOUNDS):
            if return_marg_coeff:           # vega/vega_interface.py:274-277
                return 1e100, self._random_marg_coeff
            return 1e100
        chi2 = float(chi2[0])
        assert isinstance(chi2, float)
        if return_marg_coeff:
            # the chi2 above already holds the on-the-fly template fit (folded into C^-1 when
            # marginalize-in-fit); the amplitudes themselves come from the model residual
            marg_coeff = self.compute_marg_coeff(
                self.compute_model(params, run_init=False, direct_pk=direct_pk))
            if self._random_marg_coeff is None:
                self._random_marg_coeff = marg_coeff
            return chi2, marg_coeff
        return chi2

    def log_lik(self, params=None, direct_pk=None, return_marg_coeff=False):
        """Log-likelihood for one parameter dict (vega/vega_interface.py:327-387)."""
        assert self._has_data
        if return_marg_coeff:               # vega/vega_interface.py:345-349,384-386
            chi2, marg_coeff = self.chi2(params, direct_pk, True)
            log_lik = float(self._log_norm() - 0.5 * chi2)
            if marg_coeff is None:
                return log_lik, None
            corr_names = [corr for corr in sorted(self.corr_items.keys()) if corr in marg_coeff]
            flat = np.hstack([marg_coeff[corr] for corr in corr_names]) if corr_names else np.array([])
            return log_lik, flat
        chi2 = self.chi2(params, direct_pk)
        return float(self._log_norm() - 0.5 * chi2)

    def _log_norm(self):
        """Parameter-independent part of log_lik (vega/vega_interface.py:351-369)."""
        log_norm = 0.
        for name in self.corr_items:
            log_norm -= 0.5 * self.data[name].data_size * np.log(2 * np.pi)
            if not self._use_global_cov:
                log_norm -= 0.5 * (self.data[name].scaled_log_cov_det if self.monte_carlo
                                   else self.data[name].log_cov_det)
        if self._use_global_cov:
            log_norm -= 0.5 * self.masked_global_log_cov_det
        for prior in self.priors.values():
            log_norm += self._gaussian_lik_prior(prior[1])
        return log_norm

    @property
    def output(self):
        """vega/vega_interface.py:144-146: the Output object (FITS result files)."""
        if self._output is None:
            from .output import Output
            self._output = Output(self.main_config['output'], self.data, self.corr_items, self._analysis)
        self._output.analysis = self._analysis
        return self._output

    @property
    def analysis(self):
        """vega/vega_interface.py:139-142: the Analysis object (here: the batched chi2 scan)."""
        if self._analysis is None:
            from .drivers import BatchedAnalysis
            self._analysis = BatchedAnalysis.from_vega(self, verbose=False)
        return self._analysis

    def minimize(self):
        """Best fit over the [sample] parameters (vega/vega_interface.py:581-644) with the batched
        stencil minimizer; fills bestfit_model, per-correlation chi2 statistics, chisq, p_value."""
        import scipy.stats
        from .drivers import BatchedMinimizer
        if not self.sample_params['limits']:
            print('No sampled parameters. Skipping minimization.')
            return
        if self.minimizer is None:
            self.minimizer = BatchedMinimizer.from_vega(self)
        self.minimizer.minimize()
        self.bestfit_model = self.compute_model(self.minimizer.values, run_init=False)
        self.total_data_size = 0
        self.bestfit_corr_stats = {}
        num_pars = len(self.sample_params['limits'])
        for name in self.corr_items:
            corr_data = self.data[name]
            data_size = getattr(corr_data, 'effective_data_size', corr_data.data_size)
            self.total_data_size += data_size
            if self.monte_carlo and self._use_global_cov:
                chisq = 0
            elif self.monte_carlo:
                diff = corr_data.masked_mc_mock - self.bestfit_model[name][corr_data.model_mask]
                chisq = diff.T.dot(corr_data.scaled_inv_masked_cov.dot(diff))
            else:
                diff = corr_data.masked_data_vec - self.bestfit_model[name][corr_data.model_mask]
                chisq = diff.T.dot(corr_data.inv_masked_cov.dot(diff))
            # small-scale marginalisation: the template amplitudes at the best fit, added to the
            # model that is written out (vega/vega_interface.py:612-618)
            bestfit_marg_coeff = None
            if not (self.monte_carlo and self._use_global_cov) and corr_data.marg_diff2coeff_matrix is not None:
                bestfit_marg_coeff = corr_data.marg_diff2coeff_matrix.dot(diff)
                self.bestfit_model[name] = self.bestfit_model[name] + corr_data.marg_templates.dot(bestfit_marg_coeff)
            reduced_chisq = chisq / (data_size - num_pars)
            p_value = 1 - scipy.stats.chi2.cdf(chisq, data_size - num_pars)
            print(f'{name} chi^2/(ndata-nparam): {chisq:.1f}/({data_size}-{num_pars}) '
                  f'= {reduced_chisq:.3f}, PTE={p_value:.2f}')
            self.bestfit_corr_stats[name] = {
                'masked_size': data_size, 'chisq': chisq, 'reduced_chisq': reduced_chisq,
                'p_value': p_value, 'bestfit_marg_coeff': bestfit_marg_coeff}
        self.chisq = self.minimizer.fmin.fval
        self.reduced_chisq = self.chisq / (self.total_data_size - num_pars)
        self.p_value = 1 - scipy.stats.chi2.cdf(self.chisq, self.total_data_size - num_pars)
        print(f'Total chi^2/(ndata-nparam): {self.chisq:.1f}/({self.total_data_size}-{num_pars}) '
              f'= {self.reduced_chisq:.3f}, PTE={self.p_value:.2f}')
        if not self.minimizer.fmin.is_valid:
            print('Invalid fit!!! Check data, covariance, model and priors.')

    @property
    def bestfit(self):
        """The minimizer after minimize() (the reference returns its iminuit object here,
        vega/vega_interface.py:646-663)."""
        return self.minimizer

    # -- Monte Carlo set-up (vega/vega_interface.py:448-545) ---------------------------------------
    def get_fiducial_for_monte_carlo(self, print_func=print):
        """Fiducial model the mocks are drawn around (vega/vega_interface.py:448-504): the
        [mc parameters] over the best fit of an initial fit when [sample] is not empty, from
        measured files (`use_measured_fiducial`), or from the full P(k) (`use_full_pk_for_mc`)."""
        control = self.main_config['control']
        mc_params = dict(self.mc_config['params'])
        mc_start_from_fit = control.get('mc_start_from_fit', None)
        if mc_start_from_fit is not None:
            # FitResults.params (vega/postprocess/fit_results.py:50-53): BESTFIT names -> values
            print_func(f'Reading input fit {mc_start_from_fit}')
            with fitsio.open(utils.find_file(mc_start_from_fit)) as hdul:
                best = next(h for h in hdul[1:] if h.header.get('EXTNAME') == 'BESTFIT')
                fit_params = {str(n).strip(): float(v) for n, v in zip(best.data['names'], best.data['values'])}
            mc_params = {**fit_params, **mc_params}
            print_func(f'Set template parameters to {mc_params}.')
        elif self.sample_params['limits']:
            print_func('Running initial fit')
            self.minimize()
            mc_params = {**self.minimizer.values, **mc_params}
            print_func(f'Set template parameters to {mc_params}.')
        if control.getboolean('use_measured_fiducial', False):
            fiducial_model = {}
            for name in self.corr_items:
                with fitsio.open(utils.find_file(control.get(f'mc_fiducial_{name}'))) as hdul:
                    fiducial_model[name] = np.array(hdul[1].data['DA'], dtype=float)
            return fiducial_model
        if control.getboolean('use_full_pk_for_mc', False):
            return self.compute_model(mc_params, run_init=False, direct_pk=self.fiducial['pk_full'])
        return self.compute_model(mc_params, run_init=False)

    def initialize_monte_carlo(self, scale=None, print_func=print):
        """One mock in every Data object (or one joint mock under a global covariance), the
        minimizer reset to the [monte carlo] sample set, Monte Carlo mode on
        (vega/vega_interface.py:506-545; bin/run_vega_mpi.py:64, vega/scripts/run_vega.py:24)."""
        from .drivers import BatchedMinimizer, MonteCarloFits
        assert self.mc_config is not None, 'No Monte Carlo config provided'
        fiducial_model = self.get_fiducial_for_monte_carlo(print_func)
        sample = self.mc_config['sample']
        names = list(sample['limits'].keys())
        self.minimizer = BatchedMinimizer(lambda theta: self.chi2_batch(theta, names), sample)
        control = self.main_config['control']
        forecast = control.getboolean('forecast', False)
        seed = control.getint('mc_seed', 0)
        if self._use_global_cov and scale is None and 'global_cov_rescale' in control:
            scale = control.getfloat('global_cov_rescale')
        mc = MonteCarloFits(self, fiducial_model=fiducial_model, scale=scale)
        mc.draw(1, seed=seed, forecast=forecast)        # leaves monte_carlo = True
        self._mc_fits = mc
        if self._use_global_cov:
            return mc.global_mocks[0]
        return {name: mc.full_mocks[name][0] for name in self.corr_items}

    def compute_sensitivity(self, nominal=None, frac=0.1, verbose=True):
        """Model sensitivity to each floating parameter (vega/vega_interface.py:956-1071):
        central differences at value +/- frac * error of the peak / smooth `core` components,
        distorted ([0]) and undistorted ([1]), and the Fisher information per (rp, rt) bin of
        every parameter pair.  Stored in `self.sensitivity` with the reference's keys
        (`nominal`, `partials[corr][par]` [2, 2, Nm], `fisher[corr][(p1, p2)]` [2, Nm]).

        One batch instead of the reference's 2P model evaluations: the 2P stencil points go
        through the per-evaluation xi tap (undistorted components) and, paired with
        bao_amp = 0 and 1, through vg_model_batch -- the model is bao_amp * peak + smooth with
        every later stage linear (vega/model.py:186), so the distorted components of the
        reference's save-components dictionaries are M(0) and M(1) - M(0)."""
        if nominal is None:
            if self.minimizer is None:
                raise RuntimeError('No nominal parameter values provided or saved by minimize()')
            values, errors, fixed = self.minimizer.values, self.minimizer.errors, self.minimizer.fixed
            nominal = {n: (values[n], errors[n]) for n in values if not fixed[n]}
        names = list(nominal)
        P = len(names)
        base = np.array([nominal[n][0] for n in names], dtype=np.float64)
        delta = frac * np.array([nominal[n][1] for n in names], dtype=np.float64)
        bao_amp = self.params['bao_amp']
        if 'bao_amp' in names:
            raise ValueError('compute_sensitivity weights the peak by the fixed bao_amp '
                             '(vega/vega_interface.py:994): it cannot be one of the nominal parameters.')
        if verbose:
            for i, n in enumerate(names):
                print(f'Calculating sensitivity for [{i}] {n} at {nominal[n][0]:.4f} ± {nominal[n][1]:.4f}')

        # rows 2p (+delta) and 2p+1 (-delta)
        pts = np.repeat(base[None, :], 2 * P, axis=0)
        pts[2 * np.arange(P), np.arange(P)] += delta
        pts[2 * np.arange(P) + 1, np.arange(P)] -= delta
        cols = names + ['bao_amp']
        engine = self._engine_for(tuple(cols))
        theta = np.concatenate([np.column_stack([pts, np.full(2 * P, a)]) for a in (0.0, 1.0)])
        theta = self._reorder(engine, theta, cols)
        models, status = engine.model_device(engine.theta_tensor(theta))
        status = status.cpu().numpy()
        if np.any(status & _lib.STATUS_ARINYO):
            raise utils.VegaArinyoError
        if np.any(status & _lib.STATUS_BOUNDS):
            raise utils.VegaBoundsError

        self.sensitivity = dict(nominal=copy.deepcopy(nominal), partials={}, fisher={})
        for ci, name in enumerate(self.corr_items):
            n_model = self.corr_items[name].model_coordinates.rp_grid.size
            ld = (n_model + 15) // 16 * 16
            m = models[name].cpu().numpy()
            smooth_d, peak_d = m[:2 * P], m[2 * P:] - m[:2 * P]
            xi, _ = engine.stage(ci, 3, theta[:2 * P])
            xi = xi.reshape(2 * P, -1, ld)[:, :2, :n_model]          # evaluations 0 / 1: peak / smooth
            partials = self.sensitivity['partials'][name] = {}
            for p, pname in enumerate(names):
                out = np.zeros((2, 2, n_model))
                out[0, 0] = bao_amp * (peak_d[2 * p] - peak_d[2 * p + 1])
                out[0, 1] = smooth_d[2 * p] - smooth_d[2 * p + 1]
                out[1, 0] = bao_amp * (xi[2 * p, 0] - xi[2 * p + 1, 0])
                out[1, 1] = xi[2 * p, 1] - xi[2 * p + 1, 1]
                partials[pname] = out / (2 * delta[p])

        if verbose:
            print('Computing Fisher information for each pair of parameters...')
        import torch
        for name in self.corr_items:
            mask = self.data[name].data_mask
            icov = torch.from_numpy(np.ascontiguousarray(self.data[name].inv_masked_cov)).to(engine.device)
            fisher = self.sensitivity['fisher'][name] = {}
            tot = np.stack([self.sensitivity['partials'][name][n].sum(axis=1) for n in names])   # [P, 2, Nm]
            masked = torch.from_numpy(np.ascontiguousarray(tot[:, :, mask])).to(engine.device)  # [P, 2, Nmask]
            weighted = torch.matmul(masked, icov.T).cpu().numpy()         # C^-1 . partial2
            for i1, n1 in enumerate(names):
                for i2 in range(i1, P):
                    info = np.full((2, tot.shape[2]), np.nan)
                    info[:, mask] = tot[i1][:, mask] * weighted[i2]
                    fisher[(n1, names[i2])] = info
        return self.sensitivity

    # -- batched API (new) -----------------------------------------------------------------------
    def chi2_batch(self, theta, names=None, direct_pk=None):
        """chi2 of B parameter vectors.  theta: ndarray [B, P] with columns `names`
        (default: the [sample] set, see theta_names).  Host in, host out."""
        engine = self._engine_for(tuple(names) if names else ())
        theta = self._reorder(engine, theta, names)
        with self._direct_mode(engine, direct_pk):
            chi2, status = engine.chi2_host(theta)
        chi2[(status & (_lib.STATUS_ARINYO | _lib.STATUS_BOUNDS)) != 0] = 1e100
        return chi2

    def log_lik_batch(self, theta, names=None):
        return self._log_norm() - 0.5 * self.chi2_batch(theta, names)

    def compute_model_batch(self, theta, names=None):
        """{name: ndarray [B, n_dist]} for B parameter vectors."""
        engine = self._engine_for(tuple(names) if names else ())
        theta = self._reorder(engine, theta, names)
        models, status = engine.model_device(engine.theta_tensor(theta))
        return {name: xi.cpu().numpy() for name, xi in models.items()}, status.cpu().numpy()

    def _reorder(self, engine, theta, names):
        """theta[B, len(names)] -> theta[B, len(engine.sampled)]; `names` defaults to the [sample]
        set, columns of the plan that are not named keep their [parameters] value."""
        theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
        names = self.theta_names if names is None else list(names)
        if theta.shape[1] != len(names):
            raise ValueError(f'theta has {theta.shape[1]} columns, expected {len(names)}: {names}')
        if names == engine.sampled:
            return theta
        full = np.tile([engine.builder.params[n] if n not in self.params else self.params[n]
                        for n in engine.sampled], (theta.shape[0], 1))
        for j, n in enumerate(names):
            full[:, engine.sampled.index(n)] = theta[:, j]
        return full
