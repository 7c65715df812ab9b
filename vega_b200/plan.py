"""Builds the device plan (include/vega_b200.h) from the init-time state of VegaInterface.

This is host-side plumbing only: it turns the reference's per-correlation configuration
(ini options, tracer names, parameter names) into flat descriptors -- option flags and
parameter-column indices -- plus the static arrays the kernels consume.  No model
arithmetic happens here; the kernels evaluate everything from the parameter vector.

Which reference object each descriptor block stands for:
  P(k,mu) block      PowerSpectrum.__init__/compute     vega/power_spectrum.py:18-196
  xi block           CorrelationFunction.__init__       vega/correlation_func.py:20-115
  scale parameters   ScaleParameters.get_ap_at          vega/scale_parameters.py:38-230
  metal evaluations  Metals.__init__/compute            vega/metals.py:22-142,258-336
  broadband          BroadbandPolynomials               vega/broadband_poly.py:23-198
"""
import ctypes as C
import gzip
import io

import numpy as np

from . import _lib, hankel, utils
from ._lib import EvalDesc, BBTerm

ELLS_DEFAULT = (0, 2, 4, 6)


def _dptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _legendre(ell, x):
    return np.polynomial.legendre.legval(x, [0] * ell + [1])


def _flag(cfg, key):
    return key in cfg and cfg.getboolean(key)


def resolve_scale(scale_cfg, params, is_peak, corr_name, metal_corr):
    """(form, name1, name2) of the (ap, at) rule for one component; form 0 = (1, 1).
    Mirrors vega/scale_parameters.py:38-230."""
    gb = scale_cfg.getboolean
    full_shape = gb('full-shape', False)
    full_shape_alpha = gb('full-shape-alpha', False)
    smooth_scaling = gb('smooth-scaling', False)
    metal_scaling = gb('metal-scaling', False)
    two_alpha_smooth = gb('two-alpha-smooth', False)
    func = scale_cfg.get('cosmo fit func', 'ap_at')
    form = {'ap_at': 1, 'aiso_epsilon': 2, 'phi_alpha': 3}[func]
    bao = {'ap_at': ('ap', 'at'), 'aiso_epsilon': ('aiso', 'epsilon'),
           'phi_alpha': ('phi', 'alpha')}[func]

    def fullshape():
        if func != 'phi_alpha' and not full_shape_alpha:
            raise ValueError('Only the "phi_alpha" parametrisation works with split full-shape.')
        if func == 'ap_at':
            return form, 'ap_full', 'at_full'
        if func == 'aiso_epsilon':
            return form, 'aiso_full', 'epsilon_full'
        phi_name = 'phi_full' if full_shape else 'phi_smooth'
        if full_shape_alpha:
            alpha_name = 'alpha_full'
        elif is_peak:
            alpha_name = 'alpha'
        elif two_alpha_smooth:
            alpha_name = f'alpha_smooth_{corr_name}'
        else:
            alpha_name = 'alpha_smooth'
        return form, phi_name, alpha_name

    if metal_corr and not metal_scaling:
        return 0, None, None
    if full_shape:
        return fullshape()
    if is_peak:
        return (form,) + bao
    if smooth_scaling:
        return fullshape()
    return 0, None, None


class PlanBuilder:
    """Translates a VegaInterface into a vg_plan."""

    def __init__(self, vega, sampled=None, fast_metals=None, extra_params=None):
        self.vega = vega
        self.lib = _lib.load()
        # the plan's constants are the BLINDED local parameters (vega/vega_interface.py:389-421:
        # offsets on the blinded names, BLIND_FIXED_PARS pinned to 1); sampled columns receive the
        # same transform on the device (vg_plan_set_theta_transform)
        self.unblinded = {**vega.params, **(extra_params or {})}
        self.params = dict(vega._get_lcl_prms(extra_params)) if hasattr(vega, '_get_lcl_prms') \
            else dict(self.unblinded)
        self.names = list(self.params.keys())
        self.extra = {}
        self.sampled = list(vega.sample_params['limits'].keys()) if sampled is None else list(sampled)
        for name in self.sampled:
            if name not in self.params:
                raise KeyError(f'Sampled parameter {name} is not in [parameters].')
        self.sampled_set = set(self.sampled)
        self.fast_metals_override = fast_metals
        self.grid_keys = {}
        self.keep = []          # host arrays that must outlive the ctypes calls
        self.corr_info = {}

    # -- parameter columns --------------------------------------------------------------
    def col(self, name, required=False):
        if name in self.extra:
            return len(self.names) + list(self.extra).index(name)
        if name in self.params:
            return self.names.index(name)
        if required:
            raise KeyError(f'Parameter {name} is required by the model configuration.')
        return -1

    def add_extra(self, name, value):
        if name not in self.extra:
            self.extra[name] = float(value)
        return self.col(name)

    def base_vector(self):
        return _f64([self.params[n] for n in self.names] + list(self.extra.values()))

    def value(self, name):
        return self.extra[name] if name in self.extra else self.params[name]

    def is_dyn(self, *names):
        return any(n in self.sampled_set for n in names if n is not None)

    # -- P(k,mu) block --------------------------------------------------------------------
    def fill_pk(self, d, cfg, tracer1, tracer2, dataset_name, is_peak, pk_lin_sel, fast_metals,
                name_map=None, data_binsizes=None):
        """Option flags + parameter columns of one PowerSpectrum instance
        (vega/power_spectrum.py:18-196).  name_map substitutes parameter names (metals:
        single-metal-beta, pinned growth rate: vega/metals.py:278-293)."""
        name_map = name_map or {}

        def nm(name):
            return name_map.get(name, name)

        def c(name, required=False):
            return self.col(nm(name), required)

        def dyn(*names):
            return self.is_dyn(*[nm(n) for n in names])

        n1, n2 = tracer1['name'], tracer2['name']
        pair = f'{n1}x{n2}'
        d.pk_lin_sel = pk_lin_sel
        d.is_peak = int(is_peak)
        d.fast_metals = int(fast_metals)
        d.same_tracer = int(n1 == n2)
        for t, n in enumerate((n1, n2)):
            d.idx_bias[t] = c('bias_' + n)
            d.idx_bias_eta[t] = c('bias_eta_' + n)
            d.idx_beta[t] = c('beta_' + n)
            present = [x >= 0 for x in (d.idx_bias[t], d.idx_bias_eta[t], d.idx_beta[t])]
            if sum(present) < 2:
                raise AssertionError(
                    'For each tracer, you need to specify two of these three: (bias, bias_eta, beta).'
                    f' Offending tracer: {n}')
            if present[0] and present[2]:
                d.idx_bias_eta[t] = -1          # bias and beta win (vega/utils.py:66-68)
        d.idx_growth_rate = c('growth_rate')

        # UV / HeII (power_spectrum.py:55-56,111-114,224-261)
        d.uvb = int(cfg.getboolean('UVB-fluctuations', False))
        d.heii = int(cfg.getboolean('HeII-reionization', False))
        if d.uvb or d.heii:
            d.uv_on[0], d.uv_on[1] = int(n1 == 'LYA'), int(n2 == 'LYA')
            d.idx_bias_prim = c('bias_prim', True)
            if d.uvb:
                d.idx_lambda_uv, d.idx_bias_gamma = c('lambda_uv', True), c('bias_gamma', True)
            if d.heii:
                d.idx_lambda_heii, d.idx_bias_gamma_e = c('lambda_HeII', True), c('bias_gamma_e', True)
            # none of them sampled: the scale-dependent bias term is a static function of k (moment form)
            uv_names = ['bias_prim'] + (['lambda_uv', 'bias_gamma'] if d.uvb else []) \
                + (['lambda_HeII', 'bias_gamma_e'] if d.heii else [])
            d.uv_static = int(not any(n in self.sampled_set for n in uv_names))

        # HCD (power_spectrum.py:53,116-120,263-309)
        hcd = cfg.get('model-hcd', None)
        if hcd is not None and (n1 == 'LYA' or n2 == 'LYA'):
            d.hcd_on[0], d.hcd_on[1] = int(n1 == 'LYA'), int(n2 == 'LYA')
            d.idx_bias_hcd = c(f'bias_hcd_{pair}') if c(f'bias_hcd_{pair}') >= 0 else c('bias_hcd', True)
            d.idx_beta_hcd = c(f'beta_hcd_{pair}') if c(f'beta_hcd_{pair}') >= 0 else c('beta_hcd', True)
            if 'Rogers' in hcd:
                d.hcd_model, l0_name = 1, 'L0_hcd'
                d.idx_L0 = c(l0_name, True)
            elif 'fvoigt' in hcd:
                d.hcd_model, l0_name = 2, 'L0_fvoigt'
                d.idx_L0 = c(l0_name)
                self.need_fvoigt = cfg.get('fvoigt_model')
            elif 'sinc' in hcd:
                d.hcd_model, l0_name = 3, 'L0_sinc'
                d.idx_L0 = c(l0_name)
            else:
                raise ValueError(f"Unknown hcd model {hcd}. Choose from ['Rogers', 'fvoigt', 'sinc']")
            if not dyn(l0_name):
                l0 = self.value(nm(l0_name)) if d.idx_L0 >= 0 else 1.0
                d.hcd_grid = self.grid_id(('hcd', d.hcd_model, float(l0)))

        static_key = []

        def factor(bit, names, key):
            if dyn(*names):
                d.dyn_mask |= bit
            else:
                d.static_mask |= bit
                static_key.append((bit, key, tuple(self.value(nm(n)) for n in names if c(n) >= 0)))

        skip_nl = cfg.getboolean('skip-nl-model-in-peak', False) and is_peak
        if 'small scale nl' in cfg and not skip_nl:
            kind = cfg.get('small scale nl')
            if 'arinyo' in kind:
                names = ['dnl_arinyo_' + s for s in ('q1', 'q2', 'kv', 'av', 'bv', 'kp')]
                for i, n in enumerate(names):
                    d.idx_arinyo[i] = c(n, required=(i != 1))
                two = 'LY' in n1 and 'LY' in n2
                one = 'LY' in n1 or 'LY' in n2
                d.arinyo_mode = 2 if two else (1 if one else 0)
                factor(_lib.F_ARINYO, names, d.arinyo_mode)
            elif 'mcdonald' in kind:
                assert n1 == 'LYA' and n2 == 'LYA'
                factor(_lib.F_MCDONALD, [], None)
            else:
                raise ValueError("Incorrect 'small scale nl' specified")

        if cfg.getboolean('model binning', True):
            bs = data_binsizes
            d.gk_drp = self.params.get(f'par binsize {dataset_name}', bs[0])
            d.gk_drt = self.params.get(f'per binsize {dataset_name}', bs[1])
            factor(_lib.F_GK, [], (d.gk_drp, d.gk_drt))
        if 'mock-bin-size' in cfg:
            d.mock_bin_size = cfg.getfloat('mock-bin-size')
            los = cfg.get('mock-los-smoothing', None)
            names = []
            if los == 'growth':
                d.mock_los_mode, names = 1, ['growth_rate']
                if d.idx_growth_rate < 0:
                    raise KeyError('growth_rate')
            elif los == 'amplitude':
                d.mock_los_mode, names = 2, ['los_smooth_amp']
                d.idx_los_smooth_amp = c('los_smooth_amp', True)
            elif los == 'only-los':
                d.mock_los_mode = 3
            elif los is not None:
                raise ValueError(f'Unknown mock LOS smoothing option {los}.')
            factor(_lib.F_MOCKGK, names, (d.mock_bin_size, d.mock_los_mode))

        if is_peak:
            d.idx_sigmaNL[0], d.idx_sigmaNL[1] = c('sigmaNL_par'), c('sigmaNL_per')
            if d.idx_sigmaNL[0] < 0 and d.idx_sigmaNL[1] < 0:
                raise ValueError('No parameters for peak NL found. Add sigmaNL_par and/or sigmaNL_par.')
            names = ['sigmaNL_par', 'sigmaNL_per']
            if d.idx_sigmaNL[0] < 0 or d.idx_sigmaNL[1] < 0:
                names.append('growth_rate')
                if d.idx_growth_rate < 0:
                    raise KeyError('growth_rate is needed to derive the missing sigmaNL.')
            factor(_lib.F_PEAKNL, names, None)

        if 'fullshape smoothing' in cfg and not skip_nl:
            kind = cfg.get('fullshape smoothing')
            if 'gauss' in kind:
                names = self._fs_gauss_names(c, n1, n2)
                for i, n in enumerate(names):
                    d.idx_fs_gauss[i] = c(n, True)
                factor(_lib.F_FSGAUSS, names, None)
            elif 'exp' in kind:
                names = ['par_sigma_smooth', 'per_sigma_smooth', 'par_exp_smooth', 'per_exp_smooth']
                for i, n in enumerate(names):
                    d.idx_fs_exp[i] = c(n, True)
                factor(_lib.F_FSEXP, names, None)
            else:
                raise ValueError('"fullshape smoothing" must be of type "gauss" or "exp".')

        if 'velocity dispersion' in cfg:
            kind = cfg.get('velocity dispersion')
            # 'lorentz_gauss' takes the gauss branch, like the reference (power_spectrum.py:178-186)
            if 'gauss' in kind:
                d.velo_kind, stem = 1, 'sigma_velo_disp_gauss_'
            elif 'lorentz' in kind:
                d.velo_kind, stem = 2, 'sigma_velo_disp_lorentz_'
            else:
                raise ValueError('"velocity dispersion" must be of type "gauss" or "lorentz".')
            names = []
            for t, tr in enumerate((tracer1, tracer2)):
                if tr['type'] == 'discrete':
                    d.idx_velo[t] = c(stem + tr['name'], True)
                    names.append(stem + tr['name'])
            if names:
                factor(_lib.F_VELO, names, (d.velo_kind, tuple(d.idx_velo)))

        scale = cfg.getfloat('pk-damping-scale', None)
        if scale is not None:
            d.kdamp_scale = scale
            d.kdamp_power = cfg.getint('pk-damping-power', 2)
            factor(_lib.F_KDAMP, [], (scale, d.kdamp_power))

        if d.static_mask:
            d.static_grid = self.grid_id(('static', int(d.static_mask), tuple(static_key)))

    @staticmethod
    def _fs_gauss_names(c, n1, n2):
        """The three parameter-name branches of vega/power_spectrum.py:520-553."""
        has_par, has_per = c('par_sigma_smooth') >= 0, c('per_sigma_smooth') >= 0
        if has_par or has_per:
            par = 'par_sigma_smooth' if has_par else 'per_sigma_smooth'
            per = 'per_sigma_smooth' if has_per else 'par_sigma_smooth'
            return [par, per, par, per]
        main = n1 in ('LYA', 'QSO') and n2 in ('LYA', 'QSO')
        if c('par_sigma_smooth_metals') >= 0 and c('per_sigma_smooth_metals') >= 0 and not main:
            return ['par_sigma_smooth_metals', 'per_sigma_smooth_metals'] * 2
        return [f'par_sigma_smooth_{n1}', f'per_sigma_smooth_{n1}',
                f'par_sigma_smooth_{n2}', f'per_sigma_smooth_{n2}']

    def grid_id(self, key):
        if key not in self.grid_keys:
            self.grid_keys[key] = len(self.grid_keys)
        return self.grid_keys[key]

    # -- xi block -------------------------------------------------------------------------
    def fill_xi(self, d, cfg, tracer1, tracer2, is_peak, corr_pair_name, metal_corr, name_map=None):
        """Flags + columns of one CorrelationFunction (vega/correlation_func.py:20-236)."""
        name_map = name_map or {}

        def c(name, required=False):
            return self.col(name_map.get(name, name), required)

        scale_cfg = self.vega.main_config['cosmo-fit type']
        form, a, b = resolve_scale(scale_cfg, self.params, is_peak, corr_pair_name, metal_corr)
        d.scale_form = form
        if form:
            d.idx_scale[0], d.idx_scale[1] = c(a, True), c(b, True)
        d.single_ell = cfg.getint('single_multipole', -1)
        d.rescale_sys = int(cfg.getboolean('rescale-coords-systematics', False))

        types = (tracer1['type'], tracer2['type'])
        drp_name = None
        if types[0] == 'discrete' and types[1] != 'discrete':
            drp_name = 'drp_' + tracer1['name']
        elif types[1] == 'discrete' and types[0] != 'discrete':
            drp_name = 'drp_' + tracer2['name']
        if drp_name is not None:
            d.idx_drp = c(drp_name)

        # bias evolution (correlation_func.py:276-370)
        for t, tr in enumerate((tracer1, tracer2)):
            key = 'z evol {}'.format(tr['name'])
            kind = cfg.get(key, 'standard') if key in cfg else cfg.get('z evol', 'standard')
            if 'croom' in kind:
                assert tr['name'] == 'QSO', 'Croom bias evolution only works for QSO'
                d.zevol_kind[t] = 1
                d.idx_croom[0], d.idx_croom[1] = c('croom_par0', True), c('croom_par1', True)
            else:
                d.zevol_kind[t] = 0
                d.idx_alpha[t] = c('alpha_' + tr['name'], True)

        names = [tracer1['name'], tracer2['name']]
        if _flag(cfg, 'radiation effects'):
            if not ('QSO' in names and 'LYA' in names):
                raise ValueError('You asked for QSO radiation effects, but it'
                                 ' can only be applied to the cross (QSOxLya)')
            d.radiation = 1
            for i, n in enumerate(('qso_rad_strength', 'qso_rad_asymmetry', 'qso_rad_lifetime',
                                   'qso_rad_decrease')):
                d.idx_qso_rad[i] = c(n, True)
        d.rel = int(_flag(cfg, 'relativistic correction'))
        d.asy = int(_flag(cfg, 'standard asymmetry'))
        if d.rel or d.asy:
            if ('continuous' not in types) or (types[0] == types[1]):
                raise ValueError('You asked for relativistic effects or standard asymmetry,'
                                 ' but they only work for the cross')
            form2, a2, b2 = resolve_scale(scale_cfg, self.params, is_peak, None, metal_corr)
            d.scale_form_sys = form2
            if form2:
                d.idx_scale_sys[0], d.idx_scale_sys[1] = c(a2, True), c(b2, True)
            if d.rel:
                d.idx_Arel[0], d.idx_Arel[1] = c('Arel1', True), c('Arel3', True)
            if d.asy:
                for i, n in enumerate(('Aasy0', 'Aasy2', 'Aasy3')):
                    d.idx_Aasy[i] = c(n, True)
        if _flag(cfg, 'UVB-shotnoise'):
            d.uv_shot = 1
            bg = 'bias_gamma' if c('bias_gamma') >= 0 else 'bias_gamma_e'
            d.idx_uvshot[0], d.idx_uvshot[1] = c('lambda_uv', True), c(bg, True)
            d.idx_uvshot[2] = c('uv_shotnoise_amp', True)
            self.need_uvshot = True

        # static xi stage: no parameter read by the gather is sampled, no additive term
        xi_names = [a, b] if form else []
        xi_names += [drp_name] if drp_name is not None else []
        xi_names += ['alpha_' + tr['name'] for tr in (tracer1, tracer2)] + ['croom_par0', 'croom_par1']
        sampled = any(name_map.get(n, n) in self.sampled_set for n in xi_names)
        d.static_xi = int(not sampled and not (d.radiation or d.rel or d.asy or d.uv_shot))
        evol_names = ['alpha_' + tr['name'] for tr in (tracer1, tracer2)] + ['croom_par0', 'croom_par1']
        d.static_evol = int(not any(name_map.get(n, n) in self.sampled_set for n in evol_names))

    def coord_table(self, coords, cfg, tracer1=None, tracer2=None, cosmo=None):
        """[6][n] static per-bin table + (1+z_eff)^2 (vega/correlation_func.py:46-48,238-274,
        372-444)."""
        fid = self.vega.fiducial
        n = coords.r_grid.size
        z = np.broadcast_to(np.asarray(coords.z_grid, dtype=float), (n,)).copy()
        z_eff, z_fid = fid['z_eff'], fid['z_fiducial']
        om, ode = fid.get('Omega_m', None), fid.get('Omega_de', None)
        if ode is None:
            growth = ((1 + z_fid) / (1. + z))**2
        elif cfg.getboolean('old_growth_func', False):
            growth = (utils.growth_function_old(z, om, ode) / utils.growth_function_old(z_fid, om, ode))**2
        else:
            growth = (utils.growth_function(z, om, ode) / utils.growth_function(z_fid, om, ode))**2
        rel1 = rel2 = (1. + z) / (1 + z_eff)
        types = (tracer1['type'], tracer2['type']) if tracer1 is not None else ('', '')
        if cfg.getboolean('new-bias-evolution', False) and types[0] != types[1]:
            if cosmo is None:
                print('Warning: No cosmology found in xcf files, using mean redshift evolution.')
            else:       # separate redshifts of the two tracers: rp ~ (z_F - z_Q) D_H(z)
                for t in (tracer1, tracer2):
                    key = 'z evol {}'.format(t['name'])
                    kind = cfg.get(key, 'standard') if key in cfg else cfg.get('z evol', 'standard')
                    assert 'croom' not in kind, 'Croom model is not supported with new bias evol'
                half = coords.r_grid * coords.mu_grid / (2 * cosmo.get_dist_hubble(z))
                rel_q, rel_f = (1. + z - half) / (1 + z_eff), (1. + z + half) / (1 + z_eff)
                rel1 = rel_q if types[0] == 'discrete' else rel_f
                rel2 = rel_q if types[1] == 'discrete' else rel_f
        tab = np.stack([coords.r_grid, coords.mu_grid, np.log(rel1), (1. + z)**2, growth, np.log(rel2)])
        return _f64(tab), (1 + z_eff)**2

    # -- broadband --------------------------------------------------------------------------
    def broadband_terms(self, item):
        """Terms + design tables of vega/broadband_poly.py:23-198."""
        cfg = item.config
        terms, design = [], {'pre': [], 'post': []}
        if 'broadband' not in cfg:
            return terms, design
        pos_id = {'pre-mul': 0, 'pre-add': 1, 'post-mul': 2, 'post-add': 3}
        for i, spec in enumerate(cfg['broadband'].values()):
            w = spec.split()
            if len(w) not in (5, 6):
                raise ValueError('Broadband setup must have 5 or 6 elements. Example config:'
                                 ' "add pre rp,rt 0:0:1 0:0:1 broadband_sky" (last one is optional).')
            name = f'BB-{item.name}-{i}-{w[5]}' if len(w) > 5 else f'BB-{item.name}-{i} {w[0]} {w[1]} {w[2]}'
            func = 'broadband' if len(w) == 5 else w[5]
            where = w[1]
            coords = item.model_coordinates if where == 'pre' else item.dist_model_coordinates
            t = BBTerm()
            t.position = pos_id[f'{where}-{w[0]}']
            t.idx_sky[0] = t.idx_sky[1] = -1
            if func == 'broadband':
                if w[2] == 'r,mu':
                    x1, x2 = coords.r_grid / 100., coords.mu_grid
                elif w[2] == 'rp,rt':
                    x1 = coords.r_grid / 100. * coords.mu_grid
                    x2 = coords.r_grid / 100. * np.sqrt(1 - coords.mu_grid**2)
                else:
                    raise ValueError(f'Unknown broadband coordinates {w[2]}')
                lo1, hi1, s1 = (int(x) for x in w[3].split(':'))
                lo2, hi2, s2 = (int(x) for x in w[4].split(':'))
                t.kind = 0
                t.design_offset = len(design[where])
                n = 0
                for a in np.arange(lo1, hi1 + 1, s1):
                    for b in np.arange(lo2, hi2 + 1, s2):
                        if n >= 32:
                            raise NotImplementedError('More than 32 coefficients in one broadband term.')
                        t.idx_coef[n] = self.col(f'{name} ({a},{b})', True)
                        design[where].append(x1**a * x2**b)
                        n += 1
                t.n_coef = n
            else:
                t.kind = 1
                t.idx_sky[0] = self.col(name + '-scale-sky', True)
                t.idx_sky[1] = self.col(name + '-sigma-sky', True)
            terms.append(t)
        if len(terms) > _lib.VG_MAX_BB_TERMS:
            raise NotImplementedError('Too many broadband terms.')
        return terms, design

    @staticmethod
    def sky_table(coords):
        window = (coords.rp_grid >= 0.) & (coords.rp_grid < coords.rp_binsize)
        return _f64(np.stack([coords.rt_grid, window.astype(float)]))

    # -- one correlation -----------------------------------------------------------------------
    def describe_corr(self, name):
        """Everything of one correlation that is known before the plan exists."""
        vega = self.vega
        item, data = vega.corr_items[name], vega.data[name]
        cfg = item.config
        mcfg = cfg['model']
        no_decomp = mcfg.getboolean('no-metal-decomp', True)
        if mcfg.getboolean('fht_extrap', False):
            raise NotImplementedError('fht_extrap = True is not built.')
        binsizes = (item.data_coordinates.rp_binsize, item.data_coordinates.rt_binsize)
        info = {'name': name, 'item': item, 'data': data, 'evals': [], 'coords': [], 'static_xi': {}}
        ell_max = mcfg.getint('ell_max', 6)
        info['ells'] = tuple(int(x) for x in np.arange(0, ell_max + 1, 2))
        info['legacy'] = mcfg.getboolean('old_fftlog', False)
        info['lowring'] = mcfg.getboolean('fht_lowring', True)
        pair_name = f"{item.tracer1['name']}x{item.tracer2['name']}"

        cosmo = getattr(item, 'cosmo', None)
        info['coords'].append(self.coord_table(item.model_coordinates, mcfg, item.tracer1, item.tracer2, cosmo))
        for comp, (is_peak, sel, slot) in {'peak': (True, 0, 0), 'smooth': (False, 1, 1)}.items():
            d = EvalDesc()
            self.fill_pk(d, mcfg, item.tracer1, item.tracer2, item.name, is_peak, sel, False,
                         data_binsizes=binsizes)
            self.fill_xi(d, mcfg, item.tracer1, item.tracer2, is_peak, pair_name, False)
            d.coord_set, d.out_slot = 0, slot
            info['evals'].append(d)
        d0 = info['evals'][0]
        if d0.rel or d0.asy:
            k, pk = vega.fiducial['k'], None
            for which, (ell, n_pow, on, sel) in enumerate(
                    [(1, 1., d0.rel, None), (3, 1., d0.rel, None), (0, 2., d0.asy, None), (2, 2., d0.asy, None)]):
                if on:
                    info['static_xi'][which] = (ell, n_pow)

        # metals (vega/metals.py:22-142,258-336)
        info['metal_mats'] = []
        if item.has_metals:
            metcfg = cfg['metals']
            fast_metals = mcfg.getboolean('fast_metals', False) if self.fast_metals_override is None \
                else self.fast_metals_override
            separate_auto = mcfg.getboolean('separate-metal-auto-biases', False)
            fast_bias = mcfg.getboolean('fast_metal_bias', True) or fast_metals or separate_auto
            if fast_metals and not no_decomp:
                # the reference's metal x metal cache is keyed on the pair alone (vega/metals.py:
                # 144-170), so with two components it would hand the peak term back as the smooth one
                raise NotImplementedError('fast_metals needs no-metal-decomp = True.')
            # no-metal-decomp = False (vega/model.py:119-131,178-186): the metal terms go through
            # the peak/smooth split like the core -- one evaluation per pair and component, the
            # peak one weighted by bao_amp, both through the same metal matrix
            components = [(False, 2, -1)] if no_decomp else \
                [(True, 0, self.col('bao_amp')), (False, 1, -1)]
            single_beta = mcfg.getboolean('single-metal-beta', False)
            main = [item.tracer1['name'], item.tracer2['name']]
            # fast_metals (vega/metals.py:144-207,278-336): growth_rate pinned to the fiducial value,
            # metal x metal terms frozen at the first evaluation, metal x main terms cached on
            # (beta1, beta2, every parameter) -- NOT on the pair.
            pinned = {}
            if fast_metals and 'growth_rate' in self.params and 'growth_rate' in vega.fiducial:
                self.add_extra('growth_rate@fiducial', vega.fiducial['growth_rate'])
                pinned['growth_rate'] = 'growth_rate@fiducial'
            # identity of every parameter as the reference's cache key sees it: a sampled column
            # keeps its name, a fixed one its value; single-metal-beta adds entries as the loop
            # over the pairs advances (local_pars is mutated in place, vega/metals.py:285-293)
            key_state = {n: (n if n in self.sampled_set else float(v)) for n, v in self.params.items()}
            owners = {}
            for c, pair in enumerate(item.metal_correlations):
                n1, n2 = pair
                tr1, tr2 = item.tracer_catalog[n1], item.tracer_catalog[n2]
                coords, mat = data.get_metal(pair)
                name_map = dict(pinned)
                if single_beta:
                    for n in pair:
                        if n not in main:
                            name_map[f'beta_{n}'] = 'beta_metals'
                            key_state[f'beta_{n}'] = key_state['beta_metals']
                with_main = n1 in main or n2 in main
                pair_coords = None
                for comp_peak, comp_sel, comp_weight in components:
                    d = EvalDesc()
                    self.fill_pk(d, metcfg, tr1, tr2, item.name, comp_peak, comp_sel, fast_bias,
                                 name_map=name_map, data_binsizes=binsizes)
                    owner = None
                    if fast_metals and with_main:
                        betas = tuple(key_state.get(name_map.get(f'beta_{n}', f'beta_{n}')) for n in pair)
                        key = (betas, tuple(sorted(key_state.items(), key=lambda kv: kv[0])))
                        owner = owners.get(key)
                        if owner is None:
                            owners[key] = (len(info['evals']), len(info['coords']), tr1, tr2, f'{n1}x{n2}',
                                           dict(name_map))
                    if owner is not None:
                        # cache hit: the xi of the first pair with this key (its tracers, its coordinates)
                        # is re-used; only the bias product and the metal matrix are this pair's
                        o_eval, o_coords, o_tr1, o_tr2, o_pair, o_map = owner
                        self.fill_xi(d, metcfg, o_tr1, o_tr2, comp_peak, o_pair, True, name_map=o_map)
                        d.coord_set, d.table_owner = o_coords, o_eval
                    else:
                        if pair_coords is None:
                            info['coords'].append(self.coord_table(coords, metcfg, tr1, tr2, cosmo))
                            pair_coords = len(info['coords']) - 1
                        self.fill_xi(d, metcfg, tr1, tr2, comp_peak, f'{n1}x{n2}', True, name_map=name_map)
                        d.coord_set = pair_coords
                    d.out_slot = 2 + len(info['metal_mats'])
                    d.idx_weight_extra2 = comp_weight
                    d.weight_bias = int(fast_bias)
                    d.frozen = int(fast_metals and not with_main)
                    if fast_bias and not with_main and separate_auto and n1 != n2:
                        if self.col(f'bias_{n1}_{n2}') >= 0:
                            d.idx_weight_extra = self.col(f'bias_{n1}_{n2}')
                        elif self.col(f'bias_{n2}_{n1}') >= 0:
                            d.idx_weight_extra = self.col(f'bias_{n2}_{n1}')
                        else:
                            raise ValueError(f'Separate metal auto biases is on, but no bias_{n1}_{n2}'
                                             f' or bias_{n2}_{n1} parameter found for {pair}.')
                    d.weight_const = 2.0 if (main[0] == main[1] and n1 != n2) else 1.0
                    info['evals'].append(d)
                    info['metal_mats'].append(mat)

        # Model.compute_direct (vega/model.py:189-208): the `full` component from an externally
        # supplied isotropic P(k) -- the smooth component's options with pars['peak'] = False
        d = EvalDesc()
        self.fill_pk(d, mcfg, item.tracer1, item.tracer2, item.name, False, 3, False,
                     data_binsizes=binsizes)
        self.fill_xi(d, mcfg, item.tracer1, item.tracer2, False, pair_name, False)
        d.coord_set, d.out_slot, d.direct = 0, 1, 1
        info['evals'].append(d)

        # DESI instrumental systematics (vega/correlation_func.py:553-595)
        info['inst'] = None
        if mcfg.getboolean('desi-instrumental-systematics', False):
            if item.tracer1['type'] != item.tracer2['type']:
                raise ValueError('DESI instrumental systematics model only applies '
                                 'to auto-correlation functions.')
            coords = item.model_coordinates
            rp = coords.r_grid * coords.mu_grid
            rt = coords.r_grid * np.sqrt(1 - coords.mu_grid**2)
            tab = np.genfromtxt(utils.find_file(
                'instrumental_systematics/desi-instrument-syst-for-forest-auto-correlation.csv'),
                delimiter=',', names=True)
            w = (rp > 0) & (rp < binsizes[0])
            if np.any(rt[w] < tab['RT'].min()) or np.any(rt[w] > tab['RT'].max()):
                raise ValueError('A value in x_new is outside the interpolation range.')
            table = np.zeros(rt.shape)
            table[w] = np.interp(rt[w], tab['RT'], tab['XI'])
            info['inst'] = (_f64(table), self.col('desi_inst_sys_amp'), 0.0003189935987295203)

        info['bb_terms'], info['bb_design'] = self.broadband_terms(item)
        self.share_identical_tables(info['evals'])
        self.pair_twins(info['evals'])
        return info

    # fields of vg_eval_desc that determine P(k,mu) -> knot tables (everything before the xi block)
    _PK_FIELDS = [name for name, _ in EvalDesc._fields_][
        :[name for name, _ in EvalDesc._fields_].index('coord_set')]

    def pk_signature(self, d):
        """Hashable identity of the knot tables of one evaluation: two evaluations with equal
        signatures get bit-identical P(k,mu) for every parameter vector of a batch.  A parameter
        column counts by name when it is sampled and by value when it is fixed."""
        base = self.base_vector()

        def ident(idx):
            if idx < 0:
                return None
            name = (self.names + list(self.extra))[idx]
            return ('col', idx) if name in self.sampled_set else ('val', float(base[idx]))

        skip = set()
        if d.fast_metals:
            # Kaiser term without the biases (vega/power_spectrum.py:219-220): when beta is given
            # directly and the tracer carries no HCD term, its bias never enters P(k,mu)
            for t in range(2):
                if d.idx_beta[t] >= 0 and not (d.hcd_model and d.hcd_on[t]):
                    skip.add(('idx_bias', t))
                    skip.add(('idx_bias_eta', t))
        sig = []
        for name in self._PK_FIELDS:
            v = getattr(d, name)
            if hasattr(v, '__len__'):
                vals = list(v)
                if name.startswith('idx_'):
                    vals = [None if (name, t) in skip else ident(x) for t, x in enumerate(vals)]
                sig.append((name, tuple(vals)))
            else:
                sig.append((name, ident(v) if name.startswith('idx_') else v))
        return tuple(sig)

    _TWIN_FREE = ('pk_lin_sel', 'is_peak', 'static_grid', 'static_mask')

    def pair_twins(self, evals):
        """Peak and smooth of one correlation differ only by the linear P(k) and by what is in
        their static grids (vega/model.py:176-186): when every per-vector factor is the same, the
        smooth evaluation points at the peak one (vg_eval_desc.pk_twin) and the device evaluates
        the shared Kaiser / HCD / velocity-dispersion arithmetic once for both."""
        def dynamic_part(d):
            free = self._TWIN_FREE
            if not (d.dyn_mask & _lib.F_PEAKNL):    # the peak's non-linear broadening is in its static grid
                free = free + ('idx_sigmaNL',)
            return tuple(kv for kv in self.pk_signature(d) if kv[0] not in free)
        if len(evals) < 2:
            return
        peak, smooth = evals[0], evals[1]
        if peak.table_owner >= 0 or smooth.table_owner >= 0 or peak.direct or smooth.direct:
            return
        if dynamic_part(peak) == dynamic_part(smooth):
            smooth.pk_twin = 0

    def share_identical_tables(self, evals):
        """Evaluations of one correlation whose P(k,mu) inputs are identical (e.g. QSO x metal terms
        whose metal betas are the same fixed number) read the knot tables of the first of them
        (vg_eval_desc.table_owner): the multipoles and the Hankel transform are computed once,
        every evaluation keeps its own coordinates, weight and metal matrix."""
        first = {}
        for e, d in enumerate(evals):
            if d.table_owner >= 0 or d.frozen or d.direct:
                continue
            sig = self.pk_signature(d)
            if sig in first:
                d.table_owner = first[sig]
            else:
                first[sig] = e

    # -- push everything into the library ------------------------------------------------------
    def build(self, chunk=None):
        vega, lib = self.vega, self.lib
        self.need_fvoigt = None
        self.need_uvshot = False
        infos = [self.describe_corr(name) for name in vega.corr_items]
        ells = infos[0]['ells']
        if any(i['ells'] != ells for i in infos):
            raise NotImplementedError('All correlations must share ell_max.')

        fid = vega.fiducial
        k = _f64(fid['k'])
        cfg0 = infos[0]['item'].config['model']
        nmu = cfg0.getint('num_bins_muk', 1000)
        for i in infos:
            for sect in ('model', 'metals'):
                if sect in i['item'].config and i['item'].config[sect].getint('num_bins_muk', 1000) != nmu:
                    raise NotImplementedError('All correlations must share num_bins_muk.')
        mu = (np.arange(nmu) + 0.5) / nmu
        wleg = _f64(np.stack([(1. / nmu) * (2 * ell + 1) * _legendre(ell, mu) for ell in ells]))
        pk_fid = _f64(fid['pk_full'] * ((1 + fid['z_fiducial']) / (1. + fid['z_eff']))**2)

        base = self.base_vector()
        theta_cols = np.ascontiguousarray([self.col(n) for n in self.sampled], dtype=np.int32)
        if theta_cols.size == 0:
            theta_cols = np.zeros(1, dtype=np.int32)
        handle = C.c_void_p()
        _lib.check(lib.vg_plan_create(
            base.size, _dptr(base), len(self.sampled), _dptr(theta_cols), k.size, _dptr(k),
            _dptr(_f64(fid['pk_full'])), _dptr(_f64(fid['pk_smooth'])), _dptr(pk_fid), nmu,
            _dptr(_f64(mu)), len(ells), _dptr(wleg), self.col('bao_amp'), C.byref(handle)))
        plan = DevicePlan(handle, self, infos)
        transform = vega._theta_transform(self.sampled) if hasattr(vega, '_theta_transform') else None
        if transform is not None and self.sampled:
            _lib.check(lib.vg_plan_set_theta_transform(handle, _dptr(transform[0]), _dptr(transform[1])))
            self.keep.append(transform)
        if chunk is not None:
            _lib.check(lib.vg_plan_set_chunk(handle, int(chunk)))
        if self.sampled:
            # the box theta lives in ([sample] limits, vega/vega_interface.py:737-790): the velocity-
            # dispersion tables of the device path are built over it (NaN = no limit = no tables)
            limits = getattr(vega, 'sample_params', {}).get('limits', {})
            lo = _f64([np.nan if limits.get(n, (None, None))[0] is None else limits[n][0] for n in self.sampled])
            hi = _f64([np.nan if limits.get(n, (None, None))[1] is None else limits[n][1] for n in self.sampled])
            _lib.check(lib.vg_plan_set_theta_limits(handle, _dptr(lo), _dptr(hi)))

        if self.need_fvoigt is not None:
            model = self.need_fvoigt
            if '/' in model:
                table = np.loadtxt(model)
            else:
                try:
                    path = utils.find_file(f'fvoigt_models/Fvoigt_{model}.txt')
                    table = np.loadtxt(path)
                except RuntimeError:
                    path = utils.find_file(f'fvoigt_models/Fvoigt_{model}.txt.gz')
                    with gzip.open(path, 'rt') as f:
                        table = np.loadtxt(io.StringIO(f.read()))
            x, f = _f64(table[:, 0]), _f64(table[:, 1])
            _lib.check(lib.vg_plan_set_fvoigt(handle, x.size, _dptr(x), _dptr(f)))
        if self.need_uvshot:
            tau, a = utils.uv_shotnoise_table()
            _lib.check(lib.vg_plan_set_uvshot_table(handle, tau.size, _dptr(_f64(tau)), _dptr(_f64(a))))

        if vega.priors:
            cols = np.ascontiguousarray([self.col(p, True) for p in vega.priors], dtype=np.int32)
            mean = _f64([pr[0] for pr in vega.priors.values()])
            sigma = _f64([pr[1] for pr in vega.priors.values()])
            _lib.check(lib.vg_plan_set_priors(handle, cols.size, _dptr(cols), _dptr(mean), _dptr(sigma)))

        hankel_ids = {}
        ells_arr = np.ascontiguousarray(ells, dtype=np.int32)
        for ci, info in enumerate(infos):
            item, data = info['item'], info['data']
            hkey = (info['legacy'], info['lowring'])
            if hkey not in hankel_ids:
                xk, op = hankel.build_operator_set(k, ells, info['legacy'], info['lowring'])
                hid = C.c_int32()
                _lib.check(lib.vg_plan_add_hankel(handle, len(ells), _dptr(ells_arr), xk.shape[1], _dptr(xk),
                                                  _dptr(op), 0 if info['legacy'] else 1, C.byref(hid)))
                hankel_ids[hkey] = hid.value
            info['hankel_id'] = hankel_ids[hkey]
            n_model = item.model_coordinates.r_grid.size
            n_dist = item.dist_model_coordinates.r_grid.size
            cid = C.c_int32()
            _lib.check(lib.vg_plan_add_corr(handle, hankel_ids[hkey], n_model, n_dist, C.byref(cid)))
            assert cid.value == ci
            for tab, opz2 in info['coords']:
                if tab.shape[1] != n_model:
                    raise ValueError('Metal coordinate grids must have the size of the model grid.')
                sid = C.c_int32()
                _lib.check(lib.vg_corr_add_coords(handle, ci, _dptr(tab), opz2, C.byref(sid)))
            for d in info['evals']:
                eid = C.c_int32()
                _lib.check(lib.vg_corr_add_eval(handle, ci, C.byref(d), C.byref(eid)))
            pk_sel = [_f64(fid['pk_full'] - fid['pk_smooth']), _f64(fid['pk_smooth']), _f64(fid['pk_full'])]
            for which, (ell, n_pow) in info['static_xi'].items():
                for sel in range(3):
                    tab = hankel.build_static_table(k, pk_sel[sel], ell, n_pow)
                    _lib.check(lib.vg_corr_set_static_xi(handle, ci, which, sel, tab.shape[1], _dptr(tab)))
            if any(m is not None for m in info['metal_mats']):
                rp_form = all(m is None or hasattr(m, 'rp_mat') for m in info['metal_mats'])
                for slot, mat in enumerate(info['metal_mats']):
                    if mat is not None and rp_form:     # rp-only new_metals matrices stay rp-only
                        rp_mat = _f64(mat.rp_mat)
                        _lib.check(lib.vg_corr_set_metal_matrix_rp(
                            handle, ci, slot, _dptr(rp_mat), rp_mat.shape[0], mat.rt_nbins))
                    elif mat is not None:
                        if hasattr(mat, 'rp_mat'):      # mixed with dense ones: A (x) 1_rt
                            mat = np.kron(mat.rp_mat, np.eye(mat.rt_nbins))
                        mat = _f64(mat)
                        assert mat.shape == (n_model, n_model), 'metal matrix shape'
                        _lib.check(lib.vg_corr_set_metal_matrix(handle, ci, slot, _dptr(mat)))
            if data is not None and item.has_distortion and not data.distortion_is_identity:
                dm = _f64(data.distortion_mat)
                assert dm.shape == (n_dist, n_model), 'distortion matrix shape'
                _lib.check(lib.vg_corr_set_distortion(handle, ci, _dptr(dm)))
            elif n_dist != n_model:
                raise ValueError('Distorted and undistorted grids differ but there is no distortion matrix.')
            if info['inst'] is not None:
                table, idx, default = info['inst']
                _lib.check(lib.vg_corr_set_inst_sys(handle, ci, _dptr(table), idx, default))
            if info['bb_terms']:
                terms = (BBTerm * len(info['bb_terms']))(*info['bb_terms'])
                pre = _f64(np.array(info['bb_design']['pre']).reshape(-1, n_model)) \
                    if info['bb_design']['pre'] else np.zeros((0, n_model))
                post = _f64(np.array(info['bb_design']['post']).reshape(-1, n_dist)) \
                    if info['bb_design']['post'] else np.zeros((0, n_dist))
                sky_pre = self.sky_table(item.model_coordinates)
                sky_post = self.sky_table(item.dist_model_coordinates)
                _lib.check(lib.vg_corr_set_broadband(
                    handle, ci, len(info['bb_terms']), C.cast(terms, C.c_void_p), pre.shape[0], _dptr(pre),
                    post.shape[0], _dptr(post), _dptr(sky_pre), _dptr(sky_post)))
            if data is not None:
                plan.set_data(ci, data)
        if getattr(vega, '_use_global_cov', False):
            inv = _f64(vega.masked_global_invcov)
            if getattr(vega, 'marginalize_in_fit', False):
                from scipy.linalg import block_diag
                blocks = []
                for name in vega.corr_items:
                    proj = vega.data[name].marg_projector()
                    blocks.append(np.eye(int(vega.data[name].model_mask.sum())) if proj is None else proj)
                proj = block_diag(*blocks)
                inv = _f64(proj.T @ inv @ proj)
            _lib.check(lib.vg_plan_set_global_invcov(handle, inv.shape[0], _dptr(inv)))
        plan.finalize()
        return plan


class DevicePlan:
    """Owner of a vg_plan handle + the torch-side buffers of the batched entry points."""

    def __init__(self, handle, builder, infos):
        import torch
        self.torch = torch
        self.handle = handle
        self.lib = builder.lib
        self.builder = builder
        self.sampled = list(builder.sampled)
        self.corr_names = [i['name'] for i in infos]
        self.n_dist = [i['item'].dist_model_coordinates.r_grid.size for i in infos]
        self.n_eval = [sum(1 for d in i['evals'] if not d.direct) for i in infos]
        self.infos = infos
        self.device = torch.device('cuda', torch.cuda.current_device())

    def __del__(self):
        try:
            if self.handle:
                self.lib.vg_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def set_data(self, ci, data, monte_carlo=False):
        """Masked data vector + inverse covariance of correlation ci
        (vega/vega_interface.py:307-316; MC mode swaps in the mock and the scaled C^-1).
        With marginalize-in-fit the template fit is folded into the matrix (Data.marg_projector)."""
        mask_idx = np.ascontiguousarray(np.nonzero(data.model_mask)[0], dtype=np.int32)
        vec = _f64(data.masked_mc_mock if monte_carlo else data.masked_data_vec)
        if vec.size != mask_idx.size:
            raise ValueError('Masked data and masked model have different sizes.')
        proj = data.marg_projector() if getattr(self.builder.vega, 'marginalize_in_fit', False) else None
        if data.cov_is_identity and not monte_carlo and proj is None:
            inv = None
        else:
            inv = _f64(data.scaled_inv_masked_cov if monte_carlo else data.inv_masked_cov)
            if proj is not None:
                inv = _f64(proj.T @ inv @ proj)
        # the same mask and matrix as the last upload (a mock swapped in): the data vector alone, so the
        # device keeps the matrix, its square root and the operators built on them
        key = (mask_idx.tobytes(), None if inv is None else (inv.shape, float(inv.sum()), float(np.trace(inv)),
                                                             float(inv[0, -1]), float(inv[inv.shape[0] // 2, 0])))
        if not hasattr(self, '_data_keys'):
            self._data_keys = {}
        if self._data_keys.get(ci) == key:
            _lib.check(self.lib.vg_corr_set_data_vector(self.handle, ci, mask_idx.size, _dptr(vec)))
            return
        self._data_keys[ci] = key
        _lib.check(self.lib.vg_corr_set_data(self.handle, ci, mask_idx.size, _dptr(mask_idx), _dptr(vec),
                                             _dptr(inv) if inv is not None else None))
        # a square root of the matrix just uploaded: the device folds it into its operators and drops
        # the product with C^-1 (vg_corr_set_whitening; per-correlation covariances only -- and not the
        # projected, singular matrix of marginalize-in-fit)
        vega = self.builder.vega
        if inv is not None and proj is None and not getattr(vega, '_use_global_cov', False):
            w = self._whitening(inv)
            if w is not None:
                _lib.check(self.lib.vg_corr_set_whitening(self.handle, ci, mask_idx.size, _dptr(w)))

    def _whitening(self, inv):
        """Upper Cholesky factor w of inv (inv = w^T w) or None when inv is not positive definite."""
        torch = self.torch
        try:
            low = torch.linalg.cholesky(torch.as_tensor(inv, dtype=torch.float64, device=self.device))
        except Exception:
            return None
        w = low.T.contiguous().cpu().numpy()
        return _f64(w) if np.all(np.isfinite(w)) else None

    def finalize(self):
        _lib.check(self.lib.vg_plan_finalize(self.handle, self._stream()))
        self.static_status = self.lib.vg_plan_static_status(self.handle)

    def set_templates(self, pk_full, pk_smooth):
        """New peak / smooth templates (vg_plan_set_templates: runs the finalize pass again)."""
        _lib.check(self.lib.vg_plan_set_templates(self.handle, _dptr(_f64(pk_full)), _dptr(_f64(pk_smooth)),
                                                  self._stream()))
        self.static_status = self.lib.vg_plan_static_status(self.handle)

    def set_direct_pk(self, pk):
        """Isotropic P(k) on the template's k grid for the direct mode (OPT_DIRECT_PK)."""
        pk = _f64(pk)
        if pk.size != self.builder.vega.fiducial['k'].size:
            raise ValueError('direct_pk must be sampled on the k grid of the fiducial template.')
        _lib.check(self.lib.vg_plan_set_direct_pk(self.handle, _dptr(pk)))

    def set_option(self, option, value):
        """Run-time switches of the CUDA library (include/vega_b200.h: VG_OPT_*)."""
        _lib.check(self.lib.vg_plan_set_option(self.handle, int(option), int(value)))

    def n_moment(self, ci):
        """Evaluations of correlation ci computed in moment form (csrc/vg_moments.cuh)."""
        return self.lib.vg_corr_n_moment(self.handle, ci)

    def velo_tables(self, ci):
        """(evaluations, Chebyshev orders, error measured at finalize) of the velocity-dispersion
        tables of correlation ci (csrc/vg_moments.cuh); (0, 0, 0.0) when they are off."""
        orders, err = C.c_int32(0), C.c_double(0.0)
        n = self.lib.vg_corr_velo_tables(self.handle, ci, C.byref(orders), C.byref(err))
        return n, orders.value, err.value

    def fused_metal_operator(self, ci):
        """(rows, digits) of the fused distortion x metal operator of correlation ci on the chi2 path
        (rows = masked bins; (0, 0) when it is not used)."""
        digits = C.c_int32(0)
        rows = self.lib.vg_corr_fused_metal_operator(self.handle, ci, C.byref(digits), None)
        return rows, digits.value

    def whitened(self, ci):
        """True when the chi2 path of correlation ci runs on whitened operands (VG_OPT_WHITEN)."""
        return bool(self.lib.vg_corr_whitened(self.handle, ci))

    def velo_tables_active(self):
        """True when the last batch call ran on the velocity-dispersion tables."""
        return bool(self.lib.vg_plan_velo_tables_active(self.handle))

    def n_static_xi(self, ci):
        """Evaluations of correlation ci whose xi stage is a read-out of static tables."""
        return self.lib.vg_corr_n_static_xi(self.handle, ci)

    def n_folded(self, ci):
        """Metal evaluations of correlation ci that OPT_FOLD_METALS serves from static vectors."""
        return self.lib.vg_corr_n_folded(self.handle, ci)

    def ozaki_slices(self, ci):
        """(digits, {S: bound / max|xi|}) of the split-integer metal GEMM of correlation ci."""
        bounds = (C.c_double * 9)()
        digits = self.lib.vg_corr_ozaki_slices(self.handle, ci, bounds)
        return digits, {S: bounds[S] for S in range(3, 9)}

    def ozaki_dm_slices(self, ci):
        """(digits, {S: bound / max|D xi|}) of the distortion-matrix product of correlation ci."""
        bounds = (C.c_double * 9)()
        digits = self.lib.vg_corr_ozaki_dm_slices(self.handle, ci, bounds)
        return digits, {S: bounds[S] for S in range(3, 9)}

    def device_bytes(self):
        return self.lib.vg_plan_device_bytes(self.handle)

    # -- hot path ---------------------------------------------------------------------------
    def chi2_host(self, theta):
        """theta: host ndarray [B, P] -> (chi2[B], status[B]) host arrays.  One C-ABI call:
        H2D of theta, all kernels, D2H of chi2 and status."""
        theta = _f64(theta).reshape(-1, max(len(self.sampled), 1)) if len(self.sampled) else \
            np.zeros((np.asarray(theta).shape[0], 0))
        B = theta.shape[0]
        chi2 = np.empty(B, dtype=np.float64)
        status = np.empty(B, dtype=np.int32)
        tptr = _dptr(theta) if theta.size else _dptr(np.zeros(1))
        _lib.check(self.lib.vg_chi2_batch_host(self.handle, tptr, B, _dptr(chi2), _dptr(status),
                                               self._stream()))
        return chi2, status

    def chi2_device(self, theta_t, chi2_t=None, status_t=None):
        """theta_t: CUDA float64 tensor [B, P]; returns CUDA tensors, asynchronous."""
        torch = self.torch
        B = theta_t.shape[0]
        assert theta_t.is_cuda and theta_t.dtype == torch.float64 and theta_t.is_contiguous()
        if chi2_t is None:
            chi2_t = torch.empty(B, dtype=torch.float64, device=theta_t.device)
        if status_t is None:
            status_t = torch.empty(B, dtype=torch.int32, device=theta_t.device)
        _lib.check(self.lib.vg_chi2_batch(self.handle, C.c_void_p(theta_t.data_ptr()), B,
                                          C.c_void_p(chi2_t.data_ptr()), C.c_void_p(status_t.data_ptr()),
                                          self._stream()))
        if self.static_status:
            status_t |= self.static_status
        return chi2_t, status_t

    def set_mocks(self, ci, mocks):
        """Masked Monte Carlo mocks [n_mocks, n_mask] of correlation ci for chi2_mc_device."""
        mocks = _f64(np.atleast_2d(mocks))
        _lib.check(self.lib.vg_corr_set_mocks(self.handle, ci, mocks.shape[0], _dptr(mocks)))

    def chi2_mc_device(self, theta_t, mock_idx_t, chi2_t=None, status_t=None):
        """chi2 of vector b against mock mock_idx_t[b] (int32 CUDA tensor): lock-step mock fits."""
        torch = self.torch
        B = theta_t.shape[0]
        assert theta_t.is_cuda and theta_t.dtype == torch.float64 and theta_t.is_contiguous()
        assert mock_idx_t.is_cuda and mock_idx_t.dtype == torch.int32 and mock_idx_t.shape[0] == B
        if chi2_t is None:
            chi2_t = torch.empty(B, dtype=torch.float64, device=theta_t.device)
        if status_t is None:
            status_t = torch.empty(B, dtype=torch.int32, device=theta_t.device)
        _lib.check(self.lib.vg_chi2_batch_mc(self.handle, C.c_void_p(theta_t.data_ptr()),
                                             C.c_void_p(mock_idx_t.data_ptr()), B,
                                             C.c_void_p(chi2_t.data_ptr()), C.c_void_p(status_t.data_ptr()),
                                             self._stream()))
        if self.static_status:
            status_t |= self.static_status
        return chi2_t, status_t

    def model_device(self, theta_t):
        torch = self.torch
        B = theta_t.shape[0]
        outs = [torch.empty(B, n, dtype=torch.float64, device=theta_t.device) for n in self.n_dist]
        ptrs = (C.c_void_p * len(outs))(*[o.data_ptr() for o in outs])
        status_t = torch.empty(B, dtype=torch.int32, device=theta_t.device)
        _lib.check(self.lib.vg_model_batch(self.handle, C.c_void_p(theta_t.data_ptr()), B,
                                           C.cast(ptrs, C.c_void_p), C.c_void_p(status_t.data_ptr()),
                                           self._stream()))
        if self.static_status:
            status_t |= self.static_status
        return dict(zip(self.corr_names, outs)), status_t

    def theta_tensor(self, theta):
        torch = self.torch
        theta = np.asarray(theta, dtype=np.float64)
        P = len(self.sampled)
        theta = theta.reshape(-1, P) if P else np.zeros((theta.shape[0], 1))
        return torch.from_numpy(np.ascontiguousarray(theta)).to(self.device)

    def stage(self, corr, stage, theta, eval_index=0):
        """Per-stage tap (parity tests): returns a host ndarray [B, stage_size]."""
        torch = self.torch
        t = self.theta_tensor(theta)
        B = t.shape[0]
        per = self.lib.vg_stage_size(self.handle, corr, stage)
        out = torch.zeros(B, per, dtype=torch.float64, device=self.device)
        status = torch.zeros(B, dtype=torch.int32, device=self.device)
        _lib.check(self.lib.vg_stage_batch(self.handle, corr, eval_index, stage, C.c_void_p(t.data_ptr()), B,
                                           C.c_void_p(out.data_ptr()), C.c_void_p(status.data_ptr()),
                                           self._stream()))
        torch.cuda.synchronize()
        return out.cpu().numpy(), status.cpu().numpy()
