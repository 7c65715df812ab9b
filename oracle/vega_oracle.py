"""ORACLE (test infrastructure; never imported by the product package).

Plain numpy/scipy restatement, on the CPU, of the reference's model-evaluation
and likelihood path for ONE parameter dictionary at a time:

    parameter dict -> P(k,mu) -> P_ell(k) -> FFTLog -> xi_ell(r) -> xi(rp,rt)
                   -> metals, broadband, distortion matrix -> chi2 / log-lik

Each function cites the reference lines it restates.  The oracle consumes the
plain configuration objects produced by the init-time layer (parsed ini sections,
Coordinates, Data arrays, the fiducial P(k) dict); it shares no compute code with
the CUDA path.  It is pinned against (i) the reference's own golden vectors
(tests/data/picca_bench_data.fits, the constants of its tests/test_pk.py and
tests/test_vega.py) and (ii) outputs of the unmodified reference run in the build
container (tests/golden/*.npz, generator tests/golden/make_golden.py).
The mcfit boundary itself is unpinned upstream: see oracle/hankel.py.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference)
may import this module.
"""
import copy

import numpy as np
from scipy.interpolate import interp1d
from scipy.special import legendre, expn

from .hankel import P2xi, legacy_multipoles

DEFAULT_F = 0.970386


class OracleModelError(Exception):
    """Counterpart of VegaModelError (vega/utils.py:444-453)."""


# --------------------------------------------------------------------------------------
# parameters
# --------------------------------------------------------------------------------------
def tracer_bias_beta(pars, name):
    """vega/utils.py:45-81."""
    f = pars.get('growth_rate', DEFAULT_F)
    b, be, beta = (pars.get(p + name) for p in ('bias_', 'bias_eta_', 'beta_'))
    if b is None:
        b = be * f / beta
    if beta is None:
        beta = be * f / b
    return b, beta


def pair_bias_beta(pars, name1, name2):
    """vega/utils.py:84-108."""
    b1, beta1 = tracer_bias_beta(pars, name1)
    if name1 == name2:
        return b1, beta1, b1, beta1
    b2, beta2 = tracer_bias_beta(pars, name2)
    return b1, beta1, b2, beta2


def alpha_par_perp(scale_cfg, pars, corr_name=None, metal_corr=False):
    """(ap, at) selection of vega/scale_parameters.py:38-230."""
    gb = scale_cfg.getboolean
    full_shape = gb('full-shape', False)
    full_shape_alpha = gb('full-shape-alpha', False)
    smooth_scaling = gb('smooth-scaling', False)
    metal_scaling = gb('metal-scaling', False)
    two_alpha_smooth = gb('two-alpha-smooth', False)
    form = scale_cfg.get('cosmo fit func', 'ap_at')

    def from_names(n1, n2):
        if form == 'ap_at':
            return pars[n1], pars[n2]
        if form == 'aiso_epsilon':
            a, e = pars[n1], pars[n2]
            return a * (1 + e)**2, a / (1 + e)
        if form == 'phi_alpha':
            phi, alpha = pars[n1], pars[n2]
            return alpha / np.sqrt(phi), alpha * np.sqrt(phi)
        raise ValueError(form)

    bao_names = {'ap_at': ('ap', 'at'), 'aiso_epsilon': ('aiso', 'epsilon'),
                 'phi_alpha': ('phi', 'alpha')}[form]

    def fullshape():
        if form != 'phi_alpha' and not full_shape_alpha:
            raise ValueError('Only the "phi_alpha" parametrisation works with split full-shape.')
        if form == 'ap_at':
            return from_names('ap_full', 'at_full')
        if form == 'aiso_epsilon':
            return from_names('aiso_full', 'epsilon_full')
        phi_name = 'phi_full' if full_shape else 'phi_smooth'
        if full_shape_alpha:
            alpha_name = 'alpha_full'
        elif pars['peak']:
            alpha_name = 'alpha'
        elif two_alpha_smooth:
            alpha_name = f'alpha_smooth_{corr_name}'
        else:
            alpha_name = 'alpha_smooth'
        return from_names(phi_name, alpha_name)

    if metal_corr and not metal_scaling:
        return 1., 1.
    if full_shape:
        return fullshape()
    if pars['peak']:
        return from_names(*bao_names)
    if smooth_scaling:
        return fullshape()
    return 1., 1.


# --------------------------------------------------------------------------------------
# P(k, mu)
# --------------------------------------------------------------------------------------
class PkGrid:
    """(mu, k) grid and static tables of one PowerSpectrum instance
    (vega/power_spectrum.py:18-85)."""

    def __init__(self, cfg, fiducial, tracer1, tracer2, dataset_name, fvoigt_table=None):
        self.cfg = cfg
        self.t1, self.t2 = tracer1, tracer2
        self.name = dataset_name
        self.pair = f"{tracer1['name']}x{tracer2['name']}"
        self.k = np.asarray(fiducial['k'], float)
        nmu = cfg.getint('num_bins_muk', 1000)
        self.mu = ((np.arange(nmu) + 0.5) / nmu)[:, None]
        self.kpar = self.k * self.mu
        self.kperp = self.k * np.sqrt(1 - self.mu**2)
        self.pk_fid = fiducial['pk_full'] * ((1 + fiducial['z_fiducial'])
                                             / (1. + fiducial['z_eff']))**2
        self.fvoigt = fvoigt_table
        self.cache = {}

    def cached(self, key, fn):
        if key not in self.cache:
            self.cache[key] = fn()
        return self.cache[key]


def _sinc(x):
    return np.sin(x) / x


def uv_bias(g, pars, bias, beta):
    """vega/power_spectrum.py:224-261."""
    cfg = g.cfg
    eff = bias
    if cfg.getboolean('UVB-fluctuations', False):
        w = np.arctan(g.k * pars['lambda_uv']) / (g.k * pars['lambda_uv'])
        eff = eff + pars['bias_gamma'] * w / (1 + pars['bias_prim'] * w)
    if cfg.getboolean('HeII-reionization', False):
        w = np.arctan(g.k * pars['lambda_HeII']) / (g.k * pars['lambda_HeII'])
        eff = eff + pars['bias_gamma_e'] * w / (1 + pars['bias_prim'] * w)
    return eff, beta * bias / eff


def hcd_bias(g, pars, bias, beta):
    """vega/power_spectrum.py:263-380."""
    model = g.cfg.get('model-hcd')
    b_hcd = pars.get(f'bias_hcd_{g.pair}', None)
    if b_hcd is None:
        b_hcd = pars['bias_hcd']
    beta_hcd = pars.get(f'beta_hcd_{g.pair}', None)
    if beta_hcd is None:
        beta_hcd = pars['beta_hcd']
    if 'Rogers' in model:
        L0 = pars['L0_hcd']
        F = g.cached(('rogers', L0), lambda: np.exp(-L0 * g.kpar))
    elif 'fvoigt' in model:
        L0 = pars.get('L0_fvoigt', 1)
        F = g.cached(('fvoigt', L0), lambda: np.interp(
            L0 * g.kpar, g.fvoigt[:, 0], g.fvoigt[:, 1], left=1, right=0))
    elif 'sinc' in model:
        L0 = pars.get('L0_sinc', 1)
        F = g.cached(('sinc', L0), lambda: _sinc(g.kpar * L0))
    else:
        raise ValueError(f'Unknown hcd model {model}')
    eff = bias + b_hcd * F
    return eff, (bias * beta + b_hcd * beta_hcd * F) / eff


def arinyo(g, pars):
    """vega/power_spectrum.py:435-479 (first-call behaviour of the cache)."""
    n1, n2 = g.t1['name'], g.t2['name']
    two = 'LY' in n1 and 'LY' in n2
    one = 'LY' in n1 or 'LY' in n2
    q1, kv, av, bv, kp = (pars['dnl_arinyo_' + s] for s in ('q1', 'kv', 'av', 'bv', 'kp'))
    q2 = pars.get('dnl_arinyo_q2', 0)

    def build():
        d2 = g.k**3 * g.pk_fid / (2 * np.pi**2)
        growth = q1 * d2 + q2 * d2**2
        vel = (g.k / kv)**av * np.abs(g.mu)**bv
        dnl = np.exp(growth * (1 - vel) - (g.k / kp) * (g.k / kp))
        if np.any(np.isnan(dnl)) or np.any(np.isinf(dnl)):
            raise OracleModelError('Arinyo NaN/Inf')
        if two:
            return dnl
        if one:
            return np.sqrt(dnl)
        return np.ones(dnl.shape)
    return g.cached(('arinyo', q1, q2, kv, av, bv, kp), build)


def mcdonald(g):
    """vega/power_spectrum.py:419-433."""
    kvel = 1.22 * (1 + g.k / 0.923)**0.451
    dnl = (g.k / 6.4)**0.569 - (g.k / 15.3)**2.01 - (g.k * g.mu / kvel)**1.5
    return np.exp(dnl)


def binning_kernel(g, pars):
    """vega/power_spectrum.py:481-502."""
    drp = pars.get(f'par binsize {g.name}', g.cfg.getfloat('bin_size_rp'))
    drt = pars.get(f'per binsize {g.name}', g.cfg.getfloat('bin_size_rt'))
    out = 1.
    if drp != 0:
        out = out * _sinc(g.kpar * drp / 2)
    if drt != 0:
        out = out * _sinc(g.kperp * drt / 2)
    return out


def peak_damping(g, pars):
    """vega/power_spectrum.py:382-417."""
    sp, st = pars.get('sigmaNL_par'), pars.get('sigmaNL_per')
    f = pars.get('growth_rate')
    if sp is None and st is not None:
        sp = st * (1 + f)
    elif st is None and sp is not None:
        st = sp / (1 + f)
    elif sp is None and st is None:
        raise ValueError('No parameters for peak NL found.')
    return g.cached(('peaknl', sp, st),
                    lambda: np.exp(-(g.kpar**2 * sp**2 + g.kperp**2 * st**2) / 2))


def _gauss(g, sp, st):
    return np.exp(-(g.kpar**2 * sp**2 + g.kperp**2 * st**2) / 2)


def fullshape_gauss(g, pars):
    """vega/power_spectrum.py:504-553."""
    n1, n2 = g.t1['name'], g.t2['name']
    main = n1 in ['LYA', 'QSO'] and n2 in ['LYA', 'QSO']
    if 'par_sigma_smooth' in pars or 'per_sigma_smooth' in pars:
        sp, st = pars.get('par_sigma_smooth'), pars.get('per_sigma_smooth')
        if sp is None and st is None:
            raise ValueError('fullshape gaussian smoothing without parameters')
        sp = st if sp is None else sp
        st = sp if st is None else st
        return _gauss(g, sp, st)**2
    if 'par_sigma_smooth_metals' in pars and 'per_sigma_smooth_metals' in pars and not main:
        return _gauss(g, pars['par_sigma_smooth_metals'], pars['per_sigma_smooth_metals'])**2
    return (_gauss(g, pars[f'par_sigma_smooth_{n1}'], pars[f'per_sigma_smooth_{n1}'])
            * _gauss(g, pars[f'par_sigma_smooth_{n2}'], pars[f'per_sigma_smooth_{n2}']))


def fullshape_exp(g, pars):
    """vega/power_spectrum.py:560-586."""
    gs = g.kpar**2 * pars['par_sigma_smooth']**2 + g.kperp**2 * pars['per_sigma_smooth']**2
    es = np.abs(g.kpar) * pars['par_exp_smooth']**2 + np.abs(g.kperp) * pars['per_exp_smooth']**2
    return np.exp(-gs / 2) * np.exp(-es)


def velocity_dispersion(g, pars, kind):
    """vega/power_spectrum.py:588-636."""
    out = np.ones(g.kpar.shape)
    for tr in (g.t1, g.t2):
        if tr['type'] != 'discrete':
            continue
        if kind == 'gauss':
            out = out * np.exp(-0.25 * (g.kpar * pars['sigma_velo_disp_gauss_' + tr['name']])**2)
        else:
            out = out / np.sqrt(1 + (g.kpar * pars['sigma_velo_disp_lorentz_' + tr['name']])**2)
    return out


def pk_model(g, pk_lin, pars, fast_metals=False):
    """P(k, mu) on the (n_mu, n_k) grid: vega/power_spectrum.py:87-196."""
    cfg = g.cfg
    n1, n2 = g.t1['name'], g.t2['name']
    b1, beta1, b2, beta2 = pair_bias_beta(pars, n1, n2)
    if cfg.getboolean('UVB-fluctuations', False) or cfg.getboolean('HeII-reionization', False):
        if n1 == 'LYA':
            b1, beta1 = uv_bias(g, pars, b1, beta1)
        if n2 == 'LYA':
            b2, beta2 = uv_bias(g, pars, b2, beta2)
    if cfg.get('model-hcd', None) is not None:
        if n1 == 'LYA':
            b1, beta1 = hcd_bias(g, pars, b1, beta1)
        if n2 == 'LYA':
            b2, beta2 = hcd_bias(g, pars, b2, beta2)

    kaiser = (1 + beta1 * g.mu**2) * (1 + beta2 * g.mu**2)
    if not fast_metals:
        kaiser = kaiser * (b1 * b2)
    pk = pk_lin * kaiser

    skip_nl = cfg.getboolean('skip-nl-model-in-peak', False) and pars['peak']
    if 'small scale nl' in cfg and not skip_nl:
        kind = cfg.get('small scale nl')
        if 'arinyo' in kind:
            pk = pk * arinyo(g, pars)
        elif 'mcdonald' in kind:
            pk = pk * mcdonald(g)
        else:
            raise ValueError("Incorrect 'small scale nl' specified")
    if cfg.getboolean('model binning', True):
        pk = pk * g.cached('Gk', lambda: binning_kernel(g, pars))
    if 'mock-bin-size' in cfg:
        size = cfg.getfloat('mock-bin-size')
        sm = {f'par binsize {g.name}': size, f'per binsize {g.name}': size}
        los = cfg.get('mock-los-smoothing')
        if los == 'growth':
            sm[f'par binsize {g.name}'] *= 1 + pars['growth_rate']
        elif los == 'amplitude':
            sm[f'par binsize {g.name}'] *= 1 + pars['los_smooth_amp']
        elif los == 'only-los':
            sm[f'per binsize {g.name}'] = 0
        elif los is not None:
            raise ValueError(f'Unknown mock LOS smoothing option {los}.')
        pk = pk * binning_kernel(g, sm)
    if pars['peak']:
        pk = pk * peak_damping(g, pars)
    if 'fullshape smoothing' in cfg and not skip_nl:
        kind = cfg.get('fullshape smoothing')
        if 'gauss' in kind:
            pk = pk * fullshape_gauss(g, pars)
        elif 'exp' in kind:
            pk = pk * fullshape_exp(g, pars)
        else:
            raise ValueError('"fullshape smoothing" must be of type "gauss" or "exp".')
    if 'velocity dispersion' in cfg:
        kind = cfg.get('velocity dispersion')
        # 'lorentz_gauss' contains 'gauss' and so takes the gauss branch (:178-186)
        if 'gauss' in kind:
            pk = pk * velocity_dispersion(g, pars, 'gauss')
        elif 'lorentz' in kind:
            pk = pk * velocity_dispersion(g, pars, 'lorentz')
        else:
            raise ValueError('"velocity dispersion" must be of type "gauss" or "lorentz".')
    scale = cfg.getfloat('pk-damping-scale', None)
    if scale is not None:
        pk = pk * np.exp(-scale**2 * g.k.astype(float)**cfg.getint('pk-damping-power', 2) / 2)
    return pk


# --------------------------------------------------------------------------------------
# P(k, mu) -> xi(r, mu)
# --------------------------------------------------------------------------------------
class Transform:
    """Multipole projection + Hankel transform + spline (vega/pktoxi.py:17-163)."""

    def __init__(self, g, model_cfg):
        self.k = g.k
        self.mu = g.mu
        self.dmu = 1 / len(g.mu)
        ell_max = model_cfg.getint('ell_max', 6)       # note: 'ell-max' in ini files is ignored
        self.ells = tuple(int(x) for x in np.arange(0, ell_max + 1, 2))
        self.legacy = model_cfg.getboolean('old_fftlog', False)
        assert not model_cfg.getboolean('fht_extrap', False)
        lowring = model_cfg.getboolean('fht_lowring', True)
        self.leg_k = {ell: legendre(ell)(self.mu) for ell in self.ells}
        self.hankel = {} if self.legacy else {
            ell: P2xi(self.k, l=ell, lowring=lowring) for ell in self.ells}

    def multipoles(self, pk):
        """vega/pktoxi.py:79-97."""
        return np.array([np.sum(self.dmu * self.leg_k[ell] * pk, axis=0) * (2 * ell + 1)
                         for ell in self.ells])

    def xi(self, r, mu, pk, single_ell=-1):
        """vega/pktoxi.py:99-163 (mcfit path) and :281-319 (legacy path)."""
        ells = self.ells if single_ell < 0 else [single_ell]
        if self.legacy:
            tab = legacy_multipoles(r, self.k, pk, ells, self.mu, self.dmu)
            if single_ell >= 0:
                return tab[single_ell // 2]
            return sum(tab[ell // 2] * legendre(ell)(mu) for ell in ells)

        total = np.zeros(len(r))
        nz = r != 0
        for ell in ells:
            pk_ell = np.sum(self.dmu * legendre(ell)(self.mu) * pk, axis=0) * (2 * ell + 1)
            r_fft, xi_fft = self.hankel[ell](pk_ell)
            spline = interp1d(np.log(r_fft), xi_fft, kind='cubic')
            xi_ell = np.zeros(len(r))
            try:
                xi_ell[nz] = spline(np.log(r[nz]))
            except ValueError:
                raise OracleModelError('r outside the FFTLog range')
            if single_ell >= 0:
                return xi_ell
            total = total + xi_ell * legendre(ell)(mu)
        return total


def rescale(r, mu, ap, at, delta_rp=0.):
    """vega/correlation_func.py:200-236."""
    nz = r != 0
    rp = ap * (r[nz] * mu[nz] + delta_rp)
    rt = at * (r[nz] * np.sqrt(1 - mu[nz]**2))
    r2, mu2 = np.zeros(len(r)), np.zeros(len(r))
    r2[nz] = np.sqrt(rp**2 + rt**2)
    mu2[nz] = rp / r2[nz]
    return r2, mu2


def shotnoise_table(ntau=100, nrho=10000):
    """vega/correlation_func.py:597-626 (Gontcho A Gontcho et al. eq. 19)."""
    tau = np.linspace(0.01, 5, ntau)
    rho = np.linspace(0.0001, 10, nrho)
    drho = rho[1] - rho[0]
    a = np.array([-np.sum(drho * np.exp(-rho) / rho * (
        expn(1, rho * np.sqrt(1 + (t / rho)**2)) - expn(1, rho * np.abs(1 - t / rho))))
        for t in tau])
    return tau, a


class XiModel:
    """One CorrelationFunction instance (vega/correlation_func.py:20-161)."""

    def __init__(self, cfg, fiducial, coords, scale_cfg, tracer1, tracer2, metal_corr=False,
                 growth_fn=None, inst_table=None, cosmo=None):
        self.cfg, self.scale_cfg = cfg, scale_cfg
        self.r, self.mu, self.z = coords.r_grid, coords.mu_grid, coords.z_grid
        self.t1, self.t2 = tracer1, tracer2
        self.pair = f"{tracer1['name']}x{tracer2['name']}"
        self.metal_corr = metal_corr
        self.z_eff = fiducial['z_eff']
        self.single_ell = cfg.getint('single_multipole', -1)
        self.rescale_sys = cfg.getboolean('rescale-coords-systematics', False)
        self.rel_z = (1. + self.z) / (1 + self.z_eff)
        self.inst_table = inst_table
        # vega/correlation_func.py:58,238-274: separate redshifts of the two tracers of a cross
        self.rel_z_tracers = None
        if cfg.getboolean('new-bias-evolution', False) and tracer1['type'] != tracer2['type']:
            if cosmo is None:
                print('Warning: No cosmology found in xcf files, using mean redshift evolution.')
            else:
                half = self.r * self.mu / (2 * cosmo.get_dist_hubble(self.z))
                rel = {'discrete': (1. + self.z - half) / (1 + self.z_eff),
                       'continuous': (1. + self.z + half) / (1 + self.z_eff)}
                self.rel_z_tracers = (rel[tracer1['type']], rel[tracer2['type']])

        self.drp_name = None
        if tracer1['type'] == 'discrete' and tracer2['type'] != 'discrete':
            self.drp_name = 'drp_' + tracer1['name']
        elif tracer2['type'] == 'discrete' and tracer1['type'] != 'discrete':
            self.drp_name = 'drp_' + tracer2['name']

        # growth (vega/correlation_func.py:372-444)
        z_fid = fiducial['z_fiducial']
        om, ode = fiducial.get('Omega_m', None), fiducial.get('Omega_de', None)
        if ode is None:
            self.growth = ((1 + z_fid) / (1. + self.z))**2
        elif cfg.getboolean('old_growth_func', False):
            self.growth = growth_fn('old', self.z, z_fid, om, ode)
        else:
            self.growth = growth_fn('new', self.z, z_fid, om, ode)

        def flag(key):
            return key in cfg and cfg.getboolean(key)
        self.radiation = flag('radiation effects')
        self.relativistic = flag('relativistic correction')
        self.asymmetry = flag('standard asymmetry')
        self.uv_shot = flag('UVB-shotnoise')
        self.uv_tab = shotnoise_table() if self.uv_shot else None

    def bias_evolution(self, pars):
        """vega/correlation_func.py:276-370."""
        out = 1.
        for t, tr in enumerate((self.t1, self.t2)):
            rel_z = self.rel_z if self.rel_z_tracers is None else self.rel_z_tracers[t]
            key = 'z evol {}'.format(tr['name'])
            kind = self.cfg.get(key, 'standard') if key in self.cfg \
                else self.cfg.get('z evol', 'standard')
            if 'croom' in kind:
                assert tr['name'] == 'QSO'
                assert self.rel_z_tracers is None, 'Croom model is not supported with new bias evol'
                p0, p1 = pars['croom_par0'], pars['croom_par1']
                out = out * (p0 + p1 * (1. + self.z)**2) / (p0 + p1 * (1 + self.z_eff)**2)
            else:
                out = out * rel_z**pars['alpha_{}'.format(tr['name'])]
        return out

    def compute(self, pk, pk_lin, transform, pars):
        """vega/correlation_func.py:117-198."""
        drp = 0. if self.drp_name is None else pars.get(self.drp_name, 0.)
        ap, at = alpha_par_perp(self.scale_cfg, pars, self.pair, self.metal_corr)
        r2, mu2 = rescale(self.r, self.mu, ap, at, drp)
        xi = transform.xi(r2, mu2, pk, self.single_ell)
        xi = xi * self.bias_evolution(pars)
        xi = xi * self.growth

        if self.radiation and not pars['peak']:
            xi = xi + self.qso_radiation(pars, r2, mu2)
        if self.relativistic or self.asymmetry:
            # vega/correlation_func.py:491-551: note get_ap_at WITHOUT corr_name
            drp2 = pars.get(self.drp_name, 0.)
            ap2, at2 = alpha_par_perp(self.scale_cfg, pars, None, self.metal_corr)
            r3, mu3 = rescale(self.r, self.mu, ap2, at2, drp2)
            if self.relativistic:
                t = legacy_multipoles(r3, transform.k, pk_lin, [1, 3], transform.mu,
                                      transform.dmu, tform='rel')
                xi = xi + pars['Arel1'] * t[0] * legendre(1)(mu3) \
                    + pars['Arel3'] * t[1] * legendre(3)(mu3)
            if self.asymmetry:
                t = legacy_multipoles(r3, transform.k, pk_lin, [0, 2], transform.mu,
                                      transform.dmu, tform='asy')
                xi = xi + (pars['Aasy0'] * t[0] - pars['Aasy2'] * t[1]) * r3 * legendre(1)(mu3) \
                    + pars['Aasy3'] * t[1] * r3 * legendre(3)(mu3)
        if self.uv_shot:
            xi = xi + self.uv_shotnoise(pars, r2, mu2)
        return xi

    def qso_radiation(self, pars, r2, mu2):
        """vega/correlation_func.py:446-489."""
        drp = pars.get(self.drp_name, 0.)
        if self.rescale_sys:
            rp, rt = r2 * mu2 + drp, r2 * np.sqrt(1 - mu2**2)
        else:
            rp, rt = self.r * self.mu + drp, self.r * np.sqrt(1 - self.mu**2)
        rs = np.sqrt(rp**2 + rt**2)
        ms = rp / rs
        out = pars['qso_rad_strength'] / rs**2 * (1 - pars['qso_rad_asymmetry'] * (1 - ms**2))
        return out * np.exp(-rs * ((1 + ms) / pars['qso_rad_lifetime']
                                   + 1 / pars['qso_rad_decrease']))

    def uv_shotnoise(self, pars, r2, mu2):
        """vega/correlation_func.py:628-686."""
        lam = pars['lambda_uv']
        bg = pars['bias_gamma'] if 'bias_gamma' in pars else pars['bias_gamma_e']
        r = np.sqrt(r2**2 + mu2**2) if self.rescale_sys else self.r
        tau, a = self.uv_tab
        return bg**2 * pars['uv_shotnoise_amp'] * lam / r * np.interp(
            r / lam, tau, a, left=a[0], right=0)

    def instrumental(self, pars, bin_size_rp):
        """vega/correlation_func.py:553-595."""
        rp = self.r * self.mu
        rt = self.r * np.sqrt(1 - self.mu**2)
        amp = pars.get('desi_inst_sys_amp', 0.0003189935987295203)
        w = (rp > 0) & (rp < bin_size_rp)
        out = np.zeros(rt.shape)
        out[w] = amp * interp1d(self.inst_table[0], self.inst_table[1], kind='linear')(rt[w])
        return out


# --------------------------------------------------------------------------------------
# growth (init-time)
# --------------------------------------------------------------------------------------
_GROWTH_CACHE = {}


def growth_squared(kind, z, z_fid, om, ode):
    """(D(z)/D(z_fid))^2: vega/utils.py:128-227 ('new') and
    vega/correlation_func.py:405-444 ('old')."""
    from scipy.integrate import quad

    def hub(zz):
        return np.sqrt(om * (1 + zz)**3 + ode + (1 - om - ode) * (1 + zz)**2)

    def integrand(a):
        return 1. / (a * hub(1 / a - 1))**3

    key = (kind, om, ode)
    if key not in _GROWTH_CACHE:
        if kind == 'new':
            zs = np.linspace(0, 10, 1000)
            d = [5. / 2. * om * hub(zz) * quad(integrand, 0, 1 / (1 + zz))[0] for zz in zs]
            _GROWTH_CACHE[key] = interp1d(zs, np.array(d), kind='cubic')
        else:
            zs = 5. * np.arange(100, dtype=float) / 99
            d = [5 / 2. * om * hub(zz) * quad(integrand, 0, 1 / (1 + zz))[0] for zz in zs]
            _GROWTH_CACHE[key] = interp1d(zs, np.array(d))
    fn = _GROWTH_CACHE[key]
    return (fn(z) / fn(z_fid))**2


# --------------------------------------------------------------------------------------
# broadband
# --------------------------------------------------------------------------------------
def parse_broadband(bb_section, cf_name):
    """vega/broadband_poly.py:23-72."""
    terms = {'pre-add': [], 'pre-mul': [], 'post-add': [], 'post-mul': []}
    for i, spec in enumerate(bb_section.values()):
        w = spec.split()
        if len(w) not in (5, 6):
            raise ValueError('Broadband setup must have 5 or 6 elements.')
        name = f'BB-{cf_name}-{i}-{w[5]}' if len(w) > 5 else f'BB-{cf_name}-{i} {w[0]} {w[1]} {w[2]}'
        terms[f'{w[1]}-{w[0]}'].append({
            'name': name, 'func': 'broadband' if len(w) == 5 else w[5], 'coordinates': w[2],
            'r1': tuple(int(x) for x in w[3].split(':')),
            'r2': tuple(int(x) for x in w[4].split(':'))})
    return terms


def broadband(terms, pars, pos, coords):
    """vega/broadband_poly.py:74-198."""
    total = None
    for term in terms[pos]:
        if term['func'] == 'broadband':
            if term['coordinates'] == 'r,mu':
                x1, x2 = coords.r_grid / 100., coords.mu_grid
            else:
                x1 = coords.r_grid / 100. * coords.mu_grid
                x2 = coords.r_grid / 100. * np.sqrt(1 - coords.mu_grid**2)
            lo1, hi1, s1 = term['r1']
            lo2, hi2, s2 = term['r2']
            poly = np.zeros(len(x1))
            for i in np.arange(lo1, hi1 + 1, s1):
                for j in np.arange(lo2, hi2 + 1, s2):
                    poly = poly + pars[f"{term['name']} ({i},{j})"] * x1**i * x2**j
        else:
            scale = pars[term['name'] + '-scale-sky']
            sigma = pars[term['name'] + '-sigma-sky']
            poly = scale / (sigma * np.sqrt(2. * np.pi)) * np.exp(-0.5 * (coords.rt_grid / sigma)**2)
            w = (coords.rp_grid >= 0.) & (coords.rp_grid < coords.rp_binsize)
            poly[~w] = 0.
        if total is None:
            total = 1 + poly if 'mul' in pos else poly
        elif 'mul' in pos:
            total = total * (1 + poly)
        else:
            total = total + poly
    if total is None:
        total = 1 if 'mul' in pos else 0
    return total


# --------------------------------------------------------------------------------------
# one correlation
# --------------------------------------------------------------------------------------
class CorrelationModel:
    """Peak/smooth model of one correlation incl. metals, broadband and distortion
    (vega/model.py:13-208, vega/metals.py:22-367)."""

    def __init__(self, corr_item, data, fiducial, scale_cfg, tables=None):
        tables = tables or {}
        self.item, self.data, self.fid, self.scale_cfg = corr_item, data, fiducial, scale_cfg
        cfg = corr_item.config
        mcfg = cfg['model']
        mcfg['bin_size_rp'] = str(corr_item.data_coordinates.rp_binsize)
        mcfg['bin_size_rt'] = str(corr_item.data_coordinates.rt_binsize)
        fv = tables.get('fvoigt')
        self.grid = PkGrid(mcfg, fiducial, corr_item.tracer1, corr_item.tracer2,
                           corr_item.name, fv)
        self.transform = Transform(self.grid, mcfg)
        self.xi_core = XiModel(mcfg, fiducial, corr_item.model_coordinates, scale_cfg,
                               corr_item.tracer1, corr_item.tracer2, growth_fn=growth_squared,
                               inst_table=tables.get('desi_inst'),
                               cosmo=getattr(corr_item, 'cosmo', None))
        self.bb = parse_broadband(cfg['broadband'], corr_item.name) \
            if 'broadband' in cfg else None
        self.inst_flag = mcfg.getboolean('desi-instrumental-systematics', False)
        self.no_metal_decomp = mcfg.getboolean('no-metal-decomp', True)

        self.fast_metals = mcfg.getboolean('fast_metals', False)
        self.fast_metal_bias = mcfg.getboolean('fast_metal_bias', True) or self.fast_metals \
            or mcfg.getboolean('separate-metal-auto-biases', False)
        self.separate_auto = mcfg.getboolean('separate-metal-auto-biases', False)
        self.single_metal_beta = mcfg.getboolean('single-metal-beta', False)
        self.metal_growth_rate = fiducial.get('growth_rate', None)
        self.metal = {}
        self.cache_mm = {}
        if corr_item.has_metals:
            metcfg = cfg['metals']
            metcfg['bin_size_rp'] = str(corr_item.data_coordinates.rp_binsize)
            metcfg['bin_size_rt'] = str(corr_item.data_coordinates.rt_binsize)
            for pair in corr_item.metal_correlations:
                tr1 = corr_item.tracer_catalog[pair[0]]
                tr2 = corr_item.tracer_catalog[pair[1]]
                coords, mat = data.get_metal(pair)
                g = PkGrid(metcfg, fiducial, tr1, tr2, corr_item.name, fv)
                self.metal[pair] = (g, Transform(g, mcfg), XiModel(
                    metcfg, fiducial, coords, scale_cfg, tr1, tr2, metal_corr=True,
                    growth_fn=growth_squared, cosmo=getattr(corr_item, 'cosmo', None)), mat)

    # -- metals: vega/metals.py:209-367 ------------------------------------------------
    def _metal_xi(self, pars, pk_lin, pair, fast, apply_matrix=True):
        g, tr, xim, mat = self.metal[pair]
        pk = pk_model(g, pk_lin, pars, fast_metals=fast)
        xi = xim.compute(pk, pk_lin, tr, pars)
        main = [self.item.tracer1['name'], self.item.tracer2['name']]
        if main[0] == main[1] and pair[0] != pair[1]:
            xi = xi * 2
        if apply_matrix and mat is not None:
            xi = mat.dot(xi)
        return xi

    def metals(self, pars, pk_lin):
        lp = copy.deepcopy(pars)
        if self.fast_metals and 'growth_rate' in lp and self.metal_growth_rate is not None:
            lp['growth_rate'] = self.metal_growth_rate
        main = [self.item.tracer1['name'], self.item.tracer2['name']]
        total = np.zeros(self.item.model_coordinates.rp_grid.size)
        cross_cache = {}
        for pair in self.item.metal_correlations:
            n1, n2 = pair
            if self.single_metal_beta:
                for n in pair:
                    if n not in main:
                        lp[f'beta_{n}'] = lp['beta_metals']
            b1, beta1, b2, beta2 = pair_bias_beta(lp, n1, n2)
            with_main = n1 in main or n2 in main
            weight = b1 * b2
            if not with_main and self.separate_auto and n1 != n2:
                if f'bias_{n1}_{n2}' in lp:
                    weight = weight * lp[f'bias_{n1}_{n2}']
                elif f'bias_{n2}_{n1}' in lp:
                    weight = weight * lp[f'bias_{n2}_{n1}']
                else:
                    raise ValueError(f'no bias_{n1}_{n2} parameter')
            if self.fast_metals and with_main:
                # cache keyed on (beta1, beta2, all parameters) WITHOUT the pair name
                # (vega/metals.py:194-202): the first metal x main xi with equal betas
                # is re-used for the following ones, only the matrix differs
                key = (beta1, beta2) + tuple(lp[k] for k in sorted(lp.keys()))
                if key not in cross_cache:
                    cross_cache[key] = self._metal_xi(lp, pk_lin, pair, True, apply_matrix=False)
                xi = cross_cache[key]
                mat = self.metal[pair][3]
                total = total + weight * (xi if mat is None else mat.dot(xi))
            elif self.fast_metals:
                if pair not in self.cache_mm:
                    self.cache_mm[pair] = self._metal_xi(lp, pk_lin, pair, True)
                total = total + weight * self.cache_mm[pair]
            else:
                xi = self._metal_xi(lp, pk_lin, pair, self.fast_metal_bias)
                total = total + (weight * xi if self.fast_metal_bias else xi)
        return total

    # -- vega/model.py:79-208 --------------------------------------------------------
    def component(self, pars, pk_lin, name, xi_metals=None):
        pk = pk_model(self.grid, pk_lin, pars)
        if self.item.model_pk:
            return self.transform.multipoles(pk)
        xi = self.xi_core.compute(pk, pk_lin, self.transform, pars)
        if self.item.has_metals:
            if self.no_metal_decomp and xi_metals is not None:
                xi = xi + xi_metals
            elif not self.no_metal_decomp:
                xi = xi + self.metals(pars, pk_lin)
        if self.inst_flag and name != 'peak':
            xi = xi + self.xi_core.instrumental(pars, self.item.data_coordinates.rp_binsize)
        if self.bb is not None:
            xi = xi * broadband(self.bb, pars, 'pre-mul', self.item.model_coordinates)
            xi = xi + broadband(self.bb, pars, 'pre-add', self.item.model_coordinates)
        if self.data is not None and self.item.has_distortion \
                and not self.data.distortion_is_identity:
            xi = self.data.distortion_mat.dot(xi)
        if self.bb is not None:
            xi = xi * broadband(self.bb, pars, 'post-mul', self.item.dist_model_coordinates)
            xi = xi + broadband(self.bb, pars, 'post-add', self.item.dist_model_coordinates)
        return xi

    def compute(self, pars, pk_full, pk_smooth):
        pars['peak'] = True
        peak = self.component(pars, pk_full - pk_smooth, 'peak')
        pars['peak'] = False
        xi_metals = None
        if self.item.has_metals and self.no_metal_decomp:
            xi_metals = self.metals(pars, pk_full)
        smooth = self.component(pars, pk_smooth, 'smooth', xi_metals=xi_metals)
        return pars['bao_amp'] * peak + smooth

    def compute_direct(self, pars, pk_full):
        pars['peak'] = False
        return self.component(pars, pk_full, 'full')


# --------------------------------------------------------------------------------------
# all correlations -> chi2 / log-likelihood
# --------------------------------------------------------------------------------------
class Oracle:
    """CPU counterpart of VegaInterface.compute_model / chi2 / log_lik
    (vega/vega_interface.py:208-446).

    ``setup`` is any object exposing the init-time state under the reference's names:
    corr_items, data, fiducial, params, priors, main_config, plus (optional) the global
    covariance members full_model_mask / masked_global_invcov / masked_global_log_cov_det.
    """

    def __init__(self, setup, tables=None, fast_metals=False):
        self.s = setup
        self.scale_cfg = setup.main_config['cosmo-fit type']
        self.tables = tables or {}
        self.models = {}
        for name, item in setup.corr_items.items():
            self.models[name] = CorrelationModel(
                item, setup.data[name], setup.fiducial, self.scale_cfg, self.tables)
            if fast_metals:
                self.models[name].fast_metals = True
                self.models[name].fast_metal_bias = True
        self.monte_carlo = False

    def local_params(self, params=None):
        """vega/vega_interface.py:389-421: merge, then -- on blinded data with offsets loaded
        (setup._rnsps) -- theta += pi - exp(v^2) for the blinded names (vega/utils.py:375-393) and
        the BLIND_FIXED_PARS (vega/utils.py:16-18) pinned to 1."""
        lp = copy.deepcopy(self.s.params)
        if params is not None:
            lp.update(params)
        rnsps = getattr(self.s, '_rnsps', None)
        if rnsps is not None:
            for par, val in rnsps.items():
                lp[par] += np.pi - np.exp(val**2)
            for par in lp:
                if par in ('ap_full', 'at_full', 'aiso_full', 'epsilon_full', 'phi_full'):
                    lp[par] = 1.
        return lp

    def compute_model(self, params=None, direct_pk=None):
        lp = self.local_params(params)
        out = {}
        for name in self.s.corr_items:
            if direct_pk is None:
                out[name] = self.models[name].compute(
                    lp, self.s.fiducial['pk_full'], self.s.fiducial['pk_smooth'])
            else:
                out[name] = self.models[name].compute_direct(lp, direct_pk)
        return out

    def prior_chi2(self, params=None):
        """vega/vega_interface.py:423-446,818-820."""
        lp = self.local_params(params)
        return sum((lp[p] - pr[0])**2 / pr[1]**2 for p, pr in self.s.priors.items())

    def compute_marg_coeff(self, model):
        """vega/vega_interface.py:546-579: template amplitudes = marg_diff2coeff_matrix . (d - m)."""
        s = self.s
        out = {}
        for n in s.corr_items:
            d = s.data[n]
            vec = d.masked_mc_mock if self.monte_carlo else d.masked_data_vec
            diff = vec - model[n][d.model_mask]
            if d.marg_diff2coeff_matrix is not None:
                out[n] = d.marg_diff2coeff_matrix.dot(diff)
        return out

    def chi2(self, params=None, direct_pk=None, return_marg_coeff=False):
        """vega/vega_interface.py:250-325."""
        try:
            model = self.compute_model(params, direct_pk)
        except OracleModelError:
            return (1e100, None) if return_marg_coeff else 1e100
        s = self.s
        marg_in_fit = getattr(s, 'marginalize_in_fit', False)
        marg_coeff = None
        if return_marg_coeff or marg_in_fit:          # :282-292
            marg_coeff = self.compute_marg_coeff(model)
        if marg_in_fit:
            for n in s.data:
                if s.data[n].marg_templates is not None:
                    model[n] = model[n] + s.data[n].marg_templates.dot(marg_coeff[n])
        if return_marg_coeff:
            return self._chi2_of_model(model, params), marg_coeff
        return self._chi2_of_model(model, params)

    def _chi2_of_model(self, model, params):
        s = self.s
        if getattr(s, '_use_global_cov', False):
            if self.monte_carlo:
                d = s.analysis.current_mc_mock
            else:
                d = np.concatenate([s.data[n].masked_data_vec for n in s.corr_items])
            m = np.concatenate([model[n] for n in s.corr_items])
            diff = d - m[s.full_model_mask]
            chi2 = diff.T.dot(s.masked_global_invcov.dot(diff))
        else:
            chi2 = 0
            for n in s.corr_items:
                m = model[n][s.data[n].model_mask]
                if self.monte_carlo:
                    diff = s.data[n].masked_mc_mock - m
                    chi2 += diff.T.dot(s.data[n].scaled_inv_masked_cov.dot(diff))
                else:
                    diff = s.data[n].masked_data_vec - m
                    chi2 += diff.T.dot(s.data[n].inv_masked_cov.dot(diff))
        return float(chi2 + self.prior_chi2(params))

    def log_lik(self, params=None, direct_pk=None):
        """vega/vega_interface.py:327-387."""
        s = self.s
        chi2 = self.chi2(params, direct_pk)
        log_norm = 0
        glob = getattr(s, '_use_global_cov', False)
        for n in s.corr_items:
            log_norm -= 0.5 * s.data[n].data_size * np.log(2 * np.pi)
            if not glob:
                log_norm -= 0.5 * (s.data[n].scaled_log_cov_det if self.monte_carlo
                                   else s.data[n].log_cov_det)
        if glob:
            log_norm -= 0.5 * s.masked_global_log_cov_det
        out = log_norm - 0.5 * chi2
        for pr in s.priors.values():
            out += -0.5 * np.log(2 * np.pi) - np.log(pr[1])
        return out
