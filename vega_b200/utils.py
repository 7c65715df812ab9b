"""Host-side helpers shared by the init-time layer of the B200 path.

Mirrors the parts of the reference's vega/utils.py the hot path depends on:
exceptions (vega/utils.py:444-453), file lookup (:230-268), the growth factor
(:128-227, init-time only) and the masked covariance inverse / log-determinant
(:271-318, init-time only).  The per-evaluation helpers of that file
(bias_beta, sinc, smoothing kernels) live on the device: see csrc/vg_pk.cuh.
"""
import os
from functools import lru_cache
from pathlib import Path

import numpy as np
from scipy.integrate import quad
from scipy.interpolate import interp1d

PACKAGE_DIR = Path(__file__).resolve().parent
REPO_DIR = PACKAGE_DIR.parent

# Extra directories searched by find_file (the directory of the main ini is
# appended by VegaInterface so relative paths next to it resolve)
_EXTRA_SEARCH_DIRS = []

BLIND_FIXED_PARS = ['ap_full', 'at_full', 'aiso_full', 'epsilon_full', 'phi_full']
DEFAULT_GROWTH_RATE = 0.970386
# blinded parameters and the tracers they are blinded for (vega/utils.py:21-25)
VEGA_BLINDED_PARS = {'phi_smooth': ['all'], 'growth_rate': ['all']}
# blinding strategy -> {'bao' | 'full-shape': npz file of offsets}.  The reference's table
# (vega/utils.py:331-341) points into a NERSC project directory and ships every entry as None;
# a site fills it in (tests register a synthetic file).
BLINDING_FILES = {'desi_y1': {'full-shape': None, 'bao': None},
                  'desi_y3': {'full-shape': None, 'bao': None}}


class VegaModelError(Exception):
    """Model-domain failure; chi2 turns it into the 1e100 sentinel
    (vega/vega_interface.py:268-279)."""


class VegaBoundsError(VegaModelError):
    pass


class VegaArinyoError(VegaModelError):
    pass


class VegaDeviceError(RuntimeError):
    """The CUDA library is missing or a kernel launch failed.  There is no CPU
    fallback: the product path fails loudly."""


def get_blinding(blind_pars, blinding_strat):
    """Offsets v of the blinded parameters from the strategy's file (vega/utils.py:320-372)."""
    assert blinding_strat is not None, 'Blinding failed, do not run!!!'
    print(f'Blinding parameters: {blind_pars}')
    if ('ap' in blind_pars) or ('at' in blind_pars) or ('alpha' in blind_pars):
        blinding_type = 'bao'
    elif ('growth_rate' in blind_pars) or ('phi_smooth' in blind_pars):
        blinding_type = 'full-shape'
    else:
        raise ValueError(f'No blinding implemented for parameters {blind_pars}')
    if blinding_strat not in BLINDING_FILES:
        raise ValueError(f'Unknown blinding version: {blinding_strat}.')
    blinding_file = BLINDING_FILES[blinding_strat][blinding_type]
    if blinding_file is None:
        return None
    blinding_file = Path(blinding_file)
    if not blinding_file.exists():
        raise ValueError(
            f'Blinding file not found: {blinding_file}. Full-shape analyses must be run at NERSC.')
    blinding = {}
    with np.load(blinding_file) as file:
        for par in blind_pars:
            if par not in VEGA_BLINDED_PARS:
                raise ValueError(f'Blinding for parameter {par} not implemented.')
            if par == 'alpha':
                dap, dat = float(file['ap']), float(file['at'])
                blinding[par] = np.sqrt(np.log(
                    np.pi - np.sqrt((1 + np.pi - np.exp(dap**2)) * (1 + np.pi - np.exp(dat**2))) + 1))
            else:
                blinding[par] = float(file[par])
    return blinding


def blinding_offset(value):
    """The shift a blinded parameter receives: theta += pi - exp(v^2) (vega/utils.py:375-393)."""
    return np.pi - np.exp(value**2)


def apply_blinding(params, blinding):
    """In-place, like the reference (vega/utils.py:375-393)."""
    for par, val in blinding.items():
        params[par] += blinding_offset(val)
    return params


def add_search_dir(path):
    path = Path(path)
    if path not in _EXTRA_SEARCH_DIRS:
        _EXTRA_SEARCH_DIRS.append(path)


def find_file(path):
    """Resolve an input path the way vega/utils.py:230-268 does: absolute (or
    cwd-relative) first, then the packaged models, then the tests tree, then the
    repository root; finally any directory registered with add_search_dir."""
    input_path = Path(os.path.expandvars(str(path)))
    if input_path.is_file():
        return input_path
    roots = [PACKAGE_DIR / 'models', REPO_DIR / 'tests', REPO_DIR / 'tests' / 'configs',
             REPO_DIR] + _EXTRA_SEARCH_DIRS
    for root in roots:
        candidate = root / input_path
        if candidate.is_file():
            return candidate
    raise RuntimeError('The path/file does not exists: ', input_path)


def _hubble(z, Omega_m, Omega_de):
    Omega_k = 1 - Omega_m - Omega_de
    return np.sqrt(Omega_m * (1 + z)**3 + Omega_de + Omega_k * (1 + z)**2)


def _growth_integrand(a, Omega_m, Omega_de):
    z = 1 / a - 1
    return 1. / (a * _hubble(z, Omega_m, Omega_de))**3


@lru_cache(maxsize=32)
def get_growth_interp(Omega_m, Omega_de):
    """Cubic interpolation of D(z) on 1000 nodes in z in [0, 10]
    (vega/utils.py:175-205; Dodelson eq. 7.77)."""
    z_grid = np.linspace(0, 10, 1000)
    growth = np.zeros(1000)
    for i, z in enumerate(z_grid):
        integral = quad(_growth_integrand, 0, 1 / (1 + z), args=(Omega_m, Omega_de))[0]
        growth[i] = 5. / 2. * Omega_m * _hubble(z, Omega_m, Omega_de) * integral
    return interp1d(z_grid, growth, kind='cubic')


def growth_function(z, Omega_m, Omega_de):
    return get_growth_interp(Omega_m, Omega_de)(z)


@lru_cache(maxsize=32)
def _growth_interp_old(Omega_m, Omega_de):
    """Deprecated growth table: 100 nodes in z in [0, 5], linear interpolation
    (vega/correlation_func.py:405-444, old_growth_func = True)."""
    z_grid = 5. * np.arange(100, dtype=float) / 99
    growth = np.zeros(100)
    for i, z in enumerate(z_grid):
        integral = quad(_growth_integrand, 0, 1 / (1 + z), args=(Omega_m, Omega_de))[0]
        growth[i] = 5 / 2. * Omega_m * _hubble(z, Omega_m, Omega_de) * integral
    return interp1d(z_grid, growth)


def growth_function_old(z, Omega_m, Omega_de):
    return _growth_interp_old(Omega_m, Omega_de)(z)


DEVICE_LINALG_MIN = int(os.environ.get('VEGA_B200_DEVICE_LINALG_MIN', 4096))
_factor_cache = {}


def _device_cholesky(masked_cov):
    """Cholesky factor of a large masked covariance on the GPU (cuSOLVER through torch; init
    time).  One factorisation serves the positive-definiteness check, the inverse
    and the log-determinant -- the reference factorises three times on the host
    (vega/utils.py:271-318).  Returns None below DEVICE_LINALG_MIN rows (host LAPACK, bit-identical
    to the reference), without a CUDA device, or when the matrix is not positive definite."""
    n = masked_cov.shape[0]
    if n < DEVICE_LINALG_MIN:
        return None
    import torch
    if not torch.cuda.is_available():
        return None
    # identity of the matrix: full sum, trace and a strided partial sum (one pass at memory speed)
    key = (n, float(masked_cov.sum()), float(np.trace(masked_cov)), float(masked_cov[::7, ::13].sum()),
           float(masked_cov[n // 2] @ masked_cov[n // 3]))
    if key not in _factor_cache:
        _factor_cache.clear()               # one matrix at a time: a 9540^2 factor is 728 MB
        dev = torch.from_numpy(np.ascontiguousarray(masked_cov, dtype=np.float64)).cuda()
        factor, info = torch.linalg.cholesky_ex(dev)
        _factor_cache[key] = factor if int(info) == 0 else None
    return _factor_cache[key]


def compute_masked_invcov(cov_mat, data_mask):
    """inv(cov[mask][:, mask]) (vega/utils.py:271-299)."""
    masked_cov = cov_mat[:, data_mask][data_mask, :]
    factor = _device_cholesky(masked_cov)
    if factor is not None:
        import torch
        inv = torch.cholesky_inverse(factor)
        return (0.5 * (inv + inv.T)).cpu().numpy()
    try:
        np.linalg.cholesky(masked_cov)
    except np.linalg.LinAlgError:
        print('WARNING: Reduced matrix is not positive definite')
    return np.linalg.inv(masked_cov)


def compute_log_cov_det(cov_mat, data_mask):
    """log det of the masked covariance (vega/utils.py:302-318)."""
    masked_cov = cov_mat[:, data_mask][data_mask, :]
    factor = _device_cholesky(masked_cov)
    if factor is not None:
        import torch
        return float(2.0 * torch.log(torch.diagonal(factor)).sum().cpu())
    return np.linalg.slogdet(masked_cov)[1]


def tracer_bias_beta(params, name):
    """(bias, beta) of one tracer from any two of (bias, bias_eta, beta)
    (vega/utils.py:45-81).  Works on floats and on symbolic registers."""
    growth_rate = params.get('growth_rate', DEFAULT_GROWTH_RATE)
    bias = params.get('bias_' + name, None)
    bias_eta = params.get('bias_eta_' + name, None)
    beta = params.get('beta_' + name, None)
    err_msg = ("For each tracer, you need to specify two of these three:"
               " (bias, bias_eta, beta). If all three are given, we use bias and beta. "
               f"Offending tracer: {name}")
    if bias is None:
        assert bias_eta is not None and beta is not None, err_msg
        bias = bias_eta * growth_rate / beta
    if bias_eta is None:
        assert bias is not None and beta is not None, err_msg
    if beta is None:
        assert bias is not None and bias_eta is not None, err_msg
        beta = bias_eta * growth_rate / bias
    return bias, beta


def bias_beta(params, tracer1_name, tracer2_name):
    """vega/utils.py:84-108."""
    bias1, beta1 = tracer_bias_beta(params, tracer1_name)
    if tracer1_name == tracer2_name:
        return bias1, beta1, bias1, beta1
    bias2, beta2 = tracer_bias_beta(params, tracer2_name)
    return bias1, beta1, bias2, beta2


def uv_shotnoise_table(ntau=100, nrho=10000):
    """A(tau) of the UVB shot-noise term, tabulated once at init
    (vega/correlation_func.py:597-626; Gontcho A Gontcho et al. 2014 eq. 19)."""
    from scipy.special import expn
    tau = np.linspace(0.01, 5, ntau)
    rho = np.linspace(0.0001, 10, nrho)
    drho = rho[1] - rho[0]
    amp = np.zeros(ntau)
    for i, t in enumerate(tau):
        x = rho * np.sqrt(1 + (t / rho)**2)
        y = rho * np.abs(1 - t / rho)
        amp[i] = -np.sum(drho * np.exp(-rho) / rho * (expn(1, x) - expn(1, y)))
    return tau, amp


def _on_device(n_rows):
    if n_rows < DEVICE_LINALG_MIN:
        return False
    import torch
    return torch.cuda.is_available()


def template_covariance(t, factor=1e-8):
    """sum over the non-degenerate modes of the marginalisation templates t [n_mask, n_templates]:
    SVD, modes with singular value <= factor * largest dropped, U diag(s^2) U^T
    (vega/data.py:813-828).  Returns (covariance update [n_mask, n_mask], number of modes kept).
    From DEVICE_LINALG_MIN rows on, SVD and product run on the GPU (cuSOLVER / cuBLAS through torch,
    init time); below, numpy's LAPACK as the reference."""
    if _on_device(t.shape[0]):
        import torch
        dev = torch.from_numpy(np.ascontiguousarray(t, dtype=np.float64)).cuda()
        u, sv, _ = torch.linalg.svd(dev, full_matrices=False)
        keep = sv > factor * sv[0]
        u, sv = u[:, keep], sv[keep]
        return torch.matmul(u * sv**2, u.T).cpu().numpy(), int(keep.sum())
    u, sv, _ = np.linalg.svd(t, full_matrices=False)
    keep = sv > factor * sv[0]
    u, sv = u[:, keep], sv[keep]
    return np.dot(u * sv**2, u.T), int(keep.sum())


def diff_to_coeff_matrix(t_masked, inv_cov, prior_diag=None):
    """(T^T C^-1 T + S)^-1 T^T C^-1: the matrix that turns a residual into the best-fit template
    amplitudes (vega/data.py:113-128).  t_masked [n_mask, n_templates] (dense or scipy sparse),
    inv_cov [n_mask, n_mask], prior_diag [n_templates] or None.  Device route as above."""
    dense = np.asarray(t_masked.toarray() if hasattr(t_masked, 'toarray') else t_masked, dtype=np.float64)
    if _on_device(dense.shape[0]):
        import torch
        t = torch.from_numpy(np.ascontiguousarray(dense)).cuda()
        c = torch.from_numpy(np.ascontiguousarray(inv_cov, dtype=np.float64)).cuda()
        g = torch.matmul(t.T, c)
        a = torch.matmul(g, t)
        if prior_diag is not None:
            a = a + torch.diag(torch.from_numpy(np.asarray(prior_diag, dtype=np.float64)).cuda())
        return torch.linalg.solve(a, g).cpu().numpy()
    g = np.asarray(t_masked.T @ inv_cov)
    a = np.asarray((t_masked.T @ g.T).T)
    if prior_diag is not None:
        a = a + np.diag(prior_diag)
    return np.linalg.inv(a).dot(g)
