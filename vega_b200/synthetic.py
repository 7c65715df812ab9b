"""Seeded synthetic inputs for the benchmark workloads (BASELINE.json configs C2-C5).

The reference's fixtures carry no distortion matrix, no covariance and no metal matrices
(identity everywhere), so DESI/eBOSS-shaped dense operands are synthesised
here, IN MEMORY, on the Data objects of an already constructed VegaInterface.  The oracle
and the device plan both read those same objects, so they see identical inputs.

Recipes, all FP64, bins ordered rp-major as in the data files:
  distortion matrix  I - a (A_rp (x) A_rt), A row-stochastic exponential kernels -> dense,
                     well conditioned (vega/data.py:342-346 holds the real one as CSR);
  metal matrices     A_rp (x) A_rt of row-stochastic Gaussian bands, the rp band shifted per
                     metal correlation (the structure vega/metals.py:619-622 builds), dense;
  covariance         S (R_rp (x) I_rt) S with sigma^2 ~ 1/(r^2 + r0^2), R_rp = rho^|i-j|;
  data vector        model(fiducial) + chol(C_masked) z  (vega/data.py:726-756).
"""
import numpy as np


def _row_stochastic(m):
    return m / m.sum(axis=1, keepdims=True)


def _exp_kernel(n, scale):
    i = np.arange(n)
    return _row_stochastic(np.exp(-np.abs(i[:, None] - i[None, :]) / scale))


def _gauss_band(n, sigma, shift):
    i = np.arange(n)
    return _row_stochastic(np.exp(-0.5 * ((i[:, None] - i[None, :] - shift) / sigma)**2) + 1e-12)


def distortion_matrix(n_rp, n_rt, amp=0.3):
    return np.eye(n_rp * n_rt) - amp * np.kron(_exp_kernel(n_rp, 8.), _exp_kernel(n_rt, 2.))


def metal_matrix(n_rp, n_rt, shift, rng):
    sig_p = 1.0 + 0.5 * rng.random()
    sig_t = 0.6 + 0.3 * rng.random()
    return np.kron(_gauss_band(n_rp, sig_p, shift), _gauss_band(n_rt, sig_t, 0.))


def covariance(coords, rho=0.3, sigma0=3e-3, r0=20.):
    n_rp, n_rt = coords.rp_nbins, coords.rt_nbins
    sigma = sigma0 / np.sqrt(coords.r_regular_grid**2 + r0**2)
    i = np.arange(n_rp)
    corr_p = rho ** np.abs(i[:, None] - i[None, :])
    cov = np.kron(corr_p, np.eye(n_rt))
    cov *= sigma[:, None]
    cov *= sigma[None, :]
    return cov


def inject_dense_operands(vega, seed=20261104, with_distortion=True, with_metals=True,
                          with_covariance=True):
    """Replace the identity stand-ins of every correlation by dense synthetic operands."""
    rng = np.random.default_rng(seed)
    for name, item in vega.corr_items.items():
        data = vega.data[name]
        coords = item.model_coordinates
        n_rp, n_rt = coords.rp_nbins, coords.rt_nbins
        assert n_rp * n_rt == coords.r_grid.size
        if with_distortion:
            data._distortion_mat = distortion_matrix(n_rp, n_rt)
            data.distortion_is_identity = False
        if with_metals and item.has_metals:
            for c, pair in enumerate(item.metal_correlations):
                key = pair if pair in data.metal_mats else pair[::-1]
                shift = (-1)**c * (2 + (c % 5) * 3)          # metal lines displace r_parallel
                data.metal_mats[key] = metal_matrix(n_rp, n_rt, shift, rng)
        if with_covariance:
            data._cov_mat = covariance(item.data_coordinates)
            data.cov_is_identity = False
            data._inv_masked_cov = None
            data._log_cov_det = None
            data._cholesky = None
            data.variance = data._cov_mat.diagonal()
    vega._engine, vega._engines = None, {}
    return vega


def draw_mock_data(vega, model, seed=20261104):
    """data = model + chol(C_masked) z on the masked bins (vega/data.py:726-756);
    `model` is {name: xi[n_dist]} at the fiducial parameters."""
    for m, (name, item) in enumerate(vega.corr_items.items()):
        data = vega.data[name]
        rng = np.random.default_rng(seed + 1 + m)
        fid = np.asarray(model[name], dtype=float)
        if fid.size != data.full_data_size:
            fid = fid[data.dist_model_coordinates.get_mask_to_other(data.data_coordinates)]
        vec = fid.copy()
        masked_cov = data.cov_mat[:, data.data_mask][data.data_mask, :]
        chol = np.linalg.cholesky(masked_cov)
        vec[data.data_mask] = fid[data.data_mask] + chol.dot(rng.standard_normal(int(data.data_mask.sum())))
        data._data_vec = vec
        data._masked_data_vec = None     # (the device re-uploads the data vector when it changes)
    return vega


def draw_theta(vega, n, seed=20261104, width=3.0, names=None):
    """n parameter vectors uniform in fiducial +- width*sigma, clipped to the prior box
    (workload C2).  Columns follow the [sample] order."""
    rng = np.random.default_rng(seed)
    names = list(vega.sample_params['limits']) if names is None else list(names)
    theta = np.empty((n, len(names)))
    for j, p in enumerate(names):
        lo, hi = vega.sample_params['limits'][p]
        centre = vega.sample_params['values'][p]
        err = vega.sample_params['errors'][p]
        a = centre - width * err if lo is None else max(lo, centre - width * err)
        b = centre + width * err if hi is None else min(hi, centre + width * err)
        theta[:, j] = rng.uniform(a, b, size=n)
    return theta


def inject_global_covariance(vega, coupling=0.05):
    """Joint covariance over all correlations, in memory (what read_global_cov builds from a
    file, vega/vega_interface.py:888-954): the per-correlation covariances on the diagonal
    blocks plus a weak positive-definite coupling between the blocks, then the masked inverse
    and log-determinant exactly as the reference computes them."""
    from . import utils
    names = list(vega.corr_items)
    covs = [vega.data[n].cov_mat for n in names]
    sizes = [c.shape[0] for c in covs]
    total = sum(sizes)
    cov = np.zeros((total, total))
    sig = np.concatenate([np.sqrt(c.diagonal()) for c in covs])
    off = np.cumsum([0] + sizes)
    for i, c in enumerate(covs):
        cov[off[i]:off[i + 1], off[i]:off[i + 1]] = c
    # rank-one coupling u u^T with u ~ sigma: keeps the matrix SPD
    u = np.sqrt(coupling) * sig * np.cos(np.arange(total) * 0.37)
    cov += np.outer(u, u)
    vega.global_cov = cov
    vega._use_global_cov = True
    vega.full_data_mask = np.concatenate([vega.data[n].data_mask for n in names])
    vega.full_model_mask = np.concatenate([vega.data[n].model_mask for n in names])
    vega.masked_global_invcov = utils.compute_masked_invcov(cov, vega.full_data_mask)
    vega.masked_global_log_cov_det = utils.compute_log_cov_det(cov, vega.full_data_mask)
    vega._engine, vega._engines = None, {}
    return vega
