"""The approximation behind the velocity-dispersion tables (csrc/vg_moments.cuh), checked on the CPU
oracle: xi as a function of the sampled velocity dispersion sigma_v is a Chebyshev series in
tau = log(sigma_v^2 + eps^2) / 2 over the parameter's [sample] limits, converging geometrically --
at the 64 nodes the device builds its tables from, the series reproduces the oracle at other values
of sigma_v far below the 1e-10 max|xi| parity bar.  (The device path re-checks this at finalize, on
its own tables, and picks the order count there.)"""
import numpy as np

from helpers import make_oracle
from oracle.vega_oracle import pk_model

SIGMA = 'sigma_velo_disp_lorentz_QSO'
EPS = 0.05          # VT_EPS


def test_chebyshev_series_in_sigma_converges_on_the_oracle():
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    name = next(n for n, it in vega.corr_items.items() if it.tracer1['name'] != it.tracer2['name'])
    cm = oracle.models[name]
    pars0 = oracle.local_params()
    assert SIGMA in pars0
    lo, hi = 0.0, 15.0                     # default_values.txt limits of the parameter
    pk_lin = vega.fiducial['pk_smooth']

    def xi_at(sigma):
        p = dict(pars0)
        p[SIGMA] = sigma
        p['peak'] = False
        return cm.xi_core.compute(pk_model(cm.grid, pk_lin, p), pk_lin, cm.transform, p)

    tau = lambda s: 0.5 * np.log(s * s + EPS * EPS)
    t_lo, t_hi = tau(lo), tau(hi)
    n = 64
    q = np.arange(n)
    nodes = np.cos(np.pi * (q + 0.5) / n)
    sig = np.sqrt(np.maximum(np.exp(t_lo + t_hi + (t_hi - t_lo) * nodes) - EPS * EPS, 0.0))
    values = np.array([xi_at(s) for s in sig])
    D = (2.0 / n) * np.cos(np.pi * np.outer(q, q + 0.5) / n)
    D[0] *= 0.5
    coef = D @ values
    scale = np.abs(values).max()
    size = np.abs(coef).max(axis=1) / scale
    # geometric decay: every 8 orders buy at least a factor 10 until the rounding floor
    assert size[8] < 1e-1 * size[0] and size[24] < 1e-3 * size[8] and size[56:].max() < 1e-11
    tests = [0.0, 0.013, 0.4, 2.2, 7.7724, 11.9, 14.99, 15.0]
    for order, bar in ((32, 1e-6), (48, 1e-9), (64, 5e-12)):
        worst = 0.0
        for s in tests:
            t = (2 * tau(s) - t_lo - t_hi) / (t_hi - t_lo)
            T = np.cos(np.arange(order) * np.arccos(np.clip(t, -1.0, 1.0)))
            worst = max(worst, np.abs(T @ coef[:order] - xi_at(s)).max() / scale)
        assert worst < bar, (order, worst)
