"""Batched drivers of the B200 likelihood path.

The reference drives VegaInterface.chi2 / log_lik one parameter dictionary at a time:
iminuit through Minimizer.chi2(*pars) (vega/minimizer.py:28-103), PolyChord through
log_lik(theta) -> (log_lik, derived) with a uniform prior transform
(vega/samplers/polychord.py:106-127), and the Monte Carlo scripts through one fit per mock
(vega/analysis.py:224-308, bin/run_vega_mc_fits_mpi.py:81-159).  The device path is
batched, so the drivers here produce BATCHES:

  stencil(theta0, steps)        central-difference stencil of one point: 2P+1 vectors for the
                                gradient, 2P^2+1 with the mixed terms for the Hessian;
  BatchedMinimizer              same constructor and read-out surface as the reference's
                                Minimizer (values / errors / covariance / fmin), minimising
                                with a damped Newton iteration whose gradient AND Hessian
                                come from ONE batched chi2 call per iteration;
  BatchedSamplerLikelihood      PolyChord-style adapter: log_lik(theta) -> (log_lik, []),
                                prior(hypercube), plus log_lik_batch / prior_batch over a
                                set of live points, optionally sharded over ranks;
  MonteCarloFits                many mock fits advanced in lock-step: every Newton
                                iteration is one batch of (mocks x stencil) vectors.

None of these evaluate the model on the host: every chi2 comes from the device engine.
"""
import copy
import time

import numpy as np

from . import utils


# ---------------------------------------------------------------------------------------------
# stencils
# ---------------------------------------------------------------------------------------------
def stencil(theta0, steps, hessian=False):
    """Central-difference stencil around theta0[P] with step sizes steps[P].

    Row 0 is theta0; rows 1+2j, 2+2j are theta0 +- steps[j] e_j.  With hessian=True the
    rows for every pair j<k follow, 4 per pair: (++, +-, -+, --).  Returns [n, P]."""
    theta0 = np.asarray(theta0, dtype=np.float64)
    steps = np.asarray(steps, dtype=np.float64)
    P = theta0.size
    rows = [theta0.copy()]
    for j in range(P):
        for sign in (1.0, -1.0):
            t = theta0.copy()
            t[j] += sign * steps[j]
            rows.append(t)
    if hessian:
        for j in range(P):
            for k in range(j + 1, P):
                for sj, sk in ((1., 1.), (1., -1.), (-1., 1.), (-1., -1.)):
                    t = theta0.copy()
                    t[j] += sj * steps[j]
                    t[k] += sk * steps[k]
                    rows.append(t)
    return np.array(rows)


def stencil_size(P, hessian=False):
    return 1 + 2 * P + (2 * P * (P - 1) if hessian else 0)


def derivatives_from_stencil(f, steps, hessian=False):
    """(f0, grad[P], hess[P,P] or None) from the values f[n] on a `stencil` batch."""
    f = np.asarray(f, dtype=np.float64)
    steps = np.asarray(steps, dtype=np.float64)
    P = steps.size
    f0 = f[0]
    fp, fm = f[1:1 + 2 * P:2], f[2:2 + 2 * P:2]
    grad = (fp - fm) / (2 * steps)
    if not hessian:
        return f0, grad, None
    hess = np.zeros((P, P))
    hess[np.diag_indices(P)] = (fp - 2 * f0 + fm) / steps**2
    idx = 1 + 2 * P
    for j in range(P):
        for k in range(j + 1, P):
            fpp, fpm, fmp, fmm = f[idx:idx + 4]
            hess[j, k] = hess[k, j] = (fpp - fpm - fmp + fmm) / (4 * steps[j] * steps[k])
            idx += 4
    return f0, grad, hess


# ---------------------------------------------------------------------------------------------
# damped Newton on batched stencils, M independent problems in lock-step
# ---------------------------------------------------------------------------------------------
def newton_lockstep(eval_batch, theta0, steps, lower=None, upper=None, free=None, max_iter=50,
                    edm_tol=1e-5, step_frac=0.02, verbose=False):
    """Minimise M independent functions that share one batched evaluator.

    eval_batch(theta[n, P], problem[n]) -> f[n]: value of problem problem[i] at theta[i].
    theta0 [M, P] start points, steps [P] expected 1-sigma scales (the stencil uses
    step_frac * steps, re-estimated from the Hessian once it is known), lower/upper [P]
    bounds (nan = none), free [P] bool.  Every iteration is ONE eval_batch call of
    M_active * (2F^2 + 1) vectors (F free parameters): value, gradient and Hessian of every
    active problem, followed by a Levenberg-Marquardt step per problem.
    Returns dict(theta[M,P], fval[M], hess[M,F,F], edm[M], n_iter[M], converged[M], n_eval)."""
    theta = np.array(theta0, dtype=np.float64, ndmin=2)
    M, P = theta.shape
    free = np.ones(P, bool) if free is None else np.asarray(free, bool)
    fidx = np.nonzero(free)[0]
    F = fidx.size
    lower = np.full(P, np.nan) if lower is None else np.asarray(lower, float)
    upper = np.full(P, np.nan) if upper is None else np.asarray(upper, float)
    scale = np.tile(np.asarray(steps, float)[fidx], (M, 1))     # per problem 1-sigma estimate
    lam = np.full(M, 1e-3)
    fval = np.full(M, np.inf)
    best = {'theta': theta.copy(), 'g': np.zeros((M, F)), 'H': np.tile(np.eye(F), (M, 1, 1))}
    have_best = np.zeros(M, bool)
    edm = np.full(M, np.inf)
    n_iter = np.zeros(M, int)
    done = np.zeros(M, bool)
    n_eval = 0
    n_st = stencil_size(F, hessian=True)

    def clip(t, h):
        lo = np.where(np.isnan(lower[fidx]), -np.inf, lower[fidx] + h)
        hi = np.where(np.isnan(upper[fidx]), np.inf, upper[fidx] - h)
        return np.minimum(np.maximum(t, lo), hi)

    for it in range(max_iter):
        act = np.nonzero(~done)[0]
        if act.size == 0:
            break
        h = step_frac * scale[act]                                   # [A, F]
        centre = np.array([clip(theta[m, fidx], h[i]) for i, m in enumerate(act)])
        batch = np.repeat(theta[act], n_st, axis=0)
        for i, m in enumerate(act):
            batch[i * n_st:(i + 1) * n_st][:, fidx] = stencil(centre[i], h[i], hessian=True)
        f = np.asarray(eval_batch(batch, np.repeat(act, n_st)), dtype=np.float64)
        n_eval += batch.shape[0]
        for i, m in enumerate(act):
            n_iter[m] += 1
            f0, g, H = derivatives_from_stencil(f[i * n_st:(i + 1) * n_st], h[i], hessian=True)
            if not np.isfinite(f0) or f0 >= 1e99 or (have_best[m] and f0 > fval[m]):
                lam[m] = min(lam[m] * 10., 1e8)                      # reject: back to the best point
                if lam[m] >= 1e8:
                    done[m] = True
            else:
                fval[m], have_best[m] = f0, True
                best['theta'][m, fidx] = centre[i]
                best['g'][m], best['H'][m] = g, H
                lam[m] = max(lam[m] / 10., 1e-9)
                d = np.diag(H)
                # re-estimate the 1-sigma scale (and with it the stencil step) from the curvature,
                # except for parameters resting on a bound: their centre is bound + step, so a
                # changing step would move the point under comparison
                near = (~np.isnan(lower[fidx]) & (centre[i] <= lower[fidx] + 2 * h[i])) | \
                       (~np.isnan(upper[fidx]) & (centre[i] >= upper[fidx] - 2 * h[i]))
                ok = (d > 0) & ~near
                scale[m, ok] = np.sqrt(2. / d[ok])
            if not have_best[m]:
                theta[m, fidx] = centre[i] * (1 + 1e-3)              # nudge off a failing start
                continue
            g, H = best['g'][m], best['H'][m]
            # active set: parameters sitting on a bound whose gradient pushes outwards take no
            # part in this step (MIGRAD's limits act through a variable transform instead)
            x = best['theta'][m, fidx]
            hm = step_frac * scale[m]
            at_lo = ~np.isnan(lower[fidx]) & (x <= lower[fidx] + 1.5 * hm) & (g > 0)
            at_hi = ~np.isnan(upper[fidx]) & (x >= upper[fidx] - 1.5 * hm) & (g < 0)
            mv = np.nonzero(~(at_lo | at_hi))[0]
            step = np.zeros(F)
            if mv.size:
                gm, Hm = g[mv], H[np.ix_(mv, mv)]
                dscale = np.sqrt(np.abs(np.diag(Hm))) + 1e-300
                Hs = Hm / dscale[:, None] / dscale[None, :]
                try:
                    sm = -np.linalg.solve(Hs + lam[m] * np.eye(mv.size), gm / dscale) / dscale
                    edm[m] = 0.5 * abs(gm.dot(-sm)) if lam[m] < 1e-2 else np.inf
                    pos_def = bool(np.all(np.linalg.eigvalsh(Hs) > 0))
                except np.linalg.LinAlgError:
                    sm, pos_def = -gm / (dscale**2 * (1 + lam[m])), False
                step[mv] = sm
            else:
                edm[m], pos_def = 0.0, True
            if edm[m] < edm_tol and pos_def:
                done[m] = True
                theta[m] = best['theta'][m]
                continue
            theta[m, fidx] = best['theta'][m, fidx] + step
        if verbose:
            print(f'[newton] iteration {it + 1}: {act.size} active, fval[0] = {fval[0]:.6f}, '
                  f'edm[0] = {edm[0]:.3e}')
    theta[:] = best['theta']
    return {'theta': theta, 'fval': fval, 'hess': best['H'], 'edm': edm, 'n_iter': n_iter,
            'converged': done & np.isfinite(edm) & (edm < edm_tol), 'n_eval': n_eval, 'free': fidx}


# ---------------------------------------------------------------------------------------------
# minimizer with the reference's read-out surface
# ---------------------------------------------------------------------------------------------
class FMin:
    """The few fields of iminuit's fmin object the reference reads (vega/output.py,
    vega/analysis.py:100-118)."""

    def __init__(self, fval, edm, nfcn, is_valid, has_accurate_covar):
        self.fval, self.edm, self.nfcn = float(fval), float(edm), int(nfcn)
        self.is_valid, self.has_accurate_covar = bool(is_valid), bool(has_accurate_covar)
        self.has_covariance = True

    def __repr__(self):
        return (f'FMin(fval={self.fval:.6f}, edm={self.edm:.3e}, nfcn={self.nfcn}, '
                f'is_valid={self.is_valid})')


class BatchedMinimizer:
    """Drop-in for vega.minimizer.Minimizer (vega/minimizer.py:8-190) on the batched path.

    Same constructor shape -- a chi2 function and the [sample] dictionary -- except that the
    function is BATCHED: chi2_batch(theta[B, P]) -> chi2[B] with columns in the order of
    sample_params['limits'].  minimize(params) follows the reference's two stages (a first
    fit over the bias parameters only, vega/minimizer.py:63-85, then all free parameters)
    and honours values / errors / limits / fix overrides.  The optimiser is a damped Newton
    iteration on one batched stencil per step (value + gradient + Hessian in a single device
    call) instead of MIGRAD's sequential line searches; errors are HESSE-like, from the
    inverse Hessian at the minimum with errordef = 1."""

    def __init__(self, chi2_batch, sample_params, max_iter=50, edm_tol=1e-5, verbose=True):
        self.chi2_batch = chi2_batch
        self._names = list(sample_params['limits'].keys())
        self._sample_params = sample_params
        self.max_iter, self.edm_tol, self.verbose = max_iter, edm_tol, verbose
        self._run_flag = False
        self._result = None
        self.n_eval = 0

    @classmethod
    def from_vega(cls, vega, **kw):
        names = list(vega.sample_params['limits'].keys())
        return cls(lambda theta: vega.chi2_batch(theta, names), vega.sample_params, **kw)

    def chi2(self, *pars):
        """Scalar wrapper with iminuit's calling convention (vega/minimizer.py:28-37)."""
        return float(self.chi2_batch(np.array([pars], dtype=np.float64))[0])

    def _settings(self, params):
        out = {k: copy.deepcopy(self._sample_params[k]) for k in ('values', 'errors', 'limits', 'fix')}
        if params is not None:
            for key in out:
                if key in params:
                    out[key].update(params[key])
        return out

    def _fit(self, start, errors, limits, fixed):
        lower = np.array([np.nan if limits[n][0] is None else limits[n][0] for n in self._names])
        upper = np.array([np.nan if limits[n][1] is None else limits[n][1] for n in self._names])
        res = newton_lockstep(lambda theta, prob: self.chi2_batch(theta), start[None, :],
                              np.array([errors[n] for n in self._names]), lower, upper,
                              free=~fixed, max_iter=self.max_iter, edm_tol=self.edm_tol,
                              verbose=self.verbose)
        self.n_eval += res['n_eval']
        return res

    def minimize(self, params=None):
        t0 = time.time()
        s = self._settings(params)
        start = np.array([s['values'][n] for n in self._names], dtype=np.float64)
        fixed = np.array([bool(s['fix'][n]) for n in self._names])
        self.n_eval = 0
        # initial "fast" minimisation over the biases (vega/minimizer.py:63-85)
        bias = np.array(['bias' in n for n in self._names])
        if bias.any() and (bias & ~fixed).any() and (~bias & ~fixed).any():
            first = self._fit(start, s['errors'], s['limits'], fixed | ~bias)
            if np.isfinite(first['fval'][0]):
                start = first['theta'][0]
        res = self._fit(start, s['errors'], s['limits'], fixed)
        fidx = res['free']
        cov = np.zeros((len(self._names), len(self._names)))
        accurate = True
        try:
            cov_free = 2.0 * np.linalg.inv(res['hess'][0])          # errordef = 1 for a chi2
            accurate = bool(np.all(np.linalg.eigvalsh(res['hess'][0]) > 0))
        except np.linalg.LinAlgError:
            cov_free, accurate = np.full((fidx.size, fidx.size), np.nan), False
        cov[np.ix_(fidx, fidx)] = cov_free
        self._result = {'theta': res['theta'][0], 'cov': cov, 'fixed': fixed}
        self._fmin = FMin(res['fval'][0], res['edm'][0], self.n_eval, res['converged'][0], accurate)
        self._run_flag = True
        if self.verbose:
            print(self._fmin)
            print('INFO: minimized in {:.2f} s with {} chi2 evaluations'.format(time.time() - t0, self.n_eval))

    # -- read-out (vega/minimizer.py:105-190) ------------------------------------------------
    def _need_run(self):
        if not self._run_flag:
            raise RuntimeError('Run minimize() before reading results.')

    @property
    def values(self):
        self._need_run()
        return dict(zip(self._names, self._result['theta'].tolist()))

    @property
    def errors(self):
        self._need_run()
        return dict(zip(self._names, np.sqrt(np.maximum(np.diag(self._result['cov']), 0.)).tolist()))

    @property
    def covariance(self):
        self._need_run()
        return self._result['cov']

    @property
    def fmin(self):
        self._need_run()
        return self._fmin

    @property
    def fixed(self):
        self._need_run()
        return dict(zip(self._names, self._result['fixed'].tolist()))


# ---------------------------------------------------------------------------------------------
# sampler adapter
# ---------------------------------------------------------------------------------------------
class BatchedSamplerLikelihood:
    """What vega/samplers/polychord.py:106-127 hands to PolyChord -- log_lik(theta) ->
    (log_lik, derived) and prior(hypercube) -- plus the batched forms a sampler that proposes
    whole sets of points (live points, walkers, PocoMC particles: bin/run_vega_mpi.py:94-121)
    can call: log_lik_batch(theta[B, P]) and prior_batch(cube[B, P]).

    With `sharded` (a vega_b200.sharding.ShardedChi2) the batch is split over the ranks of
    the process group, one GPU each, and the chi2 values are all-gathered."""

    def __init__(self, vega, sharded=None):
        self.vega = vega
        self.names = list(vega.sample_params['limits'].keys())
        self.limits = {n: vega.sample_params['limits'][n] for n in self.names}
        for n, (lo, hi) in self.limits.items():
            if lo is None or hi is None:
                raise ValueError(f'Sampler needs finite limits for {n}.')
        self.num_params = len(self.names)
        self.num_derived = 0
        self.sharded = sharded
        self._lo = np.array([self.limits[n][0] for n in self.names], dtype=np.float64)
        self._hi = np.array([self.limits[n][1] for n in self.names], dtype=np.float64)

    def prior(self, hypercube):
        """Uniform prior transform of one point of the unit hypercube."""
        return list(self._lo + (self._hi - self._lo) * np.asarray(hypercube, dtype=np.float64))

    def prior_batch(self, cube):
        return self._lo + (self._hi - self._lo) * np.asarray(cube, dtype=np.float64)

    def chi2_batch(self, theta):
        theta = np.atleast_2d(np.asarray(theta, dtype=np.float64))
        if self.sharded is not None:
            return self.sharded.chi2(theta)
        return self.vega.chi2_batch(theta, self.names)

    def log_lik_batch(self, theta):
        """log-likelihood of every row (vega/vega_interface.py:327-387); rows whose model
        failed carry chi2 = 1e100 exactly as the scalar path does."""
        return self.vega._log_norm() - 0.5 * self.chi2_batch(theta)

    def log_lik(self, theta):
        return float(self.log_lik_batch(np.asarray(theta, dtype=np.float64)[None, :])[0]), []


# ---------------------------------------------------------------------------------------------
# Monte Carlo mock fits in lock-step
# ---------------------------------------------------------------------------------------------
class MonteCarloFits:
    """Fits of many Gaussian mocks advanced together (vega/analysis.py:224-308 runs
    create_monte_carlo_sim + minimize once per mock; bin/run_vega_mc_fits_mpi.py:104-113
    gives every MPI rank a strided share of the mocks).

    Mocks are drawn exactly as the reference draws them -- Data.create_monte_carlo with
    np.random.seed(seed + m) per mock -- uploaded once, and every Newton iteration evaluates
    (active mocks x stencil) vectors in one vg_chi2_batch_mc call.  `mock_slice` selects this
    rank's share when the mocks are sharded over GPUs."""

    def __init__(self, vega, fiducial_model=None, scale=None):
        import torch
        self.torch = torch
        self.vega = vega
        # the [monte carlo] sample set when the configuration has one (vega/analysis.py:246)
        self.sample_params = vega.mc_config['sample'] if getattr(vega, 'mc_config', None) \
            else vega.sample_params
        self.names = list(self.sample_params['limits'].keys())
        self.fiducial_model = fiducial_model if fiducial_model is not None else vega.compute_model()
        self.scale = scale
        self.mocks = None

    def _item_scale(self, name):
        """Covariance scale of one correlation as create_monte_carlo_sim resolves it
        (vega/analysis.py:146-156): None -> the item's cov_rescale, number, or dict by name."""
        if self.scale is None:
            return getattr(self.vega.corr_items[name], 'cov_rescale', None)
        if isinstance(self.scale, dict):
            return self.scale.get(name, 1.)
        return self.scale

    def draw(self, n_mocks, seed=0, forecast=False):
        """masked mocks {name: [n_mocks, n_mask]}, drawn as run_monte_carlo draws them
        (vega/analysis.py:245,261-263): np.random.seed(seed) once, then one
        Data.create_monte_carlo per correlation per mock from the running stream
        (forecast: the fiducial itself, no noise)."""
        vega = self.vega
        mocks = {name: [] for name in vega.corr_items}
        self.full_mocks = {name: [] for name in vega.corr_items}
        if seed is not None:
            np.random.seed(seed)
        if vega._use_global_cov:
            self.mocks = self._draw_global(n_mocks, forecast)
        else:
            for m in range(n_mocks):
                for name in vega.corr_items:
                    full = vega.data[name].create_monte_carlo(self.fiducial_model[name], self._item_scale(name),
                                                              None, forecast)
                    self.full_mocks[name].append(np.array(full))
                    mocks[name].append(vega.data[name].masked_mc_mock.copy())
            self.mocks = {name: np.array(v) for name, v in mocks.items()}
        vega.monte_carlo = True
        engine = vega._engine_for(tuple(self.names))   # swaps in the scaled inverse covariance (MC mode)
        for ci, name in enumerate(vega.corr_items):
            engine.set_mocks(ci, self.mocks[name])
        self.engine = engine
        return self.mocks

    def _draw_global(self, n_mocks, forecast=False):
        """Mocks of the joint data vector under a global covariance
        (Analysis.create_global_monte_carlo, vega/analysis.py:164-222): one Cholesky factor of
        scale * cov[mask][:, mask], mock = masked fiducial + L . randn(n) from the running
        stream, one draw per mock.  The chi2 keeps the UNSCALED masked global inverse
        (vega/vega_interface.py:295-303), as the reference does.  Returned split per correlation
        in the order the residuals are concatenated."""
        vega = self.vega
        names = list(vega.corr_items)
        full_mask = np.concatenate([vega.data[n].data_mask for n in names])
        masked_cov = vega.global_cov[:, full_mask][full_mask, :]
        scale = 1 if self.scale is None else self.scale
        if isinstance(scale, dict):
            raise ValueError('A global covariance takes one scale, not one per correlation.')
        chol = None
        if masked_cov.shape[0] >= utils.DEVICE_LINALG_MIN and self.torch.cuda.is_available():
            # DESI-sized joint covariance (9540 rows): factor on the device (init time)
            dev = self.torch.from_numpy(np.ascontiguousarray(scale * masked_cov)).cuda()
            factor, info = self.torch.linalg.cholesky_ex(dev)
            chol = factor.cpu().numpy() if int(info) == 0 else None
        if chol is None:
            chol = np.linalg.cholesky(scale * masked_cov)
        fid = []
        for n in names:
            data, model = vega.data[n], np.asarray(self.fiducial_model[n])
            to_data = data.dist_model_coordinates.get_mask_to_other(data.data_coordinates)
            if data.data_mask.size == model.size:
                fid.append(model)
            elif to_data.size == model.size:
                fid.append(model[to_data])
            else:
                raise ValueError('Input fiducial has unknown size. It must match the data or the model.')
        fid = np.concatenate(fid)[full_mask]
        if forecast:                        # vega/analysis.py:214-215: the masked fiducial itself
            joint = np.tile(fid, (n_mocks, 1))
        else:
            z = np.stack([np.random.randn(fid.size) for _ in range(n_mocks)])
            joint = fid[None, :] + z @ chol.T
        self.global_mocks = joint
        edges = np.cumsum([0] + [int(vega.data[n].data_mask.sum()) for n in names])
        mocks = {n: np.ascontiguousarray(joint[:, edges[i]:edges[i + 1]]) for i, n in enumerate(names)}
        vega.analysis.current_mc_mock = joint[-1].copy()       # vega/analysis.py:217-220
        for n in names:      # what the scalar MC-mode chi2 reads (the last mock drawn, as in the reference)
            data = vega.data[n]
            data.masked_mc_mock = mocks[n][-1].copy()
            data.mc_version = getattr(data, 'mc_version', 0) + 1
            if data.scaled_inv_masked_cov is None:
                data.scaled_inv_masked_cov = data.inv_masked_cov
        return mocks

    def chi2(self, theta, mock_idx):
        """chi2 of theta[b] against mock mock_idx[b]; host arrays in and out."""
        torch, engine = self.torch, self.engine
        tt = engine.theta_tensor(self.vega._reorder(engine, theta, self.names))
        mi = torch.as_tensor(np.asarray(mock_idx, dtype=np.int32), device=engine.device)
        chi2, status = engine.chi2_mc_device(tt, mi)
        chi2 = chi2.cpu().numpy().copy()
        chi2[status.cpu().numpy() != 0] = 1e100
        return chi2

    def result_shape(self, key):
        """Trailing shape of one row of fit()[key] (what an empty shard contributes to the
        gather of sharding.fit_mocks_sharded)."""
        return (len(self.names),) if key in ('theta', 'errors') else ()

    def fit_sharded(self, **fit_kwargs):
        """fit() of this rank's contiguous share of the mocks, best fits gathered on every rank
        (one process per GPU; the C5 workload)."""
        from .sharding import fit_mocks_sharded
        n_mocks = next(iter(self.mocks.values())).shape[0]
        return fit_mocks_sharded(self, n_mocks, self.engine.device, **fit_kwargs)

    def fit(self, mock_slice=None, params=None, max_iter=50, edm_tol=1e-5, verbose=False):
        """Best fit of every mock in mock_slice (default: all).  Returns the dict of
        newton_lockstep plus 'names' and 'errors'[M, P]."""
        if self.mocks is None:
            raise RuntimeError('draw() the mocks first.')
        n_mocks = next(iter(self.mocks.values())).shape[0]
        ids = np.arange(n_mocks)[mock_slice] if mock_slice is not None else np.arange(n_mocks)
        sp = self.sample_params
        if params is not None:
            sp = {k: {**sp[k], **params.get(k, {})} for k in ('values', 'errors', 'limits', 'fix')}
        start = np.array([sp['values'][n] for n in self.names], dtype=np.float64)
        errors = np.array([sp['errors'][n] for n in self.names], dtype=np.float64)
        lower = np.array([np.nan if sp['limits'][n][0] is None else sp['limits'][n][0] for n in self.names])
        upper = np.array([np.nan if sp['limits'][n][1] is None else sp['limits'][n][1] for n in self.names])
        fixed = np.array([bool(sp['fix'][n]) for n in self.names])
        res = newton_lockstep(lambda theta, prob: self.chi2(theta, ids[prob]),
                              np.tile(start, (ids.size, 1)), errors, lower, upper, free=~fixed,
                              max_iter=max_iter, edm_tol=edm_tol, verbose=verbose)
        errs = np.zeros((ids.size, len(self.names)))
        for m in range(ids.size):
            try:
                errs[m, res['free']] = np.sqrt(np.maximum(np.diag(2.0 * np.linalg.inv(res['hess'][m])), 0.))
            except np.linalg.LinAlgError:
                errs[m, res['free']] = np.nan
        res.update(names=self.names, errors=errs, mock_ids=ids)
        return res


class BatchedAnalysis:
    """`Analysis.chi2_scan` (vega/analysis.py:53-124) with every grid point fitted at once.

    The reference minimises once per grid point of the one or two scanned parameters, those
    parameters fixed at the grid value.  The fits are independent, so here they are the M
    problems of one newton_lockstep call: each Newton iteration is one batched chi2 call of
    (active grid points) x (2F^2 + 1) vectors.  Returns the reference's list of dictionaries
    (best-fit values of every sampled parameter + 'fval'), in the reference's loop order."""

    def __init__(self, chi2_batch, sample_params, main_config, max_iter=50, edm_tol=1e-5, verbose=True):
        self._chi2_batch = chi2_batch
        self._sample_params = sample_params
        self.config = main_config
        self._names = list(sample_params['limits'].keys())
        self.max_iter, self.edm_tol, self.verbose = max_iter, edm_tol, verbose
        self.grids = {}
        self.scan_results = None
        self._vega = None
        self.has_monte_carlo = False
        self.current_mc_mock = None

    @classmethod
    def from_vega(cls, vega, **kw):
        names = list(vega.sample_params['limits'].keys())
        self = cls(lambda theta: vega.chi2_batch(theta, names), vega.sample_params, vega.main_config, **kw)
        self._vega = vega
        return self

    def run_monte_carlo(self, fiducial_model, num_mocks=1, seed=0, scale=None, forecast=False,
                        run_mc_fits=True):
        """`Analysis.run_monte_carlo` (vega/analysis.py:224-308; the loop of
        bin/run_vega_mc_mpi.py:62 and bin/run_vega_mc_fits_mpi.py:11-78): `num_mocks` Gaussian mocks
        around `fiducial_model` from np.random.seed(seed), each fitted over the [monte carlo]
        sample set.  The reference draws and fits one mock at a time; here all mocks are drawn
        from the same random stream in the same order, uploaded once and fitted in lock-step
        (MonteCarloFits).  Fills the reference's result attributes: mc_bestfits[param] ->
        [num_mocks, 2] (value, error), mc_covariances, mc_chisq, mc_valid_minima,
        mc_valid_hesse, mc_mocks, mc_failed_mask."""
        vega = self._vega
        if vega is None:
            raise RuntimeError('run_monte_carlo needs the interface: build with BatchedAnalysis.from_vega.')
        assert vega.mc_config is not None, 'No Monte Carlo config provided'
        mc = MonteCarloFits(vega, fiducial_model=fiducial_model, scale=scale)
        mc.draw(num_mocks, seed=seed, forecast=forecast)
        if vega._use_global_cov:
            self.mc_mocks = {'global': [m for m in mc.global_mocks]}
        else:
            self.mc_mocks = {name: list(v) for name, v in mc.full_mocks.items()}
        self.mc_bestfits, self.mc_covariances, self.mc_chisq = {}, [], []
        self.mc_valid_minima, self.mc_valid_hesse, self.mc_failed_mask = [], [], []
        self.mc_fits = None
        if run_mc_fits:
            res = mc.fit(max_iter=self.max_iter, edm_tol=self.edm_tol)
            self.mc_fits = res
            free = res['free']
            for j, name in enumerate(mc.names):
                self.mc_bestfits[name] = np.column_stack([res['theta'][:, j], res['errors'][:, j]])
            for m in range(num_mocks):
                cov = np.zeros((len(mc.names), len(mc.names)))
                ok = True
                try:
                    cov[np.ix_(free, free)] = 2.0 * np.linalg.inv(res['hess'][m])
                    ok = bool(np.all(np.linalg.eigvalsh(res['hess'][m]) > 0))
                except np.linalg.LinAlgError:
                    ok = False
                self.mc_covariances.append(cov)
                self.mc_chisq.append(float(res['fval'][m]))
                self.mc_valid_minima.append(bool(res['converged'][m]))
                self.mc_valid_hesse.append(ok)
                self.mc_failed_mask.append(False)
        self.has_monte_carlo = True
        return self.mc_bestfits

    def chi2_scan(self):
        if 'chi2 scan' not in self.config:
            raise ValueError('Called chi2_scan, but no config specified in'
                             ' main.ini. Add a "[chi2 scan]" section to main.')
        self.grids = {}
        for param, value in self.config.items('chi2 scan'):
            start, end, num = value.split()[:3]
            self.grids[param] = np.linspace(float(start), float(end), int(num))
        if len(self.grids) > 2:
            raise ValueError('chi2_scan only supports one/two parameter scans')
        for param in self.grids:
            if param not in self._names:
                raise ValueError(f'chi2 scan parameter {param} is not in [sample].')
        sp, names = self._sample_params, self._names
        start = np.array([sp['values'][n] for n in names], dtype=np.float64)
        errors = np.array([sp['errors'][n] for n in names], dtype=np.float64)
        lower = np.array([np.nan if sp['limits'][n][0] is None else sp['limits'][n][0] for n in names])
        upper = np.array([np.nan if sp['limits'][n][1] is None else sp['limits'][n][1] for n in names])
        free = np.array([not sp['fix'][n] and n not in self.grids for n in names])
        mesh = np.meshgrid(*self.grids.values(), indexing='ij')       # first parameter outermost
        points = np.stack([m.ravel() for m in mesh], axis=1)
        theta0 = np.tile(start, (points.shape[0], 1))
        for j, param in enumerate(self.grids):
            theta0[:, names.index(param)] = points[:, j]
        res = newton_lockstep(lambda theta, prob: self._chi2_batch(theta), theta0, errors, lower, upper,
                              free=free, max_iter=self.max_iter, edm_tol=self.edm_tol)
        self.scan_results = []
        for m in range(points.shape[0]):
            result = {n: float(res['theta'][m, j]) for j, n in enumerate(names)}
            result['fval'] = float(res['fval'][m])
            self.scan_results.append(result)
        self.converged = res['converged']
        if self.verbose:
            print('INFO: chi2 scan of {} points in {} batched chi2 evaluations, {} converged'.format(
                points.shape[0], res['n_eval'], int(np.sum(res['converged']))))
        return self.scan_results
