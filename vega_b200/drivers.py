"""Batched drivers of the B200 likelihood path.

The reference drives VegaInterface.chi2 / log_lik one parameter dictionary at a time:
iminuit through Minimizer.chi2(*pars) (vega/minimizer.py:28-103), PolyChord through
log_lik(theta) -> (log_lik, derived) with a uniform prior transform
(vega/samplers/polychord.py:106-127), and the Monte Carlo scripts through one fit per mock
(vega/analysis.py:224-308, bin/run_vega_mc_fits_mpi.py:81-159).  The device path is
batched, so the drivers here produce BATCHES:

  stencil(theta0, steps)        central-difference stencil of one point: 2P+1 vectors for the
                                gradient, P^2+P+1 with the mixed terms for the Hessian;
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

    Row 0 is theta0; rows 1+2j, 2+2j are theta0 +- steps[j] e_j.  With hessian=True two rows per
    pair j<k follow, (++, --): together with the axis points they give the mixed second
    derivative to O(h^2),  H_jk = [f(++) + f(--) - f(+j) - f(-j) - f(+k) - f(-k) + 2 f0] / (2 h_j h_k)
    -- half the points of the four-corner formula at the same order.  Returns [n, P]."""
    theta0 = np.asarray(theta0, dtype=np.float64)
    steps = np.asarray(steps, dtype=np.float64)
    return _stencil_batch(theta0[None, :], steps[None, :], hessian)[0]


def stencil_size(P, hessian=False):
    return 1 + 2 * P + (P * (P - 1) if hessian else 0)


def derivatives_from_stencil(f, steps, hessian=False):
    """(f0, grad[P], hess[P,P] or None) from the values f[n] on a `stencil` batch."""
    f = np.asarray(f, dtype=np.float64)
    steps = np.asarray(steps, dtype=np.float64)
    f0, grad, hess = _derivatives_batch(f[None, :], steps[None, :], hessian)
    return f0[0], grad[0], (hess[0] if hessian else None)


def _stencil_batch(centre, h, hessian):
    """Stencils of A points at once: centre [A, F], steps h [A, F] -> [A, n_st, F]
    (row order of `stencil`)."""
    A, F = centre.shape
    n_st = stencil_size(F, hessian)
    out = np.repeat(centre[:, None, :], n_st, axis=1)
    j = np.arange(F)
    out[:, 1 + 2 * j, j] += h
    out[:, 2 + 2 * j, j] -= h
    if hessian and F > 1:
        ja, jb = np.triu_indices(F, 1)
        rows = 1 + 2 * F + 2 * np.arange(ja.size)
        out[:, rows, ja] += h[:, ja]
        out[:, rows, jb] += h[:, jb]
        out[:, rows + 1, ja] -= h[:, ja]
        out[:, rows + 1, jb] -= h[:, jb]
    return out


def _derivatives_batch(f, h, hessian):
    """(f0 [A], grad [A, F], hess [A, F, F] -- only its diagonal without `hessian`) from the
    values f [A, n_st] on `_stencil_batch` rows."""
    A, F = h.shape
    f0 = f[:, 0]
    fp, fm = f[:, 1:1 + 2 * F:2], f[:, 2:2 + 2 * F:2]
    grad = (fp - fm) / (2 * h)
    hess = np.zeros((A, F, F))
    j = np.arange(F)
    hess[:, j, j] = (fp - 2 * f0[:, None] + fm) / h**2
    if hessian and F > 1:
        ja, jb = np.triu_indices(F, 1)
        rows = 1 + 2 * F + 2 * np.arange(ja.size)
        axis = fp + fm                                     # f(+j) + f(-j)
        v = (f[:, rows] + f[:, rows + 1] - axis[:, ja] - axis[:, jb] + 2 * f0[:, None]) / (2 * h[:, ja] * h[:, jb])
        hess[:, ja, jb] = v
        hess[:, jb, ja] = v
    return f0, grad, hess


# ---------------------------------------------------------------------------------------------
# damped Newton on batched stencils, M independent problems in lock-step
# ---------------------------------------------------------------------------------------------
class _Lockstep:
    """State shared by the two lock-step minimisers: bounds, the batched stencil evaluation and
    the bound-aware Levenberg-Marquardt step."""

    def __init__(self, eval_batch, theta0, steps, lower, upper, free, step_frac):
        self.eval_batch = eval_batch
        self.theta = np.array(theta0, dtype=np.float64, ndmin=2)
        self.M, P = self.theta.shape
        free = np.ones(P, bool) if free is None else np.asarray(free, bool)
        self.fidx = np.nonzero(free)[0]
        self.F = self.fidx.size
        lower = np.full(P, np.nan) if lower is None else np.asarray(lower, float)
        upper = np.full(P, np.nan) if upper is None else np.asarray(upper, float)
        self.lo = np.where(np.isnan(lower[self.fidx]), -np.inf, lower[self.fidx])
        self.hi = np.where(np.isnan(upper[self.fidx]), np.inf, upper[self.fidx])
        self.scale = np.tile(np.asarray(steps, float)[self.fidx], (self.M, 1))   # per problem 1-sigma estimate
        self.step_frac = step_frac
        self.n_eval = 0

    def clip(self, x, h):
        return np.minimum(np.maximum(x, self.lo + h), self.hi - h)

    def evaluate(self, act, centre, h, hessian):
        """(f0, grad, hess) of problems `act` at `centre` from ONE eval_batch call."""
        st = _stencil_batch(centre, h, hessian)
        n_st = st.shape[1]
        batch = np.repeat(self.theta[act], n_st, axis=0)
        batch[:, self.fidx] = st.reshape(-1, self.F)
        f = np.asarray(self.eval_batch(batch, np.repeat(act, n_st)), dtype=np.float64)
        self.n_eval += batch.shape[0]
        return _derivatives_batch(f.reshape(act.size, n_st), h, hessian)

    def near_bound(self, x, h):
        return (np.isfinite(self.lo) & (x <= self.lo + 2 * h)) | (np.isfinite(self.hi) & (x >= self.hi - 2 * h))

    def frozen(self, x, g, hm):
        """Parameters sitting on a bound whose gradient pushes outwards take no part in the step
        (MIGRAD's limits act through a variable transform instead)."""
        at_lo = np.isfinite(self.lo) & (x <= self.lo + 1.5 * hm) & (g > 0)
        at_hi = np.isfinite(self.hi) & (x >= self.hi - 1.5 * hm) & (g < 0)
        return at_lo | at_hi

    def rescale(self, m, x, h, diag):
        """Re-estimate the 1-sigma scale (and with it the stencil step) from the curvature, except
        for parameters resting on a bound: their centre is bound + step, so a changing step would
        move the point under comparison."""
        upd = (diag > 0) & ~self.near_bound(x, h)
        self.scale[m, upd] = np.sqrt(2. / diag[upd])


def newton_lockstep(eval_batch, theta0, steps, lower=None, upper=None, free=None, max_iter=50,
                    edm_tol=1e-5, step_frac=0.02, verbose=False, scale0=None, problems=None):
    """Minimise M independent functions that share one batched evaluator.

    eval_batch(theta[n, P], problem[n]) -> f[n]: value of problem problem[i] at theta[i].
    theta0 [M, P] start points, steps [P] expected 1-sigma scales (the stencil uses
    step_frac * steps, re-estimated from the Hessian once it is known), lower/upper [P]
    bounds (nan = none), free [P] bool.  Every iteration is ONE eval_batch call of
    M_active * (F^2 + F + 1) vectors (F free parameters): value, gradient and Hessian of every
    active problem, followed by a Levenberg-Marquardt step per problem (batched linear algebra).
    `scale0` [M, F]: curvature scales already known to the caller; `problems`: only these are
    advanced.  Returns dict(theta[M,P], fval[M], hess[M,F,F], edm[M], n_iter[M], converged[M],
    n_eval)."""
    ls = _Lockstep(eval_batch, theta0, steps, lower, upper, free, step_frac)
    M, F, fidx = ls.M, ls.F, ls.fidx
    if scale0 is not None:
        ls.scale = np.array(scale0, dtype=np.float64)
    lam = np.full(M, 1e-3 if scale0 is None else 1e-6)
    fval = np.full(M, np.inf)
    x_best = ls.theta[:, fidx].copy()
    g_best, H_best = np.zeros((M, F)), np.tile(np.eye(F), (M, 1, 1))
    have_best = np.zeros(M, bool)
    edm = np.full(M, np.inf)
    n_iter = np.zeros(M, int)
    done = np.zeros(M, bool)
    if problems is not None:
        done[:] = True
        done[np.asarray(problems, dtype=int)] = False
    trial = x_best.copy()
    for it in range(max_iter):
        act = np.nonzero(~done)[0]
        if act.size == 0 or F == 0:
            break
        h = step_frac * ls.scale[act]
        centre = ls.clip(trial[act], h)
        f0, g, H = ls.evaluate(act, centre, h, True)
        n_iter[act] += 1
        ok = np.isfinite(f0) & (f0 < 1e99) & ~(have_best[act] & (f0 > fval[act]))
        for i, m in enumerate(act):
            if ok[i]:
                fval[m], have_best[m] = f0[i], True
                x_best[m], g_best[m], H_best[m] = centre[i], g[i], H[i]
                lam[m] = max(lam[m] / 10., 1e-9)
                ls.rescale(m, centre[i], h[i], np.diag(H[i]))
            else:
                lam[m] = min(lam[m] * 10., 1e8)                  # reject: back to the best point
                if lam[m] >= 1e8:
                    done[m] = True
        live = act[have_best[act] & ~done[act]]
        if live.size:
            hm = step_frac * ls.scale[live]
            p, e, pos = _lm_steps(g_best[live], H_best[live], lam[live], ls.frozen(x_best[live], g_best[live], hm))
            edm[live] = np.where(lam[live] < 1e-2, e, np.inf)
            conv = (edm[live] < edm_tol) & pos
            done[live[conv]] = True
            trial[live] = x_best[live] + p
        cold = act[~have_best[act] & ~done[act]]
        trial[cold] = ls.clip(trial[cold], step_frac * ls.scale[cold]) * (1 + 1e-3)   # nudge off a failing start
        if verbose:
            print(f'[newton] iteration {it + 1}: {act.size} active, fval[0] = {fval[0]:.6f}, '
                  f'edm[0] = {edm[0]:.3e}')
    ls.theta[:, fidx] = x_best
    if F == 0:
        fval = np.asarray(eval_batch(ls.theta, np.arange(M)), dtype=np.float64)
        ls.n_eval += M
        edm[:] = 0.0
        done[:] = True
    return {'theta': ls.theta, 'fval': fval, 'hess': H_best, 'edm': edm, 'n_iter': n_iter,
            'converged': done & np.isfinite(edm) & (edm < edm_tol), 'n_eval': ls.n_eval, 'free': fidx,
            'scale': ls.scale}


# ---------------------------------------------------------------------------------------------
# quasi-Newton on gradient stencils, Hessian stencil only to finish
# ---------------------------------------------------------------------------------------------
def _positive_definite(H, floor=1e-4):
    """Nearest usable start Hessian(s) for a BFGS iteration: in the diagonally scaled space the
    eigenvalues are lifted to at least `floor` times the largest (a finite-difference Hessian taken
    away from the minimum can be indefinite along the bias / beta degeneracies, and BFGS keeps
    whatever definiteness it starts from)."""
    H = np.array(H, dtype=np.float64)
    single = H.ndim == 2
    H = H[None] if single else H
    d = np.sqrt(np.abs(np.einsum('aii->ai', H))) + 1e-300
    Hs = H / d[:, :, None] / d[:, None, :]
    Hs = 0.5 * (Hs + np.swapaxes(Hs, 1, 2))
    w, v = np.linalg.eigh(Hs)
    w = np.maximum(w, floor * np.max(w, axis=1, keepdims=True))
    out = np.einsum('aij,aj,akj->aik', v, w, v) * d[:, :, None] * d[:, None, :]
    return out[0] if single else out


def _lm_steps(g, H, lam, frozen):
    """Batched Levenberg-Marquardt steps in the diagonally scaled space: solves
    (Hs + lam I) p = -gs per problem; parameters flagged in `frozen` [A, F] (resting on a bound
    with the gradient pushing outwards) take no part.  Returns (step [A, F], edm [A],
    pos_def [A])."""
    A, F = g.shape
    eye = np.eye(F)
    d = np.sqrt(np.abs(np.einsum('aii->ai', H))) + 1e-300
    Hs = H / d[:, :, None] / d[:, None, :]
    gs = np.where(frozen, 0.0, g / d)
    cross = frozen[:, :, None] | frozen[:, None, :]
    Hs = np.where(cross, eye[None], Hs)
    Hs = np.where(np.isfinite(Hs), Hs, eye[None])
    mat = Hs + lam[:, None, None] * eye
    try:
        p = -np.linalg.solve(mat, gs[:, :, None])[:, :, 0]
    except np.linalg.LinAlgError:
        p = np.zeros_like(gs)
        for a in range(A):
            try:
                p[a] = -np.linalg.solve(mat[a], gs[a])
            except np.linalg.LinAlgError:
                p[a] = -gs[a] / (1 + lam[a])
    edm = 0.5 * np.abs(np.einsum('af,af->a', gs, -p))
    try:
        pos_def = np.linalg.eigvalsh(Hs).min(axis=1) > 0
    except np.linalg.LinAlgError:
        pos_def = np.zeros(A, bool)
    return np.where(frozen, 0.0, p / d), edm, pos_def


def quasi_newton_lockstep(eval_batch, theta0, steps, lower=None, upper=None, free=None, max_iter=60,
                          edm_tol=1e-5, step_frac=0.02, hess0=None, verbose=False):
    """Minimise M independent functions that share one batched evaluator, spending GRADIENT
    stencils (2F + 1 vectors) on the way down and a Hessian stencil (F^2 + F + 1) only to finish.

    Phase 1: BFGS on the Hessian, started from the diagonal second differences the gradient
    stencil yields for free (or from `hess0` [F, F] / [M, F, F]: in a set of mock fits every
    problem shares the model and the covariance, so one Hessian at the fiducial point serves all
    of them), Levenberg-Marquardt damping on rejection, the same bound handling as
    `newton_lockstep`.  Phase 2: Newton iterations on the full stencil from the phase-1 point
    until EDM < edm_tol with a positive-definite Hessian -- normally one, which is also the HESSE
    step the errors come from.  Every phase advances all active problems with ONE eval_batch call
    per iteration, and the per-problem linear algebra is batched.  Same arguments and result as
    `newton_lockstep`; at F = 16 a phase-1 iteration costs 33 vectors instead of 273."""
    ls = _Lockstep(eval_batch, theta0, steps, lower, upper, free, step_frac)
    M, F, fidx = ls.M, ls.F, ls.fidx
    if F == 0:
        return newton_lockstep(eval_batch, theta0, steps, lower, upper, free, max_iter, edm_tol, step_frac)
    n_iter = np.zeros(M, int)
    fval = np.full(M, np.inf)
    edm = np.full(M, np.inf)
    x_best = ls.theta[:, fidx].copy()
    g_best = np.zeros((M, F))
    H = np.tile(np.eye(F), (M, 1, 1))
    if hess0 is not None:
        H[:] = _positive_definite(np.asarray(hess0, dtype=np.float64))
        d0 = np.einsum('aii->ai', H)
        ls.scale = np.where(d0 > 0, np.sqrt(2. / np.where(d0 > 0, d0, 1.)), ls.scale)
    have_best = np.zeros(M, bool)
    have_H = np.full(M, hess0 is not None)
    lam = np.full(M, 1e-3)
    done = np.zeros(M, bool)
    trial = x_best.copy()
    for it in range(max_iter):
        act = np.nonzero(~done)[0]
        if act.size == 0:
            break
        h = step_frac * ls.scale[act]
        centre = ls.clip(trial[act], h)
        f0, g, Hd = ls.evaluate(act, centre, h, False)
        n_iter[act] += 1
        diag = np.einsum('aii->ai', Hd)
        ok = np.isfinite(f0) & (f0 < 1e99) & np.all(np.isfinite(g), axis=1)
        better = ok & (~have_best[act] | (f0 <= fval[act] + 1e-12 * np.abs(fval[act])))
        for i, m in enumerate(act):
            if better[i]:
                if have_best[m] and have_H[m]:           # BFGS update of the Hessian
                    sv, yv = centre[i] - x_best[m], g[i] - g_best[m]
                    Hs_ = H[m].dot(sv)
                    sy, sHs = sv.dot(yv), sv.dot(Hs_)
                    if sy > 1e-10 * np.linalg.norm(sv) * np.linalg.norm(yv) and sHs > 0:
                        H[m] += np.outer(yv, yv) / sy - np.outer(Hs_, Hs_) / sHs
                if not have_H[m]:                        # first point: diagonal second differences
                    H[m] = np.diag(np.where(diag[i] > 0, diag[i], 2.0 / ls.scale[m]**2))
                    have_H[m] = True
                if hess0 is None:
                    ls.rescale(m, centre[i], h[i], diag[i])
                x_best[m], g_best[m], fval[m], have_best[m] = centre[i], g[i], f0[i], True
                lam[m] = max(lam[m] / 10., 1e-9)
            else:
                lam[m] = min(lam[m] * 10., 1e8)
                if lam[m] >= 1e8:
                    done[m] = True
        live = act[have_best[act] & ~done[act]]
        if live.size:
            hm = step_frac * ls.scale[live]
            p, e, pos = _lm_steps(g_best[live], H[live], lam[live], ls.frozen(x_best[live], g_best[live], hm))
            edm[live] = np.where(lam[live] < 1e-2, e, np.inf)
            conv = (edm[live] < edm_tol) & pos & better[np.searchsorted(act, live)]
            done[live[conv]] = True
            trial[live] = x_best[live] + p
        cold = act[~have_best[act] & ~done[act]]
        trial[cold] = ls.clip(trial[cold], step_frac * ls.scale[cold]) * (1 + 1e-3)
        if verbose:
            print(f'[quasi-newton] iteration {it + 1}: {act.size} active, fval[0] = {fval[0]:.6f}, '
                  f'edm[0] = {edm[0]:.3e}')

    # ---- phase 2: Newton on the Hessian stencil from the phase-1 points ----
    ls.theta[:, fidx] = x_best
    lower_full = None if lower is None else lower
    polish = newton_lockstep(eval_batch, ls.theta, steps, lower_full, upper, free=free, max_iter=12,
                             edm_tol=edm_tol, step_frac=step_frac, scale0=ls.scale, verbose=verbose,
                             problems=np.nonzero(have_best)[0])
    polish['n_eval'] += ls.n_eval
    polish['n_gradient_iter'] = n_iter
    polish['n_iter'] = n_iter + polish['n_iter']
    return polish


def lockstep(method):
    """'quasi' (gradient stencils + a finishing Hessian stencil) or 'newton' (a Hessian stencil per
    iteration)."""
    if method not in ('quasi', 'newton'):
        raise ValueError(f'Unknown lock-step method {method}.')
    return quasi_newton_lockstep if method == 'quasi' else newton_lockstep


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

    def __init__(self, chi2_batch, sample_params, max_iter=50, edm_tol=1e-5, verbose=True, method='quasi'):
        self.chi2_batch = chi2_batch
        self.method = method
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

    def chi2_and_grad(self, theta, steps=None):
        """(chi2, gradient[P]) at one point from ONE batched call of the 2P + 1 central-difference
        stencil.  steps [P]: finite-difference steps (default 1e-3 of the [sample] errors -- the
        chi2 is smooth, the device evaluates it to ~1e-13 relative)."""
        theta = np.asarray(theta, dtype=np.float64)
        if steps is None:
            steps = 1e-3 * np.array([self._sample_params['errors'][n] for n in self._names], dtype=np.float64)
        f0, grad, _ = derivatives_from_stencil(self.chi2_batch(stencil(theta, steps)), steps)
        return float(f0), grad

    def grad(self, *pars):
        """Gradient callable with iminuit's signature: Minuit(self.chi2, *start, grad=self.grad,
        name=names) at a site where iminuit is installed (vega/minimizer.py:39-103 builds the
        Minuit object from the chi2 alone and lets MIGRAD difference it point by point; here each
        gradient is one batch)."""
        return self.chi2_and_grad(np.array(pars, dtype=np.float64))[1]

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
        res = lockstep(self.method)(lambda theta, prob: self.chi2_batch(theta), start[None, :],
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

    def draw(self, n_mocks, seed=0, forecast=False, keep=None):
        """masked mocks {name: [n_kept, n_mask]}, drawn as run_monte_carlo draws them
        (vega/analysis.py:245,261-263): np.random.seed(seed) once, then one
        Data.create_monte_carlo per correlation per mock from the running stream
        (forecast: the fiducial itself, no noise).

        keep: slice of the mock indices this process realises (default: all).  Every process still
        consumes the random stream of ALL n_mocks mocks, so mock m is the same vector whichever
        rank holds it; only the kept ones are multiplied by the Cholesky factor, kept and
        uploaded (`draw_sharded` keeps this rank's contiguous share)."""
        vega = self.vega
        kept = np.arange(n_mocks)[keep] if keep is not None else np.arange(n_mocks)
        self.n_total, self.mock_ids = int(n_mocks), kept
        keep_set = set(int(m) for m in kept)
        mocks = {name: [] for name in vega.corr_items}
        self.full_mocks = {name: [] for name in vega.corr_items}
        if seed is not None:
            np.random.seed(seed)
        if vega._use_global_cov:
            self.mocks = self._draw_global(n_mocks, forecast, kept)
        else:
            for m in range(n_mocks):
                for name in vega.corr_items:
                    data = vega.data[name]
                    if m not in keep_set:
                        # another rank's mock: advance the stream by what create_monte_carlo draws
                        if not forecast:
                            np.random.randn(int(data.data_mask.sum()) if data.cholesky_masked_cov
                                            else data.full_data_size)
                        continue
                    full = data.create_monte_carlo(self.fiducial_model[name], self._item_scale(name), None, forecast)
                    self.full_mocks[name].append(np.array(full))
                    mocks[name].append(data.masked_mc_mock.copy())
            self.mocks = {name: np.array(v).reshape(len(kept), -1) for name, v in mocks.items()}
        vega.monte_carlo = True
        engine = vega._engine_for(tuple(self.names))   # swaps in the scaled inverse covariance (MC mode)
        if len(kept):
            for ci, name in enumerate(vega.corr_items):
                engine.set_mocks(ci, self.mocks[name])
        self.engine = engine
        return self.mocks

    def draw_sharded(self, n_mocks, seed=0, forecast=False, group=None):
        """draw() of this rank's contiguous share of the mocks (one process per GPU)."""
        import torch.distributed as dist
        from .sharding import shard_bounds
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        rank = dist.get_rank(group) if dist.is_initialized() else 0
        lo, hi = shard_bounds(n_mocks, world, rank)
        return self.draw(n_mocks, seed=seed, forecast=forecast, keep=slice(lo, hi))

    def _draw_global(self, n_mocks, forecast=False, kept=None):
        """Mocks of the joint data vector under a global covariance
        (Analysis.create_global_monte_carlo, vega/analysis.py:164-222): one Cholesky factor of
        scale * cov[mask][:, mask], mock = masked fiducial + L . randn(n) from the running
        stream, one draw per mock.  The chi2 keeps the UNSCALED masked global inverse
        (vega/vega_interface.py:295-303), as the reference does.  Returned split per correlation
        in the order the residuals are concatenated.  The normal deviates of every mock are drawn
        (same stream on every rank); L multiplies only the `kept` ones -- on the device when the
        covariance is DESI-sized."""
        vega = self.vega
        names = list(vega.corr_items)
        kept = np.arange(n_mocks) if kept is None else np.asarray(kept, dtype=int)
        full_mask = np.concatenate([vega.data[n].data_mask for n in names])
        masked_cov = vega.global_cov[:, full_mask][full_mask, :]
        scale = 1 if self.scale is None else self.scale
        if isinstance(scale, dict):
            raise ValueError('A global covariance takes one scale, not one per correlation.')
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
            joint = np.tile(fid, (kept.size, 1))
        else:
            z = np.stack([np.random.randn(fid.size) for _ in range(n_mocks)])[kept]
            on_device = masked_cov.shape[0] >= utils.DEVICE_LINALG_MIN and self.torch.cuda.is_available()
            chol = None
            if on_device:
                # DESI-sized joint covariance (9540 rows): factor and multiply on the device (init time)
                dev = self.torch.from_numpy(np.ascontiguousarray(scale * masked_cov)).cuda()
                factor, info = self.torch.linalg.cholesky_ex(dev)
                if int(info) == 0:
                    zt = self.torch.from_numpy(np.ascontiguousarray(z)).cuda()
                    joint = fid[None, :] + self.torch.matmul(zt, factor.T).cpu().numpy()
                    chol = factor
            if chol is None:
                chol = np.linalg.cholesky(scale * masked_cov)
                joint = fid[None, :] + z @ chol.T
        self.global_mocks = joint
        edges = np.cumsum([0] + [int(vega.data[n].data_mask.sum()) for n in names])
        mocks = {n: np.ascontiguousarray(joint[:, edges[i]:edges[i + 1]]) for i, n in enumerate(names)}
        if kept.size:
            vega.analysis.current_mc_mock = joint[-1].copy()       # vega/analysis.py:217-220
        for n in names:      # what the scalar MC-mode chi2 reads (the last mock drawn, as in the reference)
            data = vega.data[n]
            if kept.size:
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
        return fit_mocks_sharded(self, self.n_total, self.engine.device, **fit_kwargs)

    def fit(self, mock_slice=None, params=None, max_iter=50, edm_tol=1e-5, verbose=False, method='quasi',
            share_hessian=True):
        """Best fit of every mock in mock_slice (default: all).  Returns the dict of the lock-step
        minimiser plus 'names' and 'errors'[M, P].

        method 'quasi': gradient stencils on the way down and one Hessian stencil per mock to finish
        (quasi_newton_lockstep); with `share_hessian` the BFGS iteration of EVERY mock starts from
        the mean Hessian of the first (up to 8) mocks at the start point -- the mocks share the model
        and the covariance, so their Hessians agree to the noise in the data.  'newton': a Hessian stencil
        per mock per iteration."""
        if self.mocks is None:
            raise RuntimeError('draw() the mocks first.')
        # mock_slice selects GLOBAL mock indices; the rows held locally are self.mock_ids
        wanted = np.arange(self.n_total)[mock_slice] if mock_slice is not None else self.mock_ids
        local = {int(m): i for i, m in enumerate(self.mock_ids)}
        missing = [int(m) for m in wanted if int(m) not in local]
        if missing:
            raise ValueError(f'Mocks {missing[:5]}... were not drawn by this process (draw(keep=...)).')
        ids = np.array([local[int(m)] for m in wanted], dtype=np.int64)
        sp = self.sample_params
        if params is not None:
            sp = {k: {**sp[k], **params.get(k, {})} for k in ('values', 'errors', 'limits', 'fix')}
        start = np.array([sp['values'][n] for n in self.names], dtype=np.float64)
        errors = np.array([sp['errors'][n] for n in self.names], dtype=np.float64)
        lower = np.array([np.nan if sp['limits'][n][0] is None else sp['limits'][n][0] for n in self.names])
        upper = np.array([np.nan if sp['limits'][n][1] is None else sp['limits'][n][1] for n in self.names])
        fixed = np.array([bool(sp['fix'][n]) for n in self.names])

        def evaluate(theta, prob):
            return self.chi2(theta, ids[prob])
        kwargs = {}
        n_extra = 0
        if method == 'quasi' and share_hessian and ids.size > 1 and (~fixed).any():
            # the Hessian every mock starts from: the Hessian stencil of the first mock AT THE START
            # POINT (the right quadratic model for the first step of every mock: they share start,
            # model and covariance; measured on the DESI-shaped model, the Hessian at a converged
            # minimum is the worse model there -- the first Newton steps overshoot along the
            # bias / beta / HCD degeneracies and are rejected).  A finite-difference Hessian away
            # from the minimum can be indefinite along those directions: quasi_newton_lockstep
            # lifts its eigenvalues (_positive_definite) before use.
            # ... taken as the MEAN over the first few mocks: one mock's finite-difference Hessian carries
            # that mock's noise (second derivatives of the model times its residual), and a rank whose
            # first mock is an outlier would hand a poor quadratic model to all of its fits (seen at
            # N = 2: one rank needed 3.4x the evaluations of the other)
            n_h = int(min(8, ids.size))
            ls = _Lockstep(lambda theta, prob: self.chi2(theta, ids[prob]), np.tile(start, (n_h, 1)), errors, lower,
                           upper, ~fixed, 0.02)
            h = 0.02 * ls.scale
            f0, g, hess0 = ls.evaluate(np.arange(n_h), ls.clip(np.tile(start[~fixed], (n_h, 1)), h), h, True)
            n_extra = ls.n_eval
            ok = np.all(np.isfinite(hess0.reshape(n_h, -1)), axis=1)
            if ok.any():
                kwargs['hess0'] = hess0[ok].mean(axis=0)
        if method == 'quasi':
            res = quasi_newton_lockstep(evaluate, np.tile(start, (ids.size, 1)), errors, lower, upper, free=~fixed,
                                        max_iter=max_iter, edm_tol=edm_tol, verbose=verbose, **kwargs)
        else:
            res = newton_lockstep(evaluate, np.tile(start, (ids.size, 1)), errors, lower, upper, free=~fixed,
                                  max_iter=max_iter, edm_tol=edm_tol, verbose=verbose)
        res['n_eval'] += n_extra
        self.last_n_eval = res['n_eval']
        errs = np.zeros((ids.size, len(self.names)))
        free_idx = res['free']
        try:
            cov = 2.0 * np.linalg.inv(res['hess'])
            errs[:, free_idx] = np.sqrt(np.maximum(np.einsum('mii->mi', cov), 0.))
        except np.linalg.LinAlgError:
            for m in range(ids.size):
                try:
                    errs[m, free_idx] = np.sqrt(np.maximum(np.diag(2.0 * np.linalg.inv(res['hess'][m])), 0.))
                except np.linalg.LinAlgError:
                    errs[m, free_idx] = np.nan
        res.update(names=self.names, errors=errs, mock_ids=np.asarray(wanted))
        return res


class BatchedAnalysis:
    """`Analysis.chi2_scan` (vega/analysis.py:53-124) with every grid point fitted at once.

    The reference minimises once per grid point of the one or two scanned parameters, those
    parameters fixed at the grid value.  The fits are independent, so here they are the M
    problems of one newton_lockstep call: each Newton iteration is one batched chi2 call of
    (active grid points) x (F^2 + F + 1) vectors.  Returns the reference's list of dictionaries
    (best-fit values of every sampled parameter + 'fval'), in the reference's loop order."""

    def __init__(self, chi2_batch, sample_params, main_config, max_iter=50, edm_tol=1e-5, verbose=True,
                 method='quasi'):
        self._chi2_batch = chi2_batch
        self.method = method
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
        res = lockstep(self.method)(lambda theta, prob: self._chi2_batch(theta), theta0, errors, lower, upper,
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
