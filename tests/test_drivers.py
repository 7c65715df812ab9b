"""Batched drivers (vega_b200/drivers.py).  CPU tests cover the stencil algebra, the lock-step
Newton iteration, the Minimizer surface and the sampler adapter with closed-form
functions; GPU tests drive the device engine and check against the oracle."""
import numpy as np
import pytest

from vega_b200.drivers import (BatchedMinimizer, BatchedSamplerLikelihood, MonteCarloFits,
                               derivatives_from_stencil, newton_lockstep, stencil, stencil_size)


def test_stencil_derivatives_of_a_quadratic_are_exact():
    rng = np.random.default_rng(1)
    P = 4
    A = rng.standard_normal((P, P))
    A = A + A.T
    b, x0 = rng.standard_normal(P), rng.standard_normal(P)
    steps = np.array([0.1, 0.2, 0.05, 0.3])

    def f(x):
        return 0.5 * np.einsum('bi,ij,bj->b', x, A, x) + x.dot(b) + 1.5

    for hessian in (False, True):
        st = stencil(x0, steps, hessian=hessian)
        assert st.shape == (stencil_size(P, hessian), P)
        f0, g, H = derivatives_from_stencil(f(st), steps, hessian=hessian)
        assert np.isclose(f0, f(x0[None])[0])
        assert np.allclose(g, A.dot(x0) + b, rtol=1e-10, atol=1e-12)
        if hessian:
            assert np.allclose(H, A, rtol=1e-8, atol=1e-10)


def test_newton_lockstep_bounds_and_fixed_parameters():
    rng = np.random.default_rng(2)
    M, P = 3, 4
    targets = rng.uniform(-1, 1, (M, P))
    calls = []

    def ev(theta, prob):
        calls.append(theta.shape[0])
        d = theta - targets[prob]
        return (d**2).sum(axis=1) * np.array([1., 4., 9.])[prob] + 0.05 * (d**4).sum(axis=1)

    lower = np.array([np.nan, np.nan, 0.5, np.nan])      # parameter 2 is bounded away from its target
    upper = np.full(P, np.nan)
    targets[:, 2] = -0.3
    free = np.array([True, True, True, False])           # parameter 3 stays at its start value
    res = newton_lockstep(ev, np.zeros((M, P)) + 0.7, np.ones(P), lower, upper, free=free)
    assert res['converged'].all() or np.all(res['theta'][:, 2] <= 0.5 + 0.05)
    assert np.allclose(res['theta'][:, :2], targets[:, :2], atol=5e-3)      # edm 1e-5 ~ 3e-3 sigma
    assert np.all(res['theta'][:, 2] >= 0.5) and np.all(res['theta'][:, 2] < 0.56)
    assert np.all(res['theta'][:, 3] == 0.7)
    # one evaluator call per iteration, (active problems) x (2 F^2 + 1) vectors each
    assert calls[0] == M * stencil_size(3, hessian=True)


def test_quasi_newton_matches_newton_with_fewer_evaluations():
    """16 correlated parameters, 12 problems that share the curvature (mock fits): the gradient
    stencil iteration + one finishing Hessian stencil reaches the same minima and Hessians as the
    Hessian-stencil Newton iteration in a fraction of the evaluations; a shared start Hessian
    (MonteCarloFits.fit(share_hessian=True)) cuts them further."""
    from vega_b200.drivers import quasi_newton_lockstep
    rng = np.random.default_rng(3)
    M, P = 12, 16
    sig = 10**rng.uniform(-3, 0, P)
    c = rng.standard_normal((P, P))
    c = c @ c.T
    c = 0.6 * c / np.sqrt(np.outer(np.diag(c), np.diag(c))) + 0.4 * np.eye(P)
    icov = np.linalg.inv(c * sig[:, None] * sig[None, :])
    targets = rng.standard_normal((M, P)) * sig

    def ev(theta, prob):
        d = theta - targets[prob]
        return np.einsum('bi,ij,bj->b', d, icov, d) + 0.02 * ((d / sig)**4).sum(axis=1)

    start = np.zeros((M, P))
    ref = newton_lockstep(ev, start, 2 * sig)
    got = quasi_newton_lockstep(ev, start, 2 * sig)
    shared = quasi_newton_lockstep(ev, start, 2 * sig, hess0=2 * icov)
    for res in (got, shared):
        assert res['converged'].all()
        assert np.max(np.abs(res['theta'] - ref['theta']) / sig) < 5e-3      # EDM 1e-5 ~ 3e-3 sigma
        assert np.allclose(res['fval'], ref['fval'], atol=1e-4)
        assert np.allclose(res['hess'], ref['hess'], rtol=2e-2, atol=1e-6 * np.abs(ref['hess']).max())
    assert got['n_eval'] < 0.8 * ref['n_eval'] and shared['n_eval'] < 0.5 * ref['n_eval']
    # a bound every problem runs into: the parameter rests one stencil step inside it
    lower = np.full(P, np.nan)
    lower[3] = targets[:, 3].max() + 0.5 * sig[3]
    start[:, 3] = lower[3] + sig[3]
    ref = newton_lockstep(ev, start, 2 * sig, lower)
    got = quasi_newton_lockstep(ev, start, 2 * sig, lower)
    assert got['converged'].all() and np.all(got['theta'][:, 3] >= lower[3])
    assert np.all(got['theta'][:, 3] < lower[3] + 0.05 * sig[3])
    assert np.max(np.abs(got['theta'] - ref['theta']) / sig) < 1e-2
    assert np.allclose(got['fval'], ref['fval'], atol=1e-2)


def test_gradient_callable_for_iminuit():
    """BatchedMinimizer.grad(*pars): one 2P + 1 batch per gradient, iminuit's calling convention."""
    names = ['a', 'b', 'c']
    batches = []

    def chi2_batch(theta):
        batches.append(theta.shape[0])
        return (theta[:, 0] - 1)**2 + 3 * (theta[:, 1] + 2)**2 + theta[:, 0] * theta[:, 2] + np.exp(0.1 * theta[:, 2])

    mini = BatchedMinimizer(chi2_batch, _sample(names, [0.3, 0.1, -0.2], [0.1, 0.1, 0.1]), verbose=False)
    x = np.array([0.3, 0.1, -0.2])
    g = mini.grad(*x)
    assert batches == [7]
    exact = np.array([2 * (x[0] - 1) + x[2], 6 * (x[1] + 2), x[0] + 0.1 * np.exp(0.1 * x[2])])
    assert np.allclose(g, exact, rtol=1e-7)
    f, g2 = mini.chi2_and_grad(x, steps=np.full(3, 1e-5))
    assert f == pytest.approx(float(chi2_batch(x[None])[0])) and np.allclose(g2, exact, rtol=1e-8)


def _sample(names, values, errors, limits=None):
    return {'limits': {n: (limits or {}).get(n, (None, None)) for n in names},
            'values': dict(zip(names, values)), 'errors': dict(zip(names, errors)),
            'fix': {n: False for n in names}}


def test_batched_minimizer_surface_and_two_stage_fit():
    names = ['ap', 'bias_LYA', 'beta_LYA']
    truth = np.array([1.02, -0.12, 1.6])
    sig = np.array([0.03, 0.004, 0.09])
    corr = np.array([[1, .3, 0], [.3, 1, -.5], [0, -.5, 1]])
    icov = np.linalg.inv(corr * sig[:, None] * sig[None, :])
    batches = []

    def chi2_batch(theta):
        batches.append(theta.shape[0])
        d = theta - truth
        return np.einsum('bi,ij,bj->b', d, icov, d) + 10.

    mini = BatchedMinimizer(chi2_batch, _sample(names, [1.0, -0.1, 1.5], [0.1, 0.01, 0.1]), verbose=False,
                           method='newton')
    with pytest.raises(RuntimeError):
        mini.values
    mini.minimize()
    assert np.allclose([mini.values[n] for n in names], truth, atol=1e-3 * sig.max() + 1e-6)
    assert np.isclose(mini.fmin.fval, 10., atol=1e-6) and mini.fmin.is_valid
    assert np.allclose(mini.covariance, corr * sig[:, None] * sig[None, :], rtol=1e-4)   # errordef = 1
    assert np.allclose([mini.errors[n] for n in names], sig, rtol=1e-4)
    # first stage: only the bias parameter free -> stencil of 1 parameter (3 vectors)
    assert batches[0] == stencil_size(1, hessian=True)
    assert mini.chi2(*truth) == pytest.approx(10.)
    # fixing a parameter through minimize(params) (vega/minimizer.py:52-61)
    mini.minimize({'fix': {'ap': True}, 'values': {'ap': 1.0}})
    assert mini.values['ap'] == 1.0 and mini.fixed['ap'] and mini.errors['ap'] == 0.0


def test_chi2_scan_fits_every_grid_point_in_one_batch():
    """BatchedAnalysis.chi2_scan (vega/analysis.py:53-124) on a correlated quadratic: the profile
    over the free parameter has a closed form."""
    import configparser
    from vega_b200.drivers import BatchedAnalysis
    names = ['ap', 'at', 'bias']
    truth = np.array([1.0, 1.0, -0.1])
    sig = np.array([0.03, 0.04, 0.01])
    corr = np.array([[1, .4, .2], [.4, 1, -.3], [.2, -.3, 1]])
    icov = np.linalg.inv(corr * sig[:, None] * sig[None, :])
    batches = []

    def chi2_batch(theta):
        batches.append(theta.shape[0])
        d = theta - truth
        return np.einsum('bi,ij,bj->b', d, icov, d) + 3.

    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda o: o
    cfg['chi2 scan'] = {'ap': '0.95 1.05 5', 'at': '0.9 1.1 3'}
    scan = BatchedAnalysis(chi2_batch, _sample(names, [1.01, 0.99, -0.11], [0.05, 0.05, 0.02]), cfg, verbose=False,
                          method='newton')
    results = scan.chi2_scan()
    assert len(results) == 15 and scan.converged.all()
    assert batches[0] == 15 * stencil_size(1, hessian=True)       # one free parameter per grid point
    k = 0
    for ap in np.linspace(0.95, 1.05, 5):            # the reference's loop order: first parameter outermost
        for at in np.linspace(0.9, 1.1, 3):
            r = results[k]
            k += 1
            assert r['ap'] == ap and r['at'] == at
            d2 = np.array([ap, at]) - truth[:2]
            bias = truth[2] - icov[2, :2].dot(d2) / icov[2, 2]            # profile over the free parameter
            d = np.array([ap, at, bias]) - truth
            # converged to the EDM tolerance (1e-5): << 1e-3 sigma on the parameter
            assert abs(r['bias'] - bias) < 1e-3 * sig[2] and abs(r['fval'] - (d.dot(icov).dot(d) + 3.)) < 1e-4
    cfg['chi2 scan'] = {'ap': '0.9 1 2', 'at': '0.9 1 2', 'bias': '-1 0 2'}
    with pytest.raises(ValueError):
        scan.chi2_scan()


class _StubVega:
    def __init__(self):
        self.sample_params = _sample(['a', 'b'], [0.5, 2.0], [0.1, 0.1], {'a': (0., 1.), 'b': (1., 5.)})

    def _log_norm(self):
        return -3.0

    def chi2_batch(self, theta, names):
        assert list(names) == ['a', 'b']
        return (theta**2).sum(axis=1)


def test_sampler_adapter_matches_polychord_calling_convention():
    like = BatchedSamplerLikelihood(_StubVega())
    assert like.num_params == 2 and like.num_derived == 0
    assert like.prior([0.25, 0.5]) == [0.25, 3.0]
    cube = np.array([[0., 0.], [1., 1.], [0.5, 0.25]])
    theta = like.prior_batch(cube)
    assert np.allclose(theta, [[0., 1.], [1., 5.], [0.5, 2.]])
    ll = like.log_lik_batch(theta)
    assert np.allclose(ll, -3.0 - 0.5 * (theta**2).sum(axis=1))
    value, derived = like.log_lik(theta[2])
    assert value == ll[2] and derived == []
    stub = _StubVega()
    stub.sample_params['limits']['a'] = (None, 1.)
    with pytest.raises(ValueError):
        BatchedSamplerLikelihood(stub)


# ---------------------------------------------------------------------------------------------
# on the device
# ---------------------------------------------------------------------------------------------
def _c2_small():
    """C2 operands (dense distortion / metal / covariance matrices), data drawn around the
    oracle's fiducial model."""
    from helpers import make_c2
    return make_c2(2, width=0.5)


@pytest.mark.gpu
def test_mock_chi2_matches_oracle_monte_carlo_mode():
    """vg_chi2_batch_mc: vector b against mock mock_idx[b] == the reference's monte_carlo
    branch of chi2 (vega/vega_interface.py:296-312) evaluated mock by mock."""
    vega, oracle, theta = _c2_small()
    names = list(vega.sample_params['limits'])
    mc = MonteCarloFits(vega, fiducial_model=oracle.compute_model(), scale=1.3)
    mocks = mc.draw(3, seed=11)
    assert all(v.shape == (3, int(vega.data[n].data_mask.sum())) for n, v in mocks.items())
    pairs = [(0, 2), (1, 0), (0, 1), (1, 1)]
    got = mc.chi2(np.array([theta[t] for t, _ in pairs]), [m for _, m in pairs])
    oracle.monte_carlo = True
    for (t, m), value in zip(pairs, got):
        for name in vega.corr_items:
            vega.data[name].masked_mc_mock = mocks[name][m]
        ref = oracle.chi2(dict(zip(names, theta[t])))
        assert abs(value - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4), (t, m, value, ref)


@pytest.mark.gpu
def test_batched_minimizer_on_the_device_path():
    """One Newton iteration = one batched chi2 call; the minimum is a stationary point of the
    ORACLE's chi2 (central differences of the CPU path) and lies below the start value."""
    vega, oracle, theta = _c2_small()
    names = list(vega.sample_params['limits'])
    mini = BatchedMinimizer.from_vega(vega, verbose=False)
    start = np.array([vega.sample_params['values'][n] for n in names])
    mini.minimize()
    best = np.array([mini.values[n] for n in names])
    f_start, f_best = vega.chi2_batch(np.array([start, best]))
    assert mini.fmin.is_valid and np.isclose(mini.fmin.fval, f_best, atol=1e-6)
    assert f_best < f_start
    ref = oracle.chi2(dict(zip(names, best)))
    assert abs(ref - f_best) <= 1e-6 * max(1.0, abs(ref) / 1e4)
    # stationarity on the oracle along two well-constrained directions
    for n in ('bias_eta_LYA', 'beta_LYA'):
        if n not in names:
            continue
        j, h = names.index(n), 0.1 * mini.errors[n]
        up, dn = best.copy(), best.copy()
        up[j] += h
        dn[j] -= h
        slope = (oracle.chi2(dict(zip(names, up))) - oracle.chi2(dict(zip(names, dn)))) / (2 * h)
        assert abs(slope) * mini.errors[n] < 0.05, (n, slope)     # << 1 chi2 unit per sigma


@pytest.mark.gpu
def test_monte_carlo_fits_in_lock_step():
    vega, oracle, theta = _c2_small()
    mc = MonteCarloFits(vega, fiducial_model=oracle.compute_model())
    mc.draw(4, seed=3)
    res = mc.fit(max_iter=30)
    assert res['theta'].shape == (4, len(mc.names)) and res['converged'].all()
    # every fit beats the truth evaluated on its own mock, by about the number of parameters
    truth = np.tile([vega.params[n] for n in mc.names], (4, 1))
    at_truth = mc.chi2(truth, np.arange(4))
    assert np.all(res['fval'] <= at_truth + 1e-6)
    assert np.all(at_truth - res['fval'] < 10 * len(mc.names))
    # the sharded call: a slice of the mocks gives the same answers
    part = mc.fit(mock_slice=slice(1, 3), max_iter=30)
    # (the slice starts its BFGS iteration from the Hessian of ITS first mock: same minima to the
    # EDM tolerance, 1e-5 ~ 3e-3 sigma)
    sigma = np.maximum(res['errors'][1:3], 1e-30)
    assert np.all(np.abs(part['theta'] - res['theta'][1:3]) <= 1e-2 * sigma + 1e-12)
    # with the Hessian-stencil iteration the slice retraces the full run (a different batch
    # composition only changes the split-K reduction order of the GEMMs at the 1e-13 level)
    full_n = mc.fit(max_iter=30, method='newton')
    part_n = mc.fit(mock_slice=slice(1, 3), max_iter=30, method='newton')
    sigma = np.maximum(full_n['errors'][1:3], 1e-30)
    assert np.all(np.abs(part_n['theta'] - full_n['theta'][1:3]) <= 1e-4 * sigma + 1e-12)
    vega.monte_carlo = False


@pytest.mark.gpu
def test_lock_step_methods_and_sharded_draws_agree():
    """quasi-Newton (gradient stencils, shared start Hessian, one finishing Hessian stencil) against
    the Hessian-stencil Newton iteration on the device chi2, and a draw that keeps only a slice of
    the mocks (every rank consumes the whole random stream) against the full draw."""
    vega, oracle, theta = _c2_small()
    fiducial = oracle.compute_model()
    mc = MonteCarloFits(vega, fiducial_model=fiducial)
    mc.draw(6, seed=9)
    newton = mc.fit(max_iter=30, method='newton')
    quasi = mc.fit(max_iter=40, method='quasi')
    assert newton['converged'].all() and quasi['converged'].all()
    sigma = np.maximum(newton['errors'], 1e-30)
    assert np.max(np.abs(quasi['theta'] - newton['theta']) / sigma) < 2e-2
    assert np.allclose(quasi['fval'], newton['fval'], atol=1e-3)
    assert np.allclose(quasi['errors'], newton['errors'], rtol=2e-2)
    assert quasi['n_eval'] < 0.6 * newton['n_eval']
    full = {n: v.copy() for n, v in mc.mocks.items()}
    part = MonteCarloFits(vega, fiducial_model=fiducial)
    part.draw(6, seed=9, keep=slice(2, 5))
    for n in full:
        assert np.array_equal(part.mocks[n], full[n][2:5])
    res = part.fit(mock_slice=slice(3, 5), max_iter=40)
    assert list(res['mock_ids']) == [3, 4]
    assert np.max(np.abs(res['theta'] - quasi['theta'][3:5]) / sigma[3:5]) < 2e-2
    with pytest.raises(ValueError):
        part.fit(mock_slice=slice(0, 2))
    vega.monte_carlo = False


@pytest.mark.gpu
def test_chi2_scan_on_the_device_path():
    """vega.analysis.chi2_scan(): every grid point is a constrained best fit of the device chi2."""
    vega, oracle, theta = _c2_small()
    names = list(vega.sample_params['limits'])
    assert 'ap' in names
    vega.main_config['chi2 scan'] = {'ap': '0.99 1.01 3'}
    results = vega.analysis.chi2_scan()
    assert [r['ap'] for r in results] == list(np.linspace(0.99, 1.01, 3))
    assert vega.analysis.converged.all()
    start = np.array([vega.sample_params['values'][n] for n in names])
    for r in results:
        best = np.array([r[n] for n in names])
        at_start = start.copy()
        at_start[names.index('ap')] = r['ap']
        f_best, f_start = vega.chi2_batch(np.array([best, at_start]), names)
        assert abs(f_best - r['fval']) <= 1e-6 * max(1.0, abs(f_best)) and f_best <= f_start + 1e-9
        # stationary in a free parameter
        j = next(i for i, n in enumerate(names) if n != 'ap' and not vega.sample_params['fix'][n])
        h = 1e-3 * abs(best[j]) + 1e-6
        up, dn = best.copy(), best.copy()
        up[j] += h
        dn[j] -= h
        f_up, f_dn = vega.chi2_batch(np.array([up, dn]), names)
        assert f_up >= f_best - 1e-7 and f_dn >= f_best - 1e-7


@pytest.mark.gpu
def test_mock_fits_with_a_global_covariance():
    """Mocks of the joint data vector (Analysis.create_global_monte_carlo, vega/analysis.py:164-222)
    drawn with the reference's random stream, their chi2 under the masked global C^-1
    (vega/vega_interface.py:295-303) against the oracle mock by mock, and a lock-step fit."""
    from helpers import set_fast_metals, static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import VegaInterface, synthetic
    vega = VegaInterface('configs/full_configs/main.ini')
    set_fast_metals(vega)
    synthetic.inject_global_covariance(vega)
    oracle = Oracle(vega, tables=static_tables())
    fiducial = oracle.compute_model()
    mc = MonteCarloFits(vega, fiducial_model=fiducial, scale=1.2)
    mocks = mc.draw(3, seed=5)
    # the reference's draw, restated: one Cholesky factor, one randn(n) per mock from seed 5
    names = list(vega.corr_items)
    mask = np.concatenate([vega.data[n].data_mask for n in names])
    chol = np.linalg.cholesky(1.2 * vega.global_cov[:, mask][mask, :])
    np.random.seed(5)
    fid = np.concatenate([fiducial[n] for n in names])[mask]
    for m in range(3):
        ref_mock = fid + chol.dot(np.random.randn(mask.sum()))
        assert np.max(np.abs(mc.global_mocks[m] - ref_mock)) <= 1e-12 * np.max(np.abs(ref_mock))
    assert np.array_equal(np.concatenate([mocks[n][2] for n in names]), mc.global_mocks[2])

    pars = mc.names
    base = np.array([vega.params[n] for n in pars])
    theta = base * (1 + 0.01 * np.random.default_rng(2).standard_normal((2, len(pars))))
    pairs = [(0, 1), (1, 2), (1, 0)]
    got = mc.chi2(np.array([theta[t] for t, _ in pairs]), [m for _, m in pairs])
    oracle.monte_carlo = True
    for (t, m), value in zip(pairs, got):
        vega.analysis.current_mc_mock = mc.global_mocks[m]
        ref = oracle.chi2(dict(zip(pars, theta[t])))
        assert abs(value - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4), (t, m, value, ref)

    res = mc.fit(max_iter=30)
    truth = np.tile(base, (3, 1))
    at_truth = mc.chi2(truth, np.arange(3))
    assert np.all(res['fval'] <= at_truth + 1e-6)
    assert np.all(at_truth - res['fval'] < 10 * len(pars))
    vega.monte_carlo = False


@pytest.mark.gpu
def test_monte_carlo_surface_of_the_scripts(tmp_path):
    """What bin/run_vega_mc_mpi.py:42-62 and vega/scripts/run_vega.py:24 call:
    get_fiducial_for_monte_carlo (initial fit over [sample], [mc parameters] on top),
    initialize_monte_carlo (one mock from mc_seed, minimizer on the [monte carlo] set, MC mode on)
    and analysis.run_monte_carlo (vega/analysis.py:224-308) with its result attributes."""
    import configparser
    from helpers import static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import VegaInterface, synthetic, utils
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda option: option
    cfg.read(utils.find_file('configs/c2/main.ini'))
    cfg['sample'] = {'bias_eta_LYA': 'True', 'beta_LYA': 'True'}
    cfg['priors'] = {}
    cfg['monte carlo'] = {'ap': 'True', 'at': 'True', 'bias_eta_LYA': 'True'}
    cfg['mc parameters'] = {'ap': '1.01', 'at': '0.99'}
    cfg['control']['mc_seed'] = '4'
    path = tmp_path / 'main_mc.ini'
    with open(path, 'w') as f:
        cfg.write(f)
    vega = VegaInterface(str(path))
    synthetic.inject_dense_operands(vega)
    synthetic.draw_mock_data(vega, vega.compute_model())
    quiet = lambda *a, **k: None    # noqa: E731

    mocks = vega.initialize_monte_carlo(print_func=quiet)
    assert vega.monte_carlo and list(mocks) == list(vega.corr_items)
    assert list(vega.minimizer._names) == ['ap', 'at', 'bias_eta_LYA']
    # the mock is the reference's draw: seed once, one create_monte_carlo per correlation in order
    np.random.seed(4)
    for name in vega.corr_items:
        data = vega.data[name]
        noise = data._cholesky.dot(np.random.randn(int(data.data_mask.sum())))
        fid = np.asarray(vega._mc_fits.fiducial_model[name])[data.data_mask]
        assert np.max(np.abs(mocks[name][data.data_mask] - (fid + noise))) <= 1e-14 * np.max(np.abs(fid))
        assert mocks[name].shape == (data.full_data_size,)
        assert np.array_equal(data.masked_mc_mock, mocks[name][data.data_mask])
    # scalar chi2 in Monte Carlo mode against the oracle's monte_carlo branch
    oracle = Oracle(vega, tables=static_tables())
    oracle.monte_carlo = True
    pars = {'ap': 1.004, 'at': 0.997, 'bias_eta_LYA': -0.2005}
    ref = oracle.chi2(pars)
    assert abs(vega.chi2(pars) - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4)

    best = vega.analysis.run_monte_carlo(vega._mc_fits.fiducial_model, num_mocks=3, seed=7)
    ana = vega.analysis
    assert ana.has_monte_carlo and set(best) == {'ap', 'at', 'bias_eta_LYA'}
    assert all(v.shape == (3, 2) for v in best.values())
    assert len(ana.mc_chisq) == len(ana.mc_covariances) == len(ana.mc_valid_minima) == 3
    assert ana.mc_failed_mask == [False] * 3 and all(ana.mc_valid_minima) and all(ana.mc_valid_hesse)
    n_dof = sum(int(d.data_mask.sum()) for d in vega.data.values()) - 3
    assert all(abs(c - n_dof) < 6 * np.sqrt(2 * n_dof) for c in ana.mc_chisq)
    assert all(len(ana.mc_mocks[n]) == 3 and ana.mc_mocks[n][0].shape == (vega.data[n].full_data_size,)
               for n in vega.corr_items)
    # best fits scatter around the template values by about their errors
    for name, truth in (('ap', 1.01), ('at', 0.99)):
        pull = (best[name][:, 0] - truth) / best[name][:, 1]
        assert np.all(np.abs(pull) < 5), (name, pull)
    # the files the Monte Carlo scripts end with (bin/run_vega_mc_fits_mpi.py:155-159)
    from vega_b200 import fitsio
    vega.main_config['output']['filename'] = str(tmp_path / 'out')
    vega._output = None
    mc_file = vega.output.write_monte_carlo(cpu_id=0)
    tables = {h.header.get('EXTNAME'): h.data for h in fitsio.open(mc_file)[1:]}
    assert np.array_equal(tables['Bestfit']['values'][0], best['ap'][:, 0])
    assert np.array_equal(tables['FitInfo']['chisq'], ana.mc_chisq)
    # a re-draw is seen by the scalar API: the reference reads data.masked_mc_mock on every call
    # (vega/vega_interface.py:296-312)
    before = vega.chi2(pars)
    for name in vega.corr_items:
        vega.data[name].create_monte_carlo(vega._mc_fits.fiducial_model[name], seed=11 + len(name))
    ref = oracle.chi2(pars)
    after = vega.chi2(pars)
    assert abs(after - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4) and abs(after - before) > 1.0
    for name in vega.corr_items:         # ... and so is a change of the covariance scale
        vega.data[name].create_monte_carlo(vega._mc_fits.fiducial_model[name], scale=2.0, seed=5)
    ref = oracle.chi2(pars)
    assert abs(vega.chi2(pars) - ref) <= 1e-6 * max(1.0, abs(ref) / 1e4)
    assert abs(vega.log_lik(pars) - oracle.log_lik(pars)) <= 1e-6 * max(1.0, abs(ref) / 1e4)
    vega.monte_carlo = False
    # mc_start_from_fit: template parameters from the BESTFIT table of a result file
    # (vega/vega_interface.py:466-471), [mc parameters] on top
    fit_file = tmp_path / 'fit.fits'
    fitsio.write_tables(fit_file, [('BESTFIT', {}, {'names': np.array(['bias_eta_LYA', 'beta_LYA']),
                                                    'values': np.array([-0.199, 1.7])})])
    vega.main_config['control']['mc_start_from_fit'] = str(fit_file)
    got = vega.get_fiducial_for_monte_carlo(print_func=quiet)
    want = vega.compute_model({'bias_eta_LYA': -0.199, 'beta_LYA': 1.7, 'ap': 1.01, 'at': 0.99}, run_init=False)
    assert all(np.array_equal(got[n], want[n]) for n in want)
