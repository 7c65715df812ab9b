"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle and the
reference's golden vectors.  Tolerances are BASELINE.json's: model xi within
1e-10 * max|xi| per correlation, chi2 within 1e-6 absolute (FP64)."""
import numpy as np
import pytest
import torch

from vega_b200 import fitsio, utils, synthetic
from helpers import make_oracle, make_c2

pytestmark = pytest.mark.gpu
XI_RTOL = 1e-10
CHI2_ATOL = 1e-6


def assert_xi_close(got, ref, label=''):
    for name in ref:
        scale = np.max(np.abs(ref[name]))
        err = np.max(np.abs(got[name] - ref[name]))
        assert err <= XI_RTOL * scale, f'{label} {name}: {err / scale:.3e}'


@pytest.mark.parametrize('kind,main', [('auto', 'configs/picca_bench/main.ini'),
                                       ('cross', 'configs/picca_bench/main_cross.ini')])
def test_picca_golden_vectors_on_gpu(kind, main):
    """reference tests/test_vega.py:21-44 (old_fftlog path) on the CUDA path."""
    vega, oracle = make_oracle(main, omega_de_none=True)
    xi = vega.compute_model()
    hdul = fitsio.open(utils.find_file('data/picca_bench_data.fits'))
    hdu = hdul[1] if kind == 'auto' else hdul[2]
    for i in (0, 1, 2, 4, 5, 6, 7):
        gold = np.array(hdu.data[f'{kind}_test_{i}'])
        assert np.allclose(xi[f'test_{i}'], gold), (kind, i)
    assert_xi_close(xi, oracle.compute_model(), kind)


def test_c1_auto_chi2():
    vega, oracle = make_oracle('configs/configs/main.ini')
    assert_xi_close(vega.compute_model(), oracle.compute_model(), 'c1')
    assert abs(vega.chi2() - oracle.chi2()) <= CHI2_ATOL


def test_full_config_known_answer():
    """reference tests/test_vega.py:13-14: log_lik == -8766.997108462287 (rel 1e-9)."""
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    assert_xi_close(vega.compute_model(), oracle.compute_model(), 'full')
    assert np.isclose(vega.log_lik(), -8766.997108462287, rtol=1e-9, atol=0)
    assert abs(vega.chi2() - oracle.chi2()) <= CHI2_ATOL
    assert abs(vega.log_lik() - oracle.log_lik()) <= CHI2_ATOL


def test_full_config_batch_matches_oracle():
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'L0_hcd', 'drp_QSO',
             'sigma_velo_disp_lorentz_QSO', 'bao_amp', 'alpha_LYA',
             'BB-lyalya_lyalya-0 add post r,mu (1,2)']
    rng = np.random.default_rng(5)
    base = np.array([vega.params[n] for n in names])
    theta = base * (1 + 0.05 * rng.standard_normal((6, len(names))))
    chi2 = vega.chi2_batch(theta, names)
    models, status = vega.compute_model_batch(theta, names)
    assert not status.any()
    for b in range(theta.shape[0]):
        pars = dict(zip(names, theta[b]))
        ref = oracle.compute_model(pars)
        assert_xi_close({n: models[n][b] for n in ref}, ref, f'vector {b}')
        assert abs(chi2[b] - oracle.chi2(pars)) <= CHI2_ATOL


def test_bounds_error_gives_sentinel():
    """vega/vega_interface.py:268-279: a spline-bounds failure makes chi2 return 1e100."""
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    pars = {'ap': 1e6, 'at': 1e6}
    assert oracle.chi2(pars) == 1e100
    assert vega.chi2(pars) == 1e100
    with pytest.raises(utils.VegaBoundsError):
        vega.compute_model(pars)


def test_c2_dense_operands_match_oracle():
    """eBOSS-shaped auto+cross with dense distortion / metal / covariance matrices."""
    vega, oracle, theta = make_c2(3, width=0.2)
    names = list(vega.sample_params['limits'])
    chi2 = vega.chi2_batch(theta)
    models, status = vega.compute_model_batch(theta)
    assert not status.any()
    for b in range(theta.shape[0]):
        pars = dict(zip(names, theta[b]))
        ref = oracle.compute_model(pars)
        assert_xi_close({n: models[n][b] for n in ref}, ref, f'vector {b}')
        ref_chi2 = oracle.chi2(pars)
        # 1e-6 absolute (BASELINE.json) for chi2 of the order of the number of bins
        assert abs(chi2[b] - ref_chi2) <= CHI2_ATOL * max(1.0, abs(ref_chi2) / 1e4), (chi2[b], ref_chi2)
    # scalar API == batch of one
    assert abs(vega.chi2(dict(zip(names, theta[0]))) - chi2[0]) <= CHI2_ATOL


def test_dgemm_against_torch():
    import ctypes as C
    from vega_b200 import _lib
    lib = _lib.load()
    torch.manual_seed(1)
    # every tile height (8..64 in steps of 8, 3 x 32, 128), ragged edges, K % 32 == 16, split-K
    shapes = [(1, 130, 32, 1), (37, 1590, 2512, 1), (300, 257, 4096, 5), (1024, 814, 816, 1)]
    shapes += [(M, 200, 272, 1 + M % 3) for M in (5, 8, 9, 16, 17, 24, 25, 33, 41, 49, 57, 64, 65, 96, 97, 129, 130, 192, 193, 513)]
    for (M, N, K, splits) in shapes:
        A = torch.randn(M, K, device='cuda', dtype=torch.float64)
        B = torch.randn(N, K, device='cuda', dtype=torch.float64)
        ldc = N + (N & 1)
        Cm = torch.zeros(M, ldc, device='cuda', dtype=torch.float64)
        ws = torch.zeros(max(splits, 1), M, ldc, device='cuda', dtype=torch.float64)
        _lib.check(lib.vg_dgemm_tn(A.data_ptr(), B.data_ptr(), Cm.data_ptr(), M, N, K, K, K, ldc, 1.0, 0.0,
                                   splits, ws.data_ptr(), None))
        torch.cuda.synchronize()
        ref = A @ B.T
        assert (Cm[:, :N] - ref).abs().max() <= 1e-12 * ref.abs().max()


def test_moment_form_matches_per_point_pass():
    """csrc/vg_moments.cuh: knot tables as a 3/9-term combination of static tables must equal
    the per-point P(k,mu) pass + Hankel GEMM (exact algebra, only the summation order
    differs) -- and both must match the oracle."""
    from vega_b200 import _lib
    vega, oracle, theta = make_c2(4, width=1.0)
    names = list(vega.sample_params['limits'])
    engine = vega._engine_for()
    assert sum(engine.n_moment(ci) for ci in range(len(engine.infos))) > 0
    chi2_mom = vega.chi2_batch(theta)
    models_mom, _ = vega.compute_model_batch(theta)
    engine.set_option(_lib.OPT_MOMENTS, 0)
    assert sum(engine.n_moment(ci) for ci in range(len(engine.infos))) == 0
    chi2_full = vega.chi2_batch(theta)
    models_full, _ = vega.compute_model_batch(theta)
    engine.set_option(_lib.OPT_MOMENTS, 1)
    for name in models_full:
        scale = np.max(np.abs(models_full[name]))
        assert np.max(np.abs(models_mom[name] - models_full[name])) <= 1e-12 * scale
    assert np.max(np.abs(chi2_mom - chi2_full) / np.maximum(1.0, np.abs(chi2_full) / 1e4)) <= CHI2_ATOL
    ref = oracle.compute_model(dict(zip(names, theta[0])))
    assert_xi_close({n: models_mom[n][0] for n in ref}, ref, 'moment form')


def test_moment_form_with_the_uv_bias_term_matches_per_point_pass():
    """The scale-dependent UV bias (vega/power_spectrum.py:224-261) at FIXED UV parameters is a static
    function of k: five more moment terms (G, G F, G^2, G mu^2, G F mu^2).  Every evaluation of the
    reference's 4-correlation configuration (UVB-fluctuations on) is then in moment form; same model
    as the per-point pass and as the oracle."""
    from vega_b200 import _lib
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    names = ['bias_eta_LYA', 'beta_LYA', 'bias_hcd', 'beta_hcd', 'ap', 'at', 'bias_eta_SiII(1260)']
    names = [n for n in names if n in vega.params]
    base = np.array([vega.params[n] for n in names])
    theta = base * (1 + 0.05 * np.random.default_rng(5).standard_normal((12, len(names))))
    models_mom, _ = vega.compute_model_batch(theta, names)
    engine = vega._engine_for(tuple(names))
    n_mom = [engine.n_moment(ci) for ci in range(len(engine.infos))]
    for ci, info in enumerate(engine.infos):
        own = [d for d in info['evals'] if d.table_owner < 0 and not d.frozen and not d.direct]
        assert any(d.uvb and (d.uv_on[0] or d.uv_on[1]) and d.uv_static for d in own)
        assert n_mom[ci] == len(own), (info['item'].name, n_mom[ci], len(own))
    engine.set_option(_lib.OPT_MOMENTS, 0)
    models_pp, _ = vega.compute_model_batch(theta, names)
    engine.set_option(_lib.OPT_MOMENTS, 1)
    for name in models_pp:
        scale = np.max(np.abs(models_pp[name]))
        assert np.max(np.abs(models_mom[name] - models_pp[name])) <= 1e-12 * scale, name
    for b in (0, 11):
        ref = oracle.compute_model(dict(zip(names, theta[b])))
        assert_xi_close({n: models_mom[n][b] for n in ref}, ref, f'uv moment form, row {b}')


def test_whitened_chi2_matches_the_product_with_the_inverse_covariance():
    """VG_OPT_WHITEN: with C^-1 = W^T W (vg_corr_set_whitening) the chi2 of
    vega/vega_interface.py:294-319 is |W d - W m|^2 with W folded into the masked distortion rows, the
    fused metal operator, the data vector and the mocks.  Same values as the product with C^-1 (option
    0), for large and small batches, and as the oracle."""
    from vega_b200 import _lib
    vega, oracle, theta = make_c2(48, width=1.0)
    names = list(vega.sample_params['limits'])
    engine = vega._engine_for()
    got = vega.chi2_batch(theta)
    assert all(engine.whitened(ci) for ci in range(len(engine.infos)))
    small = vega.chi2_batch(theta[:3])
    engine.set_option(_lib.OPT_WHITEN, 0)
    ref = vega.chi2_batch(theta)
    assert not any(engine.whitened(ci) for ci in range(len(engine.infos)))
    ref_small = vega.chi2_batch(theta[:3])
    engine.set_option(_lib.OPT_WHITEN, 1)
    assert np.max(np.abs(got - ref) / np.abs(ref)) <= 1e-10
    assert np.max(np.abs(small - ref_small) / np.abs(ref_small)) <= 1e-10
    np.testing.assert_allclose(vega.chi2_batch(theta), got, rtol=1e-13)      # rebuilt after the option flip
    for b in (0, 47):
        want = oracle.chi2(dict(zip(names, theta[b])))
        assert abs(got[b] - want) <= CHI2_ATOL * max(1.0, abs(want) / 1e4), (b, got[b], want)


@pytest.mark.parametrize('case', ['c2_dense', 'c2_fast_metals', 'rp_only'])
def test_folded_metals_match_the_matrix_product(case, tmp_path):
    """VG_OPT_FOLD_METALS: metal terms whose xi stage is in moment form are served from static
    vectors Y_m = M_c X_m instead of the per-vector matrix product (vega/metals.py:350-358).
    Same model and chi2 as with the option off, and as the oracle."""
    from vega_b200 import _lib
    if case == 'rp_only':
        from helpers import make_new_metals_config
        from oracle.vega_oracle import Oracle
        from helpers import static_tables
        from vega_b200 import VegaInterface
        vega = VegaInterface(make_new_metals_config(tmp_path, True))
        oracle = Oracle(vega, tables=static_tables())
        names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_eta_SiII(1260)', 'bias_eta_SiIII(1207)']
        names = [n for n in names if n in vega.params]
        rng = np.random.default_rng(11)
        base = np.array([vega.params[n] for n in names])
        theta = base * (1 + 0.02 * rng.standard_normal((9, len(names))))
    else:
        vega, oracle, theta = make_c2(9, width=0.5, fast_metals=case == 'c2_fast_metals')
        names = list(vega.sample_params['limits'])
        if case == 'c2_fast_metals':
            oracle.compute_model()
    engine = vega._engine_for(names)
    n_folded = sum(engine.n_folded(ci) for ci in range(len(engine.infos)))
    assert n_folded > 0
    chi2_off = vega.chi2_batch(theta, names)
    models_off, _ = vega.compute_model_batch(theta, names)
    engine.set_option(_lib.OPT_FOLD_METALS, 1)
    try:
        chi2_on = vega.chi2_batch(theta, names)
        models_on, status = vega.compute_model_batch(theta, names)
    finally:
        engine.set_option(_lib.OPT_FOLD_METALS, 0)
    assert not status.any()
    for name in models_off:
        scale = np.max(np.abs(models_off[name]))
        assert np.max(np.abs(models_on[name] - models_off[name])) <= 1e-12 * scale, name
    assert np.max(np.abs(chi2_on - chi2_off) / np.maximum(1.0, np.abs(chi2_off) / 1e4)) <= CHI2_ATOL
    ref = oracle.compute_model(dict(zip(names, theta[0])))
    assert_xi_close({n: models_on[n][0] for n in ref}, ref, 'folded metals')


def test_c2_fast_metals_matches_oracle():
    """fast_metals on the dense C2 operands: frozen metal x metal vectors (M_c xi_c evaluated
    once at the plan parameters through the same kernels), shared metal x main tables."""
    vega, oracle, theta = make_c2(3, width=0.2, fast_metals=True)
    names = list(vega.sample_params['limits'])
    oracle.compute_model()          # the reference freezes its metal x metal cache at the first call
    chi2 = vega.chi2_batch(theta)
    models, status = vega.compute_model_batch(theta)
    assert not status.any()
    for b in range(theta.shape[0]):
        pars = dict(zip(names, theta[b]))
        ref = oracle.compute_model(pars)
        assert_xi_close({n: models[n][b] for n in ref}, ref, f'vector {b}')
        ref_chi2 = oracle.chi2(pars)
        assert abs(chi2[b] - ref_chi2) <= CHI2_ATOL * max(1.0, abs(ref_chi2) / 1e4), (chi2[b], ref_chi2)


def test_c3_shaped_four_correlations_global_covariance_fast_metals():
    """The ingredients of the DESI-shaped workload C3 on the reference's own 4-correlation
    configuration: dense distortion / metal matrices, fast_metals, post-distortion broadband
    polynomials, ONE global covariance over all masked bins (vega/vega_interface.py:295-304)."""
    from helpers import set_fast_metals
    from oracle.vega_oracle import Oracle
    from helpers import static_tables
    from vega_b200 import VegaInterface
    vega = VegaInterface('configs/full_configs/main.ini')
    set_fast_metals(vega)
    synthetic.inject_dense_operands(vega)
    synthetic.inject_global_covariance(vega)
    oracle = Oracle(vega, tables=static_tables())
    synthetic.draw_mock_data(vega, oracle.compute_model())
    names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'drp_QSO',
             'sigma_velo_disp_lorentz_QSO', 'bao_amp', 'bias_eta_SiII(1260)',
             'BB-lyalya_lyalya-0 add post r,mu (1,2)']
    rng = np.random.default_rng(8)
    base = np.array([vega.params[n] for n in names])
    theta = base * (1 + 0.03 * rng.standard_normal((3, len(names))))
    chi2 = vega.chi2_batch(theta, names)
    log_lik = vega.log_lik_batch(theta, names)
    for b in range(theta.shape[0]):
        pars = dict(zip(names, theta[b]))
        ref = oracle.chi2(pars)
        assert abs(chi2[b] - ref) <= CHI2_ATOL * max(1.0, abs(ref) / 1e4), (chi2[b], ref)
        assert abs(log_lik[b] - oracle.log_lik(pars)) <= CHI2_ATOL * max(1.0, abs(ref) / 1e4)


def test_chunked_batches_agree_with_single_vectors():
    """C4-shaped call: more vectors than one workspace chunk (ragged last chunk); every chi2
    must equal the batch-of-one result of the same vector, sentinel rows included."""
    vega, oracle = make_oracle('configs/configs/main.ini')
    vega._batch_chunk = 64
    names = list(vega.sample_params['limits'])
    theta = synthetic.draw_theta(vega, 150, seed=20261105, width=2.0)
    theta[17, names.index('ap')] = 1e6          # out of the FFTLog range -> 1e100
    chi2 = vega.chi2_batch(theta)
    assert chi2[17] == 1e100 and np.sum(chi2 == 1e100) == 1
    for b in (0, 63, 64, 127, 128, 149):
        one = vega.chi2(dict(zip(names, theta[b])))
        assert abs(one - chi2[b]) <= 1e-9 * max(1.0, abs(one))
    assert np.allclose(vega.log_lik_batch(theta[:3]), [vega.log_lik(dict(zip(names, t))) for t in theta[:3]],
                       rtol=0, atol=1e-8)


@pytest.mark.parametrize('los', [None, 'growth', 'only-los'])
def test_mock_bin_size_smoothing(los):
    """vega/power_spectrum.py:143-160: the extra G(k) of mocks generated on a coarse grid."""
    from oracle.vega_oracle import Oracle
    from helpers import static_tables
    from vega_b200 import VegaInterface
    vega = VegaInterface('configs/configs/main.ini')
    for item in vega.corr_items.values():
        item.config['model']['mock-bin-size'] = '2.19'
        if los is not None:
            item.config['model']['mock-los-smoothing'] = los
    oracle = Oracle(vega, tables=static_tables())
    assert_xi_close(vega.compute_model(), oracle.compute_model(), f'mock-bin-size {los}')
    pars = {'ap': 1.02, 'beta_LYA': 1.5}
    assert_xi_close(vega.compute_model(pars), oracle.compute_model(pars), f'mock-bin-size {los}')
    assert abs(vega.chi2(pars) - oracle.chi2(pars)) <= CHI2_ATOL


def test_desi_shaped_model_grid_larger_than_the_data_grid(tmp_path):
    """The distortion-matrix file carries its own grid (vega/data.py:441-473): the model and
    the distorted model live on 55x55 bins out to 220 Mpc/h while the data stay on 50x50 bins
    out to 200 -- the DESI layout (5625/11250 model bins for 2500/5000 data bins), reduced.
    The masks must pick the same bins from both grids and chi2 only needs the masked rows."""
    import configparser
    from oracle.vega_oracle import Oracle
    from helpers import static_tables
    from vega_b200 import VegaInterface
    n_p = n_t = 55
    size = 4.0
    rp = (np.arange(n_p) + 0.5) * size
    rt = (np.arange(n_t) + 0.5) * size
    rp_grid, rt_grid = [g.ravel() for g in np.meshgrid(rp, rt, indexing='ij')]
    dmat = synthetic.distortion_matrix(n_p, n_t)
    dmat_path = tmp_path / 'dmat_55.fits'
    fitsio.write_tables(dmat_path, [
        ('DMAT', {'RPMIN': 0., 'RPMAX': n_p * size, 'RTMAX': n_t * size, 'NP': n_p, 'NT': n_t,
                  'COEFMOD': 1}, {'DM': dmat}),
        ('ATTRI', {}, {'RP': rp_grid, 'RT': rt_grid, 'Z': np.full(rp_grid.size, 2.33)})])
    corr = configparser.ConfigParser()
    corr.optionxform = lambda o: o
    corr.read(utils.find_file('configs/configs/lya_lyaxlya_lya.ini'))
    corr['data']['distortion-file'] = str(dmat_path)
    corr_path = tmp_path / 'lyaxlya.ini'
    with open(corr_path, 'w') as f:
        corr.write(f)
    main = configparser.ConfigParser()
    main.optionxform = lambda o: o
    main.read(utils.find_file('configs/configs/main.ini'))
    main['data sets']['ini files'] = str(corr_path)
    main_path = tmp_path / 'main.ini'
    with open(main_path, 'w') as f:
        main.write(f)

    vega = VegaInterface(str(main_path))
    data = vega.data['lyaxlya']
    assert data.model_coordinates.r_grid.size == n_p * n_t and data.full_data_size == 2500
    assert int(data.model_mask.sum()) == int(data.data_mask.sum())
    synthetic.inject_dense_operands(vega, with_distortion=False, with_metals=False)   # covariance only
    oracle = Oracle(vega, tables=static_tables())
    synthetic.draw_mock_data(vega, oracle.compute_model())
    names = list(vega.sample_params['limits'])
    theta = synthetic.draw_theta(vega, 3, width=0.3)
    models, status = vega.compute_model_batch(theta)
    chi2 = vega.chi2_batch(theta)
    assert not status.any() and models['lyaxlya'].shape == (3, n_p * n_t)
    for b in range(3):
        pars = dict(zip(names, theta[b]))
        ref = oracle.compute_model(pars)
        assert_xi_close({n: models[n][b] for n in ref}, ref, f'vector {b}')
        ref_chi2 = oracle.chi2(pars)
        assert abs(chi2[b] - ref_chi2) <= CHI2_ATOL * max(1.0, abs(ref_chi2) / 1e4), (chi2[b], ref_chi2)


def test_c3_full_size_desi_shaped(tmp_path):
    """C3 at its full size: four correlations, data on 2500/5000 bins, model and
    distorted model on the distortion matrices' own 5625/11250-bin grids (dense 253 MB / 1 GB
    matrices), new_metals rp-only matrices built on the device on those grids, fast_metals
    (frozen metal x metal terms, shared metal x main tables), post-distortion broadband, ONE
    global covariance over the 9540 masked bins.  A gradient stencil of 2P+1 vectors -- the
    batch an iminuit step consumes -- against the oracle on the same operands."""
    from helpers import make_c3_config, static_tables
    from oracle.vega_oracle import Oracle
    from vega_b200 import VegaInterface
    from vega_b200.drivers import stencil
    vega = VegaInterface(make_c3_config(tmp_path))
    sizes = {n: (d.model_coordinates.r_grid.size, d.full_data_size) for n, d in vega.data.items()}
    assert sorted(sizes.values()) == [(5625, 2500), (5625, 2500), (11250, 5000), (11250, 5000)]
    synthetic.inject_dense_operands(vega, with_distortion=False, with_metals=False)   # covariances
    synthetic.inject_global_covariance(vega)
    oracle = Oracle(vega, tables=static_tables(), fast_metals=True)
    synthetic.draw_mock_data(vega, oracle.compute_model())
    names = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'beta_hcd', 'drp_QSO',
             'sigma_velo_disp_lorentz_QSO', 'bao_amp', 'bias_eta_SiII(1260)',
             'BB-lyalya_lyalya-0 add post r,mu (1,2)']
    base = np.array([vega.params[n] for n in names])
    theta = stencil(base, 0.02 * np.abs(base) + 1e-4, hessian=False)
    assert theta.shape[0] == 2 * len(names) + 1
    chi2 = vega.chi2_batch(theta, names)
    engine = vega._engine_for(tuple(names))
    assert sum(engine.n_static_xi(ci) for ci in range(4)) > 0
    for b in (0, 3, 2 * len(names)):
        ref = oracle.chi2(dict(zip(names, theta[b])))
        assert abs(chi2[b] - ref) <= CHI2_ATOL * max(1.0, abs(ref) / 1e4), (b, chi2[b], ref)
    # the same vectors in a larger batch (different CTA scheduling, split-K choices and chunking):
    # results do not depend on the batch composition
    big = vega.chi2_batch(np.tile(theta, (5, 1)), names)
    assert np.max(np.abs(big.reshape(5, -1) - chi2[None, :])) <= 1e-7
    model = vega.compute_model(dict(zip(names, theta[3])))
    assert_xi_close(model, oracle.compute_model(dict(zip(names, theta[3]))), 'C3 stencil point')


def stage_objects_match_the_oracle(vega, oracle, pars):
    """The reference's per-stage call shapes on CALLER-SUPPLIED arrays:
    PowerSpectrum.compute(pk_lin, params, fast_metals) (vega/power_spectrum.py:87),
    PktoXi.compute_pk_ells(pk) / compute(r_grid, mu_grid, pk, single_ell) / init_from_Pk
    (vega/pktoxi.py:61-163) and CorrelationFunction.compute(pk, pk_lin, PktoXi_obj, params)
    (vega/correlation_func.py:117), each against the oracle's stage of the same name."""
    from oracle.vega_oracle import pk_model
    from vega_b200.pktoxi import PktoXi
    fid = vega.fiducial
    lp = oracle.local_params(pars)
    pk_lin = fid['pk_smooth'] * (1 + 0.05 * np.cos(3 * np.log(fid['k'])))        # not one of the templates
    for name in ('lyalya_lyalya', 'lyalya_qso'):
        model, cm = vega.models[name], oracle.models[name]
        for peak in (True, False):
            lp['peak'] = peak
            for fast in (False, True):
                pk_ref = pk_model(cm.grid, pk_lin, lp, fast_metals=fast)
                pk_got = model.Pk_core.compute(pk_lin, lp, fast_metals=fast)
                assert pk_got.shape == pk_ref.shape
                # (without b1 b2 the Kaiser term of a tracer is divided by its effective bias, which
                # crosses zero at some k once the UV term is on: the grid points next to that pole carry
                # the cancellation error of BOTH sides and are left out of the comparison)
                regular = np.abs(pk_ref) <= 1e3 * np.median(np.abs(pk_ref[:, :400]))
                assert regular.mean() > 0.99
                assert np.max(np.abs(pk_got - pk_ref)[regular]) <= 1e-12 * np.max(np.abs(pk_ref[regular])), \
                    (name, peak, fast)
            pk_ref = pk_model(cm.grid, pk_lin, lp)
            # the transform object: multipoles, xi on arbitrary coordinates, one multipole alone
            pktoxi = model.PktoXi
            ells = pktoxi.compute_pk_ells(pk_ref)
            ells_ref = cm.transform.multipoles(pk_ref)
            assert np.max(np.abs(ells - ells_ref)) <= 1e-12 * np.max(np.abs(ells_ref))
            r = np.exp(np.linspace(np.log(3.), np.log(290.), 777))
            r[5] = 0.
            mu = np.linspace(-1, 1, 777)
            xi = pktoxi.compute(r, mu, pk_ref)
            xi_ref = cm.transform.xi(r, mu, pk_ref)
            assert xi[5] == 0. and np.max(np.abs(xi - xi_ref)) <= XI_RTOL * np.max(np.abs(xi_ref))
            xi2 = pktoxi.compute(r, mu, pk_ref, single_ell=2)
            assert np.max(np.abs(xi2 - cm.transform.xi(r, mu, pk_ref, 2))) <= XI_RTOL * np.max(np.abs(xi_ref))
            with pytest.raises(utils_errors().VegaBoundsError):
                pktoxi.compute(np.array([1e-6, 10.]), np.array([0.5, 0.5]), pk_ref)
            # CorrelationFunction.compute on the caller's P(k,mu)
            xi_got = model.Xi_core.compute(pk_ref, fid['pk_smooth'], pktoxi, lp)
            xi_want = cm.xi_core.compute(pk_ref, fid['pk_smooth'], cm.transform, lp)
            assert np.max(np.abs(xi_got - xi_want)) <= XI_RTOL * np.max(np.abs(xi_want)), (name, peak)
        # a metal correlation's objects
        pair = next(iter(model.metals.Pk_metal))
        g, tr, xim, mat = cm.metal[pair]
        lp['peak'] = False
        pk_ref = pk_model(g, fid['pk_full'], lp, fast_metals=True)
        pk_got = model.metals.Pk_metal[pair].compute(fid['pk_full'], lp, fast_metals=True)
        assert np.max(np.abs(pk_got - pk_ref)) <= 1e-12 * np.max(np.abs(pk_ref)), (name, pair)
        xi_got = model.metals.Xi_metal[pair].compute(pk_ref, fid['pk_full'], model.metals.PktoXi[pair], lp)
        xi_want = xim.compute(pk_ref, fid['pk_full'], tr, lp)
        assert np.max(np.abs(xi_got - xi_want)) <= XI_RTOL * np.max(np.abs(xi_want)), (name, pair)
    # a stand-alone transform object (bin/make_template-style use): PktoXi.init_from_Pk
    model, cm = vega.models['lyalya_lyalya'], oracle.models['lyalya_lyalya']
    alone = PktoXi.init_from_Pk(model.Pk_core, vega.corr_items['lyalya_lyalya'].config['model'])
    lp['peak'] = False
    pk_ref = pk_model(cm.grid, pk_lin, lp)
    r, mu = np.linspace(5., 250., 300), np.full(300, 0.3)
    assert np.max(np.abs(alone.compute(r, mu, pk_ref) - cm.transform.xi(r, mu, pk_ref))) <= \
        XI_RTOL * np.max(np.abs(cm.transform.xi(r, mu, pk_ref)))
    ells_ref = cm.transform.multipoles(pk_ref)
    assert np.max(np.abs(alone.compute_pk_ells(pk_ref) - ells_ref)) <= 1e-12 * np.max(np.abs(ells_ref))


def utils_errors():
    from vega_b200 import utils
    return utils


def test_model_call_shapes_and_stage_components():
    """VegaInterface.models[name]: Model.compute / compute_direct (vega/model.py:157-208) and the
    per-stage components against the oracle's own stages (PowerSpectrum.compute,
    PktoXi.compute_pk_ells, CorrelationFunction.compute)."""
    from oracle.vega_oracle import pk_model
    vega, oracle = make_oracle('configs/full_configs/main.ini')
    fid = vega.fiducial
    pars = dict(vega.params)
    pars.update({'bias_eta_LYA': -0.21, 'beta_LYA': 1.5, 'ap': 0.98, 'sigma_velo_disp_lorentz_QSO': 5.0})
    ref = oracle.compute_model({k: pars[k] for k in ('bias_eta_LYA', 'beta_LYA', 'ap', 'sigma_velo_disp_lorentz_QSO')})
    for name in ('lyalya_lyalya', 'lyalya_qso'):
        model = vega.models[name]
        got = model.compute(pars, fid['pk_full'], fid['pk_smooth'])
        assert_xi_close({name: got}, {name: ref[name]}, 'Model.compute')
        # other templates than the fiducial file's (vega/model.py:157-187 takes them as arguments)
        tilt = 1 + 0.03 * np.sin(2 * np.log(fid['k']))
        got = model.compute(pars, fid['pk_full'] * tilt, fid['pk_smooth'] * tilt**0.5)
        want = oracle.models[name].compute(oracle.local_params(pars), fid['pk_full'] * tilt, fid['pk_smooth'] * tilt**0.5)
        assert_xi_close({name: got}, {name: want}, 'Model.compute, other templates')
        again = vega.compute_model({k: pars[k] for k in ('bias_eta_LYA', 'beta_LYA', 'ap', 'sigma_velo_disp_lorentz_QSO')})
        assert_xi_close({name: again[name]}, {name: ref[name]}, 'back on the fiducial templates')
        cm = oracle.models[name]
        lp = oracle.local_params(pars)
        for component, pk_lin, peak in (('peak', fid['pk_full'] - fid['pk_smooth'], True),
                                        ('smooth', fid['pk_smooth'], False)):
            lp['peak'] = peak
            pk_ref = pk_model(cm.grid, pk_lin, lp)
            pk_got = model.pk_grid(pars, component)
            assert np.max(np.abs(pk_got - pk_ref)) <= 1e-12 * np.max(np.abs(pk_ref)), (name, component)
            ells_ref = cm.transform.multipoles(pk_ref)
            ells_got = model.pk_ells(pars, component)
            assert np.max(np.abs(ells_got - ells_ref)) <= 1e-11 * np.max(np.abs(ells_ref)), (name, component)
            xi_ref = cm.xi_core.compute(pk_ref, pk_lin, cm.transform, lp)
            xi_got = model.xi_of(pars, component)
            assert np.max(np.abs(xi_got - xi_ref)) <= XI_RTOL * np.max(np.abs(xi_ref)), (name, component)
    stage_objects_match_the_oracle(vega, oracle, pars)
    direct = fid['pk_full'] * (1 + 0.02 * np.cos(np.log(fid['k'])))
    got = vega.models['lyalya_qso'].compute_direct(pars, direct)
    want = oracle.compute_model({k: pars[k] for k in ('bias_eta_LYA', 'beta_LYA', 'ap', 'sigma_velo_disp_lorentz_QSO')},
                                direct_pk=direct)['lyalya_qso']
    assert np.max(np.abs(got - want)) <= XI_RTOL * np.max(np.abs(want))


@pytest.mark.gpu
def test_device_covariance_inverse_and_log_det():
    """Init-time masked C^-1 and log|C| of a DESI-sized covariance on the device (one Cholesky
    factor, vega_b200/utils.py) against the host LAPACK route the reference takes
    (vega/utils.py:271-318), and the chi2-like quadratic form both give."""
    from vega_b200 import utils
    rng = np.random.default_rng(11)
    n_all, n = 5200, 4600
    idx = np.arange(n_all)
    cov = 0.4 ** (np.abs(idx[:, None] - idx[None, :]) / 3.0) * np.outer(1 + 0.1 * rng.random(n_all),
                                                                         1 + 0.1 * rng.random(n_all))
    cov = 0.5 * (cov + cov.T) * 1e-8 + 1e-9 * np.eye(n_all)
    mask = np.zeros(n_all, dtype=bool)
    mask[rng.choice(n_all, n, replace=False)] = True
    assert n >= utils.DEVICE_LINALG_MIN
    got_inv = utils.compute_masked_invcov(cov, mask)
    got_det = utils.compute_log_cov_det(cov, mask)
    assert utils._factor_cache and next(iter(utils._factor_cache.values())) is not None
    masked = cov[:, mask][mask, :]
    ref_inv, ref_det = np.linalg.inv(masked), np.linalg.slogdet(masked)[1]
    assert np.max(np.abs(got_inv - ref_inv)) <= 1e-10 * np.max(np.abs(ref_inv))
    assert abs(got_det - ref_det) <= 1e-8
    r = rng.standard_normal(n) * 1e-4
    assert abs(r @ got_inv @ r - r @ ref_inv @ r) <= 1e-6


def test_device_marginalisation_matrices():
    """Init-time small-scale marginalisation at DESI size on the device (vega/data.py:96-128,
    762-828): SVD compression of the template covariance and the residual -> amplitudes matrix
    (T^T C^-1 T + S)^-1 T^T C^-1, against the host LAPACK route the reference takes."""
    from vega_b200 import utils
    rng = np.random.default_rng(12)
    n, nt = 4300, 60
    assert n >= utils.DEVICE_LINALG_MIN
    r = np.linspace(0, 1, n)
    t = np.stack([np.exp(-0.5 * ((r - c) / 0.01)**2) for c in rng.random(nt - 4)] +
                 [np.exp(-0.5 * ((r - 0.5) / 0.01)**2)] * 4, axis=1) * 5.0        # 4 degenerate templates
    got, modes = utils.template_covariance(t)
    u, sv, _ = np.linalg.svd(t, full_matrices=False)
    keep = sv > 1e-8 * sv[0]
    ref = np.dot(u[:, keep] * sv[keep]**2, u[:, keep].T)
    assert modes == int(keep.sum()) and nt - 8 <= modes <= nt - 3      # the 4 equal templates are one mode
    assert np.max(np.abs(got - ref)) <= 1e-10 * np.max(np.abs(ref))
    idx = np.arange(n)
    cov = 0.3 ** (np.abs(idx[:, None] - idx[None, :]) / 2.0) * 1e-8 + 1e-9 * np.eye(n)
    inv_cov = np.linalg.inv(cov)
    t_u = np.stack([np.exp(-0.5 * ((r - c) / 0.01)**2) for c in np.linspace(0.05, 0.95, 40)], axis=1) * 5.0
    prior = np.full(40, 5.0**-2)
    got = utils.diff_to_coeff_matrix(t_u, inv_cov, prior)
    g = t_u.T @ inv_cov
    ref = np.linalg.inv(g @ t_u + np.diag(prior)).dot(g)
    assert np.max(np.abs(got - ref)) <= 1e-8 * np.max(np.abs(ref))
    diff = rng.standard_normal(n) * 1e-4
    assert np.max(np.abs(got @ diff - ref @ diff)) <= 1e-10 * np.max(np.abs(ref @ diff))


@pytest.mark.gpu
def test_component_tables_of_write_pk_and_write_cf(tmp_path):
    """The save-components dictionaries (vega/model.py:113-116,151-153) and the PK_ / Xi_ tables
    write_pk / write_cf add to the result file (vega/output.py:100-110,351-440), against the
    oracle's per-stage outputs; peak and smooth recombine to the model."""
    import configparser
    from vega_b200 import VegaInterface, fitsio, utils
    cfg = configparser.ConfigParser()
    cfg.optionxform = lambda option: option
    cfg.read(utils.find_file('configs/full_configs/main.ini'))
    cfg['output'] = {'filename': str(tmp_path / 'components'), 'write_cf': 'True', 'write_pk': 'True'}
    path = tmp_path / 'main.ini'
    with open(path, 'w') as f:
        cfg.write(f)
    vega = VegaInterface(str(path))
    pars = {'bias_eta_LYA': -0.21, 'beta_LYA': 1.55, 'ap': 1.01, 'bao_amp': 0.9}
    model = vega.compute_model(pars)
    out = vega.output.write_results(model, vega._get_lcl_prms(pars), models=vega.models)
    hdul = fitsio.open(out)
    by_name = {h.header.get('EXTNAME'): h for h in hdul[1:]}
    for name in vega.corr_items:
        m = vega.models[name]
        pk, xi = by_name['PK_' + name], by_name['Xi_' + name]
        assert np.array_equal(pk.data['peak_core'], m.pk['peak']['core'])
        assert np.array_equal(xi.data['raw_smooth_core'], m.xi['smooth']['core'])
        assert np.array_equal(xi.data['distorted_peak_core'], m.xi_distorted['peak']['core'])
        # bao_amp * peak + smooth is the model (vega/model.py:186)
        total = 0.9 * m.xi_distorted['peak']['core'] + m.xi_distorted['smooth']['core']
        assert np.max(np.abs(total - model[name])) <= 1e-13 * np.max(np.abs(model[name]))
        # the same arrays as the per-stage accessors (parity-tested against the oracle's stages above;
        # the distorted pair is pinned against the unmodified reference by the sensitivity golden)
        assert np.array_equal(m.xi['peak']['core'], m.xi_of(pars, 'peak'))
        assert pk.data['smooth_core'].shape == m.pk['smooth']['core'].shape == (1000, vega.fiducial['k'].size)


def test_small_batches_replay_a_cuda_graph():
    """chi2 passes of <= 96 vectors: two ordinary calls per batch size, then a captured CUDA graph
    on fixed staging buffers.  Replays with NEW parameter vectors must give what the ordinary path
    gives (the 1024-vector path never takes a graph), also after a larger batch has re-allocated
    work buffers, after an option changed, and in Monte Carlo mode."""
    from vega_b200 import _lib
    vega, oracle, theta = make_c2(1100, width=1.0)
    engine = vega._engine_for()
    lib = engine.lib
    ref = vega.chi2_batch(theta)                      # chunks of 1024 + 76: the tail goes through chi2_chunk_small
    for size in (1, 33, 76):
        for rep in range(5):
            rows = np.arange(rep * size, (rep + 1) * size)
            got = vega.chi2_batch(theta[rows])
            assert np.max(np.abs(got - ref[rows])) <= 1e-9 * np.max(np.abs(ref[rows])), (size, rep)
    assert lib.vg_plan_graph_replays(engine.handle) >= 6
    n0 = lib.vg_plan_graph_replays(engine.handle)
    vega.chi2_batch(theta)                            # a large batch in between
    got = vega.chi2_batch(theta[:33])
    assert np.max(np.abs(got - ref[:33])) <= 1e-9 * np.max(np.abs(ref[:33]))
    engine.set_option(_lib.OPT_FOLD_METALS, 1)        # an option drops the graphs
    a = vega.chi2_batch(theta[:33])
    engine.set_option(_lib.OPT_FOLD_METALS, 0)
    for _ in range(4):
        b = vega.chi2_batch(theta[:33])
    assert np.max(np.abs(a - ref[:33])) <= 1e-9 * np.max(np.abs(ref[:33]))
    assert np.max(np.abs(b - ref[:33])) <= 1e-9 * np.max(np.abs(ref[:33]))
    assert lib.vg_plan_graph_replays(engine.handle) > n0
    # the scalar API against the oracle after the graph is in place
    names = vega.theta_names
    for i in range(4):
        value = vega.chi2(dict(zip(names, theta[i])))
    want = oracle.chi2(dict(zip(names, theta[3])))
    assert abs(value - want) <= 1e-6 * max(1.0, abs(want) / 1e4)


def test_second_lane_of_thin_batches_changes_nothing():
    """VG_OPT_LANES: the correlations of a chunk are spread over up to four streams, each with its own
    workspace (digit planes of the split-integer GEMMs included: the 100-vector case).  Same chi2 as with every correlation on one stream -- two correlations with their
    own covariances (C2), four under one global covariance, first calls (ordinary launches) and
    CUDA-graph replays alike."""
    from helpers import set_fast_metals
    from vega_b200 import _lib, VegaInterface
    vega, oracle, theta = make_c2(104, width=1.0)
    cases = [(vega, None, theta)]
    vega4 = VegaInterface('configs/full_configs/main.ini')
    set_fast_metals(vega4)
    synthetic.inject_dense_operands(vega4)
    synthetic.inject_global_covariance(vega4)
    synthetic.draw_mock_data(vega4, vega4.compute_model())
    names4 = ['bias_eta_LYA', 'beta_LYA', 'ap', 'at', 'bias_hcd', 'drp_QSO', 'bao_amp', 'bias_eta_SiII(1260)']
    base = np.array([vega4.params[n] for n in names4])
    cases.append((vega4, names4, base * (1 + 0.03 * np.random.default_rng(3).standard_normal((104, len(names4))))))
    for v, names, th in cases:
        engine = v._engine_for(tuple(names)) if names else v._engine_for()
        for size in (1, 7, 33, 100):
            out = {}
            for lanes in (1, 0, 1):
                engine.set_option(_lib.OPT_LANES, lanes)
                for rep in range(4):                    # two ordinary passes, the capture, a replay
                    rows = np.arange(rep, rep + size)
                    got = v.chi2_batch(th[rows], names) if names else v.chi2_batch(th[rows])
                    out.setdefault(rep, []).append(got)
            for rep, (a, b, c) in out.items():
                assert np.max(np.abs(a - b) / np.abs(b)) <= 1e-14, (size, rep)
                assert np.array_equal(a, c), (size, rep)
