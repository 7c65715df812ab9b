"""Split-integer GEMM on tcgen05 (csrc/vg_ozaki.cuh) against FP64 matmul, through the C ABI."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def run(A, B, slices):
    import torch
    from vega_b200 import _lib
    lib = _lib.load()
    a = torch.from_numpy(np.ascontiguousarray(A)).cuda()
    b = torch.from_numpy(np.ascontiguousarray(B)).cuda()
    M, K = A.shape
    N = B.shape[0]
    ldc = (N + 15) // 16 * 16
    c = torch.zeros(M, ldc, dtype=torch.float64, device='cuda')
    err = C.c_int32(0)
    _lib.check(lib.vg_ozaki_gemm(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(c.data_ptr()),
                                 M, N, K, K, K, ldc, slices, C.byref(err), None))
    assert err.value == 0, 'a pipeline wait of ozaki_gemm_kernel timed out'
    ref = torch.matmul(a, b.T).cpu().numpy()
    return c.cpu().numpy()[:, :N], ref


def bound(A, B, slices):
    """(S + 1) K 2^(-7S) max|A_row| max|B_row| plus the FP64 rounding of the recombination."""
    K = A.shape[1]
    am, bm = np.max(np.abs(A), axis=1), np.max(np.abs(B), axis=1)
    return ((slices + 1) * K * 2.0**(-7 * slices) + 8 * 2.0**-53) * np.outer(am, bm)


@pytest.mark.parametrize('shape', [(128, 64, 64), (128, 64, 128), (96, 40, 208), (300, 200, 1040), (257, 129, 4112)])
def test_small_integers_are_exact(shape):
    """Operands that are single radix-128 digits times a row scale: every product is exact."""
    M, N, K = shape
    rng = np.random.default_rng(5)
    A = rng.integers(-64, 65, size=(M, K)).astype(float) * 2.0**rng.integers(-20, 20, size=(M, 1))
    B = rng.integers(-64, 65, size=(N, K)).astype(float) * 2.0**rng.integers(-20, 20, size=(N, 1))
    A[:, 0] = np.sign(A[:, 0] + 0.5) * 64 * 2.0**np.floor(np.log2(np.max(np.abs(A), axis=1) / 64 + 1e-300))
    got, ref = run(A, B, 3)
    exact = A @ B.T          # sums of < 2^31 integers times powers of two: exact in FP64 as well
    assert np.array_equal(ref, exact)
    assert np.array_equal(got, exact)


@pytest.mark.parametrize('slices', [5, 6, 7, 8])
@pytest.mark.parametrize('shape', [(128, 64, 64), (200, 100, 1000 // 16 * 16), (1024, 2500, 37680)])
def test_truncation_bound(slices, shape):
    M, N, K = shape
    rng = np.random.default_rng(slices)
    # correlation-function-like rows: a smooth envelope over four decades, signs mixed
    A = rng.standard_normal((M, K)) * np.exp(-9.2 * rng.random((1, K))) * np.exp(rng.uniform(-3, 3, (M, 1)))
    B = np.abs(rng.standard_normal((N, K))) * (rng.random((N, K)) < 0.05) * np.exp(rng.uniform(-2, 2, (N, 1)))
    got, ref = run(A, B, slices)
    err = np.abs(got - ref)
    assert np.all(err <= bound(A, B, slices) + 4 * 2.0**-53 * np.abs(A) @ np.abs(B).T), \
        (slices, float(np.max(err / bound(A, B, slices))))


def test_zero_rows_and_nonfinite_rows():
    rng = np.random.default_rng(0)
    A = rng.standard_normal((130, 256))
    B = rng.standard_normal((70, 256))
    A[3] = 0.0
    B[5] = 0.0
    A[7, 11] = np.nan
    got, ref = run(A, B, 7)
    assert np.all(got[3] == 0.0) and np.all(got[:, 5][np.arange(130) != 7] == 0.0)
    assert np.all(np.isnan(got[7]))
    ok = np.arange(130) != 7
    assert np.max(np.abs(got[ok] - ref[ok])) <= np.max(bound(A[ok], B, 7)) + 1e-13


def test_metal_gemm_option_matches_dmma():
    """VG_OPT_OZAKI on the dense C2 workload (metal matrices AND the distortion matrix, VG_OPT_OZAKI_DM
    = 1, the default): chi2 and the model against the FP64 DMMA path (option 0), for the automatic
    digit count (the default) and for fixed ones; then the metal product alone (VG_OPT_OZAKI_DM = 0)."""
    from helpers import make_c2
    from vega_b200 import _lib
    vega, oracle, theta = make_c2(64, width=1.0)
    engine = vega._engine_for()
    engine.set_option(_lib.OPT_OZAKI, 0)
    ref_model, _ = vega.compute_model_batch(theta)
    ref = vega.chi2_batch(theta)
    for slices in (-1, 6, 7, 8):
        engine.set_option(_lib.OPT_OZAKI, slices)
        for ci in range(len(engine.infos)):
            digits, bounds = engine.ozaki_slices(ci)
            assert 3 <= digits <= 8
            if slices == -1:        # the automatic choice is the fewest digits under the budget
                assert bounds[digits] <= 5e-11 and (digits == 3 or bounds[digits - 1] > 5e-11)
            digits, bounds = engine.ozaki_dm_slices(ci)
            assert 3 <= digits <= 8
            if slices == -1:
                assert bounds[digits] <= 5e-11 and (digits == 3 or bounds[digits - 1] > 5e-11)
        got = vega.chi2_batch(theta)
        model, status = vega.compute_model_batch(theta)
        assert not status.any()
        for name in ref_model:
            scale = np.max(np.abs(ref_model[name]))
            # fixed digit counts below the automatic ones (7 / 8 here) give up accuracy on purpose
            tol = {6: 2e-10, 7: 2e-11}.get(slices, 1e-11)
            assert np.max(np.abs(model[name] - ref_model[name])) <= tol * scale, (slices, name)
        # (the chi2 path applies the metal matrices through the fused distortion x metal operator, whose
        # own budget asks for 8 digits on the auto-correlation: forcing 6 or 7 gives up 128x per digit)
        tol_chi2 = {6: 1e-2, 7: 1e-4}.get(slices, 1e-6)
        err = np.max(np.abs(got - ref)) / max(1.0, np.max(np.abs(ref)) / 1e4)
        assert err <= tol_chi2, (slices, err)
    engine.set_option(_lib.OPT_OZAKI, -1)
    engine.set_option(_lib.OPT_OZAKI_DM, 0)
    assert all(engine.ozaki_dm_slices(ci)[0] == 0 for ci in range(len(engine.infos)))
    model, _ = vega.compute_model_batch(theta)
    for name in ref_model:
        assert np.max(np.abs(model[name] - ref_model[name])) <= 1e-11 * np.max(np.abs(ref_model[name])), name
    assert np.max(np.abs(vega.chi2_batch(theta) - ref)) <= 1e-6 * max(1.0, np.max(np.abs(ref)) / 1e4)
    engine.set_option(_lib.OPT_OZAKI_DM, 1)
    # the oracle itself, on the vectors whose chi2 is of the order of the degrees of freedom (taken
    # from the whole batch: fewer than 64 vectors would go to the DMMA kernel)
    got = vega.chi2_batch(theta)[:4]
    for b in range(4):
        want = oracle.chi2(dict(zip(vega.theta_names, theta[b])))
        assert abs(got[b] - want) <= 1e-6 * max(1.0, abs(want) / 1e4)


@pytest.mark.gpu
def test_fused_distortion_metal_operator_matches_the_two_products():
    """VG_OPT_FUSE_DM_METALS: on the chi2 path the metal matrices are applied as (masked rows of the
    distortion matrix) x M_c, built at finalize (vega/model.py:142-143 with vega/metals.py:350-358:
    D (M_c xi_c) = (D M_c) xi_c).  Same chi2 as the two-step form -- with the split-integer GEMM and
    with the FP64 DMMA kernel, for a large and for a small batch -- and as the oracle."""
    from helpers import make_c2
    from vega_b200 import _lib
    vega, oracle, theta = make_c2(64, width=1.0)
    engine = vega._engine_for()
    fused = [engine.fused_metal_operator(ci) for ci in range(len(engine.infos))]
    for (rows, digits), info in zip(fused, engine.infos):
        assert rows == int(info['data'].model_mask.sum()) and 3 <= digits <= 8
    results = {}
    for ozaki in (-1, 0):
        engine.set_option(_lib.OPT_OZAKI, ozaki)
        for fuse in (1, 0):
            engine.set_option(_lib.OPT_FUSE_DM_METALS, fuse)
            results[ozaki, fuse] = (vega.chi2_batch(theta), vega.chi2_batch(theta[:5]))
    engine.set_option(_lib.OPT_OZAKI, -1)
    engine.set_option(_lib.OPT_FUSE_DM_METALS, 1)
    ref_big, ref_small = results[0, 0]                     # FP64 DMMA, two products
    for key, (big, small) in results.items():
        assert np.max(np.abs(big - ref_big) / np.abs(ref_big)) <= 1e-10, key
        assert np.max(np.abs(small - ref_small) / np.abs(ref_small)) <= 1e-10, key
    got = results[-1, 1][0]
    for b in (0, 33, 63):
        want = oracle.chi2(dict(zip(vega.theta_names, theta[b])))
        assert abs(got[b] - want) <= 1e-6 * max(1.0, abs(want) / 1e4), (b, got[b], want)
