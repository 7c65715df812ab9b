"""vg_dgemm_tn_static: the split-K FP64 GEMM whose partials live in a file-scope device array, so the
call allocates nothing (include/vega_b200.h; no counterpart in the reference).
Collected last on purpose.

Checked through the C ABI against vg_dgemm_tn with a caller workspace -- the existing entry point,
the same kernel and the same summation order, so the tolerance is that of test_dgemm_against_torch
(1e-12 of the largest element) and in practice the outputs are bit-identical -- on two calls in a
row that share the scratch array, the second with a different shape and different data."""
import pytest

torch = pytest.importorskip('torch')
pytestmark = pytest.mark.gpu

# (M, N, K, splits): ragged tiles, an odd N (the partials' row stride is N + 1), a split count that
# leaves the last K range empty, a head + tail launch (M % 128 in 1..64)
SHAPES = [(300, 257, 4096, 5), (37, 1590, 2512, 3), (513, 200, 272, 2), (8, 130, 48, 4)]


def both(lib, _lib, M, N, K, splits, seed):
    gen = torch.Generator(device='cuda').manual_seed(seed)
    A = torch.randn(M, K, device='cuda', dtype=torch.float64, generator=gen)
    B = torch.randn(N, K, device='cuda', dtype=torch.float64, generator=gen)
    ldc = N + 3                                     # the caller's own row stride, pad columns untouched
    sentinel = -7.25
    out_static = torch.full((M, ldc), sentinel, device='cuda', dtype=torch.float64)
    ldw = N + (N & 1)
    out_ws = torch.zeros(M, ldw, device='cuda', dtype=torch.float64)
    ws = torch.zeros(splits, M, ldw, device='cuda', dtype=torch.float64)
    torch.cuda.synchronize()
    _lib.check(lib.vg_dgemm_tn_static(A.data_ptr(), B.data_ptr(), out_static.data_ptr(), M, N, K, K, K, ldc,
                                      0.75, splits, None))
    _lib.check(lib.vg_dgemm_tn(A.data_ptr(), B.data_ptr(), out_ws.data_ptr(), M, N, K, K, K, ldw, 0.75, 0.0,
                               splits, ws.data_ptr(), None))
    torch.cuda.synchronize()
    ref = 0.75 * (A @ B.T)
    scale = ref.abs().max()
    assert (out_static[:, :N] - out_ws[:, :N]).abs().max() <= 1e-12 * scale, (M, N, K, splits)
    assert (out_static[:, :N] - ref).abs().max() <= 1e-12 * scale, (M, N, K, splits)
    assert bool((out_static[:, N:] == sentinel).all()), 'pad columns of C were written'


def test_static_scratch_gemm_matches_workspace_gemm_twice_in_a_row():
    from vega_b200 import _lib
    lib = _lib.load()
    for seed, (M, N, K, splits) in enumerate(SHAPES):
        both(lib, _lib, M, N, K, splits, seed)
    # the same shape again on new data: the array still holds the previous call's partials
    both(lib, _lib, *SHAPES[0], seed=11)
    both(lib, _lib, *SHAPES[0], seed=12)


def test_static_scratch_gemm_rejects_what_does_not_fit():
    from vega_b200 import _lib
    lib = _lib.load()
    M, N, K, splits = 2048, 1024, 32, 2           # 2 * 2048 * 1024 doubles = 2x the array
    assert splits * M * N > _lib.STATIC_SCRATCH_DOUBLES
    A = torch.zeros(M, K, device='cuda', dtype=torch.float64)
    B = torch.zeros(N, K, device='cuda', dtype=torch.float64)
    out = torch.zeros(M, N, device='cuda', dtype=torch.float64)
    rc = lib.vg_dgemm_tn_static(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, K, K, N, 1.0, splits, None)
    assert rc == _lib.ERR_ARG
    assert b'exceeds the static scratch' in lib.vg_last_error()
    # exactly at the limit is accepted
    M = _lib.STATIC_SCRATCH_DOUBLES // (splits * N)
    _lib.check(lib.vg_dgemm_tn_static(A.data_ptr(), B.data_ptr(), out.data_ptr(), M, N, K, K, K, N, 1.0, splits, None))
    torch.cuda.synchronize()
    assert float(out.abs().max()) == 0.0
    # splits = 1 has no partials: that is vg_dgemm_tn's job
    assert lib.vg_dgemm_tn_static(A.data_ptr(), B.data_ptr(), out.data_ptr(), 8, N, K, K, K, N, 1.0, 1, None) == _lib.ERR_ARG
