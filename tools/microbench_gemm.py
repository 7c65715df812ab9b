"""GPU microbench: own DMMA DGEMM vs torch (cuBLAS) FP64 matmul; correctness + TFLOP/s."""
import ctypes
import json
import sys
from pathlib import Path

import torch

lib = ctypes.CDLL(str(Path(__file__).resolve().parents[1] / 'vega_b200' / 'libvega_b200.so'))
lib.vg_dgemm_tn.argtypes = [ctypes.c_void_p] * 3 + [ctypes.c_int] * 3 + [ctypes.c_longlong] * 3 + \
    [ctypes.c_double] * 2 + [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]
lib.vg_last_error.restype = ctypes.c_char_p


def own(A, B, C, splits=1, ws=None):
    M, K = A.shape
    N = B.shape[0]
    rc = lib.vg_dgemm_tn(A.data_ptr(), B.data_ptr(), C.data_ptr(), M, N, K, A.stride(0), B.stride(0),
                         C.stride(0), 1.0, 0.0, splits, ws.data_ptr() if ws is not None else None,
                         torch.cuda.current_stream().cuda_stream)
    assert rc == 0, lib.vg_last_error()


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


out = {}
torch.manual_seed(0)
for (M, N, K, splits) in [(1024, 2500, 37504, 1), (1024, 2500, 37504, 15), (1024, 2512, 2512, 1),
                          (23552, 1632, 816, 1), (8192, 8192, 8192, 1), (1184, 2500, 37504, 1),
                          (33, 1590, 2512, 1), (1024, 5000, 20032, 4),
                          (1, 2500, 37504, 14), (35, 2500, 37504, 14), (64, 2500, 37504, 14), (35, 5000, 20032, 7)]:
    A = torch.randn(M, K, device='cuda', dtype=torch.float64)
    B = torch.randn(N, K, device='cuda', dtype=torch.float64)
    ldc = N + (N & 1)
    C = torch.zeros(M, ldc, device='cuda', dtype=torch.float64)
    ws = torch.zeros(splits, M, ldc, device='cuda', dtype=torch.float64) if splits > 1 else None
    own(A, B, C, splits, ws)
    ref = A @ B.T
    err = ((C[:, :N] - ref).abs().max() / ref.abs().max()).item()
    t_own = timeit(lambda: own(A, B, C, splits, ws))
    t_ref = timeit(lambda: torch.matmul(A, B.T))
    fl = 2.0 * M * N * K
    out[f'{M}x{N}x{K}/s{splits}'] = {'relerr': err, 'own_ms': t_own, 'own_tflops': fl / t_own / 1e9,
                                     'cublas_ms': t_ref, 'cublas_tflops': fl / t_ref / 1e9}
    print(f'{M}x{N}x{K} s{splits}', out[f'{M}x{N}x{K}/s{splits}'], flush=True)

# FP64 FMA peak via torch elementwise chain is not reliable; report cuBLAS DGEMM as the tensor peak
Path('gpurun_out').mkdir(exist_ok=True)
json.dump(out, open('gpurun_out/microbench_gemm.json', 'w'), indent=1)
