// Shared declarations of the vega_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define VG_OK 0
#define VG_ERR_CUDA 1
#define VG_ERR_ARG 2
#define VG_ERR_STATE 3

// per-vector status bits written by the kernels (vega/utils.py:444-453 exceptions)
#define VG_STATUS_BOUNDS 1   // r outside the FFTLog range  -> VegaBoundsError
#define VG_STATUS_ARINYO 2   // NaN/Inf in the Arinyo D_NL  -> VegaArinyoError
#define VG_STATUS_NONFINITE 4

#define VG_CUDA_CHECK(expr)                                                        \
  do {                                                                             \
    cudaError_t err__ = (expr);                                                    \
    if (err__ != cudaSuccess) {                                                    \
      vg_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                   \
                   cudaGetErrorString(err__));                                     \
      return VG_ERR_CUDA;                                                          \
    }                                                                              \
  } while (0)

void vg_set_error(const char* fmt, ...);

static inline int vg_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }
