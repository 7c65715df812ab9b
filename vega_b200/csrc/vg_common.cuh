// Shared declarations of the vega_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

// VG_OK / VG_ERR_* / VG_STATUS_*: public, in include/vega_b200.h
#include "../../include/vega_b200.h"

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
