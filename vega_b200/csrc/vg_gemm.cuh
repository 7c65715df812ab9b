// FP64 tensor-core GEMM for sm_100a:  C[m,n] = alpha * sum_k A[m,k] * B[n,k]  (+ beta*C)
//
// Both operands are K-contiguous ("TN"): A is the batch of model vectors (one row per
// parameter vector), B is a static operator stored row-per-output exactly as the
// reference holds it (distortion matrix rows, metal-matrix rows, C^-1 rows, Hankel
// operator rows).  This one kernel carries every contraction of the path:
//   - Hankel/spline operator  (vega/pktoxi.py:141-152, mcfit.P2xi + interp1d)
//   - metal matrices          (vega/metals.py:338-367)
//   - distortion matrix       (vega/model.py:143-144)
//   - r^T C^-1 r              (vega/vega_interface.py:294-316)   [ROWDOT epilogue]
//
// Math is DMMA (mma.sync m8n8k4 f64: tcgen05 has no FP64 kind), fed from shared memory
// filled by a 3-stage cp.async pipeline.  CTA tile 128x128x16, 8 warps as 2x4, warp tile
// 64x32 = 8x4 DMMA fragments; smem rows padded to 20 doubles so the 64-bit fragment
// loads of a half-warp hit 16 distinct bank pairs.
#pragma once
#include "vg_common.cuh"

namespace vg {

constexpr int GEMM_BM = 128, GEMM_BN = 128, GEMM_BK = 16;
constexpr int GEMM_LDS = 20;          // padded smem row stride (doubles)
constexpr int GEMM_STAGES = 3;
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_SMEM = GEMM_STAGES * (GEMM_BM + GEMM_BN) * GEMM_LDS * 8;

enum GemmEpilogue { GEMM_STORE = 0, GEMM_ROWDOT = 1 };

struct GemmArgs {
  const double* A; const double* B; double* C;
  int M, N, K;                 // K must be a multiple of 16 (operands are zero padded)
  long long lda, ldb, ldc;     // in doubles, even
  long long strideA, strideB, strideC;   // per batch (blockIdx.z / splits)
  int batch, splits;           // grid.z = batch*splits; split s covers K range s*Kper..
  long long split_stride;      // C offset per split (doubles)
  double alpha, beta;
  // ROWDOT: out[m, blockIdx.y*splits + split] = sum_n acc[m,n] * D[m,n]   (C = out, ldc = n_tiles*splits)
  const double* D; long long ldd;
};

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem, bool pred) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  int bytes = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(s), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// BM = 128: 2 x 4 warps, warp tile 64 x 32.  BM = 32 (batches of a few vectors: the scalar API,
// gradient stencils): 1 x 8 warps, warp tile 32 x 16 -- a 128-row tile would spend its DMMA
// cycles on rows that do not exist, and the launch is bound by streaming the operator anyway.
template <int EPI, int BM>
__global__ void __launch_bounds__(GEMM_THREADS, BM == 128 ? 1 : 2) dgemm_tn_kernel(GemmArgs g) {
  constexpr int MI = BM == 128 ? 8 : 4;                     // 8-row fragments per warp
  constexpr int WARPS_M = BM / (MI * 8), WARPS_N = 8 / WARPS_M;
  constexpr int NJ = GEMM_BN / (WARPS_N * 8);               // 8-column fragments per warp
  constexpr int A_CHUNKS = BM * 8 / GEMM_THREADS;           // 16-byte chunks of the A tile per thread
  extern __shared__ __align__(16) double smem[];
  double* As = smem;                                        // [STAGES][BM][LDS]
  double* Bs = smem + GEMM_STAGES * BM * GEMM_LDS;          // [STAGES][BN][LDS]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / WARPS_N, wn = warp % WARPS_N;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * GEMM_BN;
  const int bz = blockIdx.z / g.splits, sp = blockIdx.z % g.splits;

  const int ktiles_total = g.K / GEMM_BK;
  const int kt_per = (ktiles_total + g.splits - 1) / g.splits;
  const int kt_begin = sp * kt_per;
  const int kt_end = min(ktiles_total, kt_begin + kt_per);
  const int nkt = max(0, kt_end - kt_begin);

  const double* A = g.A + (long long)bz * g.strideA;
  const double* B = g.B + (long long)bz * g.strideB;

  // global -> smem mapping: 128 rows x 8 chunks of 16 B per operand, 4 chunks per thread
  auto load_stage = [&](int stage, int kt) {
    const int k0 = (kt_begin + kt) * GEMM_BK;
    double* as = As + stage * BM * GEMM_LDS;
    double* bs = Bs + stage * GEMM_BN * GEMM_LDS;
#pragma unroll
    for (int i = 0; i < A_CHUNKS; ++i) {
      int c = tid + i * GEMM_THREADS;
      int row = c >> 3, ch = c & 7;
      int gm = m0 + row;
      bool p = gm < g.M;
      cp_async16(as + row * GEMM_LDS + ch * 2, A + (long long)(p ? gm : 0) * g.lda + k0 + ch * 2, p);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int c = tid + i * GEMM_THREADS;        // 0..1023
      int row = c >> 3, ch = c & 7;
      int gn = n0 + row;
      bool q = gn < g.N;
      cp_async16(bs + row * GEMM_LDS + ch * 2, B + (long long)(q ? gn : 0) * g.ldb + k0 + ch * 2, q);
    }
  };

  double acc[MI][NJ][2];
#pragma unroll
  for (int i = 0; i < MI; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
  for (int s = 0; s < GEMM_STAGES - 1; ++s) {
    if (s < nkt) load_stage(s, s);
    cp_async_commit();
  }

  const int frow = lane >> 2, fcol = lane & 3;
  for (int kt = 0; kt < nkt; ++kt) {
    cp_async_wait<GEMM_STAGES - 2>();
    __syncthreads();
    {   // prefetch tile kt+STAGES-1 into the slot freed at iteration kt-1
      int nk = kt + GEMM_STAGES - 1;
      if (nk < nkt) load_stage(nk % GEMM_STAGES, nk);
      cp_async_commit();
    }
    const double* as = As + (kt % GEMM_STAGES) * BM * GEMM_LDS + (wm * MI * 8 + frow) * GEMM_LDS + fcol;
    const double* bs = Bs + (kt % GEMM_STAGES) * GEMM_BN * GEMM_LDS + (wn * NJ * 8 + frow) * GEMM_LDS + fcol;
#pragma unroll
    for (int kk = 0; kk < GEMM_BK; kk += 4) {
      double a[MI], b[NJ];
#pragma unroll
      for (int i = 0; i < MI; ++i) a[i] = as[i * 8 * GEMM_LDS + kk];
#pragma unroll
      for (int j = 0; j < NJ; ++j) b[j] = bs[j * 8 * GEMM_LDS + kk];
#pragma unroll
      for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
  }
  cp_async_wait<0>();

  if (EPI == GEMM_STORE) {
    double* C = g.C + (long long)bz * g.strideC + (long long)sp * g.split_stride;
#pragma unroll
    for (int i = 0; i < MI; ++i) {
      int m = m0 + wm * MI * 8 + i * 8 + frow;
      if (m >= g.M) continue;
#pragma unroll
      for (int j = 0; j < NJ; ++j) {
        int n = n0 + wn * NJ * 8 + j * 8 + fcol * 2;
        double* c = C + (long long)m * g.ldc + n;
        double v0 = g.alpha * acc[i][j][0], v1 = g.alpha * acc[i][j][1];
        if (n + 1 < g.N) {
          if (g.beta != 0.0) { double2 o = *reinterpret_cast<double2*>(c); v0 += g.beta * o.x; v1 += g.beta * o.y; }
          *reinterpret_cast<double2*>(c) = make_double2(v0, v1);
        } else if (n < g.N) {
          if (g.beta != 0.0) v0 += g.beta * c[0];
          c[0] = v0;
        }
      }
    }
  } else {
    // ROWDOT: deterministic reduction of acc[m,n]*D[m,n] over this CTA's 128 columns
    __syncthreads();
    double* red = smem;   // [BM][WARPS_N] partials per n-warp
#pragma unroll
    for (int i = 0; i < MI; ++i) {
      int ml = wm * MI * 8 + i * 8 + frow;
      int m = m0 + ml;
      double s = 0.0;
      if (m < g.M) {
#pragma unroll
        for (int j = 0; j < NJ; ++j) {
          int n = n0 + wn * NJ * 8 + j * 8 + fcol * 2;
          const double* d = g.D + (long long)m * g.ldd + n;
          if (n < g.N) s += acc[i][j][0] * d[0];
          if (n + 1 < g.N) s += acc[i][j][1] * d[1];
        }
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (fcol == 0) red[ml * WARPS_N + wn] = s;
    }
    __syncthreads();
    if (tid < BM) {
      int m = m0 + tid;
      if (m < g.M) {
        double s = 0.0;
#pragma unroll
        for (int w = 0; w < WARPS_N; w += 4)
          s += (red[tid * WARPS_N + w] + red[tid * WARPS_N + w + 1]) +
               (red[tid * WARPS_N + w + 2] + red[tid * WARPS_N + w + 3]);
        // linear in the K range: each split contributes its own partial column
        g.C[(long long)bz * g.strideC + (long long)m * g.ldc + blockIdx.y * g.splits + sp] = g.alpha * s;
      }
    }
  }
}

// sums `splits` partial slices: C[i] = sum_s W[s*stride + i] (fixed order => deterministic)
__global__ void reduce_splits_kernel(const double* __restrict__ W, double* __restrict__ C,
                                     long long n, int splits, long long stride) {
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = W[i];
  for (int k = 1; k < splits; ++k) s += W[(long long)k * stride + i];
  C[i] = s;
}

constexpr int GEMM_BM_SMALL = 32;
constexpr int GEMM_SMEM_SMALL = GEMM_STAGES * (GEMM_BM_SMALL + GEMM_BN) * GEMM_LDS * 8;
// rows of A at or below which the 32-row tile is used
constexpr int GEMM_SMALL_M = 96;
inline int gemm_tile_m(int M) { return M <= GEMM_SMALL_M ? GEMM_BM_SMALL : GEMM_BM; }

inline int launch_dgemm(const GemmArgs& g, int epilogue, cudaStream_t stream) {
  if (g.K % GEMM_BK != 0 || (g.lda & 1) || (g.ldb & 1) || g.M <= 0 || g.N <= 0) {
    vg_set_error("dgemm: bad shape M=%d N=%d K=%d lda=%lld ldb=%lld", g.M, g.N, g.K, g.lda, g.ldb);
    return VG_ERR_ARG;
  }
  const int bm = gemm_tile_m(g.M);
  dim3 grid(vg_cdiv(g.M, bm), vg_cdiv(g.N, GEMM_BN), g.batch * g.splits);
  static bool attr_set = false;
  if (!attr_set) {
    VG_CUDA_CHECK(cudaFuncSetAttribute(dgemm_tn_kernel<GEMM_STORE, GEMM_BM>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
    VG_CUDA_CHECK(cudaFuncSetAttribute(dgemm_tn_kernel<GEMM_ROWDOT, GEMM_BM>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM));
    VG_CUDA_CHECK(cudaFuncSetAttribute(dgemm_tn_kernel<GEMM_STORE, GEMM_BM_SMALL>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_SMALL));
    VG_CUDA_CHECK(cudaFuncSetAttribute(dgemm_tn_kernel<GEMM_ROWDOT, GEMM_BM_SMALL>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_SMALL));
    attr_set = true;
  }
  if (bm == GEMM_BM) {
    if (epilogue == GEMM_STORE)
      dgemm_tn_kernel<GEMM_STORE, GEMM_BM><<<grid, GEMM_THREADS, GEMM_SMEM, stream>>>(g);
    else
      dgemm_tn_kernel<GEMM_ROWDOT, GEMM_BM><<<grid, GEMM_THREADS, GEMM_SMEM, stream>>>(g);
  } else {
    if (epilogue == GEMM_STORE)
      dgemm_tn_kernel<GEMM_STORE, GEMM_BM_SMALL><<<grid, GEMM_THREADS, GEMM_SMEM_SMALL, stream>>>(g);
    else
      dgemm_tn_kernel<GEMM_ROWDOT, GEMM_BM_SMALL><<<grid, GEMM_THREADS, GEMM_SMEM_SMALL, stream>>>(g);
  }
  VG_CUDA_CHECK(cudaGetLastError());
  return VG_OK;
}

}  // namespace vg
