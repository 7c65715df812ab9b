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
// filled by a 3-stage cp.async pipeline.  CTA tile 128x128x32 (216 KB of shared memory, one
// CTA per SM), 8 warps as 2x4, warp tile 64x32 = 8x4 DMMA fragments; smem rows padded to
// BK+4 doubles so the 64-bit fragment loads of a half-warp hit 16 distinct bank pairs.
// One CTA barrier per 32 k (256 DMMAs per warp): with 16 k per barrier ncu showed the
// barrier + cp.async wait + LDS latency stalls at ~16 % of the DMMA pipe's cycles.
#pragma once
#include "vg_common.cuh"

namespace vg {

constexpr int GEMM_BM = 128, GEMM_BN = 128;
constexpr int GEMM_BK = 16;           // K granularity of the operands (zero padded to it)
constexpr int GEMM_BK_BIG = 32;       // k per pipeline stage of the 128-row tile
constexpr int GEMM_STAGES = 3;
constexpr int GEMM_THREADS = 256;
constexpr int gemm_smem_bytes(int bm, int bk) { return GEMM_STAGES * (bm + GEMM_BN) * (bk + 4) * 8; }
constexpr int GEMM_SMEM = gemm_smem_bytes(GEMM_BM, GEMM_BK_BIG);

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
  // ROWDOT with A = D and B a SYMMETRIC N x N matrix (the quadratic form r^T B r): sym = 1 makes the
  // column tile [n0, n0 + 128) read only k >= n0 -- its diagonal block once, everything to the right
  // of it twice -- so half of B is streamed and half of the DMMAs are issued; the splits of a tile
  // divide that tile's own k range
  int sym;
};

// shared-memory operands are addressed by their 32-bit shared-window offsets, computed once
// per CTA: with generic pointers the compiler re-derived the window base (S2UR of
// SR_CgaCtaId, a long-scoreboard read) inside the k loop -- 3 % of the stall samples
__device__ __forceinline__ void cp_async16(unsigned smem_addr, const void* gmem, bool pred) {
  int bytes = pred ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(smem_addr), "l"(gmem), "r"(bytes));
}
__device__ __forceinline__ double lds_f64(unsigned smem_addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];\n" : "=d"(v) : "r"(smem_addr));
  return v;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}
__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// BM = 128: 2 x 4 warps, warp tile 64 x 32.  BM = 8..64 in steps of 8 (batches of a few
// vectors: the scalar API is 1 row, a gradient stencil 2P+1 ~ 35): 1 x 8 warps, warp tile
// BM x 16 -- the tile is as tall as the batch (rounded up to the 8 rows of a DMMA fragment),
// so no DMMA cycles go to rows that do not exist and the launch is bound by streaming the
// operator from HBM.
template <int EPI, int BM, int BK>
__global__ void __launch_bounds__(GEMM_THREADS, BM == 128 ? 1 : 2) dgemm_tn_kernel(GemmArgs g) {
  constexpr int GEMM_LDS = BK + 4;                          // padded smem row stride (doubles)
  constexpr int CPR = BK / 2;                               // 16-byte chunks per tile row
  constexpr int MI = BM == 128 ? 8 : BM / 8;                // 8-row fragments per warp
  constexpr int WARPS_M = BM / (MI * 8), WARPS_N = 8 / WARPS_M;
  constexpr int NJ = GEMM_BN / (WARPS_N * 8);               // 8-column fragments per warp
  constexpr int A_CHUNKS = (BM * CPR + GEMM_THREADS - 1) / GEMM_THREADS;   // 16-byte chunks of the A tile per thread
  constexpr int B_CHUNKS = GEMM_BN * CPR / GEMM_THREADS;
  extern __shared__ __align__(16) double smem[];
  double* As = smem;                                        // [STAGES][BM][LDS]
  double* Bs = smem + GEMM_STAGES * BM * GEMM_LDS;          // [STAGES][BN][LDS]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / WARPS_N, wn = warp % WARPS_N;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * GEMM_BN;
  const int bz = blockIdx.z / g.splits, sp = blockIdx.z % g.splits;

  const int ktiles_total = (g.K + BK - 1) / BK;
  const bool sym = EPI == GEMM_ROWDOT && g.sym != 0;
  const int kt_lo = sym ? n0 / BK : 0;                       // (n0 is a multiple of 128, BK of 16 or 32)
  const int kt_per = (ktiles_total - kt_lo + g.splits - 1) / g.splits;
  const int kt_begin = kt_lo + sp * kt_per;
  const int kt_end = min(ktiles_total, kt_begin + kt_per);
  const int nkt = max(0, kt_end - kt_begin);
  const int kt_diag_end = kt_lo + GEMM_BN / BK;              // sym: first k tile right of the diagonal block

  const double* A = g.A + (long long)bz * g.strideA;
  const double* B = g.B + (long long)bz * g.strideB;

  // global -> smem mapping: a pass of the 256 threads covers RPP rows x CPR chunks of 16 B;
  // each thread keeps ONE row pointer per operand (its chunk column is the same in every
  // pass), the row-in-range bits are loop invariant, and chunks past K (a last 32-wide tile
  // over a K that is a multiple of 16 only) are zero filled through the src-size operand.
  constexpr int RPP = GEMM_THREADS / CPR;                   // tile rows per pass
  const unsigned as_base = (unsigned)__cvta_generic_to_shared(As);
  const unsigned bs_base = (unsigned)__cvta_generic_to_shared(Bs);
  const int lrow = tid / CPR, lch = (tid % CPR) * 2;
  const double* a_src = A + (long long)(m0 + lrow) * g.lda + lch;
  const double* b_src = B + (long long)(n0 + lrow) * g.ldb + lch;
  const long long a_pass = (long long)RPP * g.lda, b_pass = (long long)RPP * g.ldb;
  unsigned a_rows = 0, b_rows = 0;
#pragma unroll
  for (int i = 0; i < A_CHUNKS; ++i) a_rows |= (m0 + lrow + i * RPP < g.M ? 1u : 0u) << i;
#pragma unroll
  for (int i = 0; i < B_CHUNKS; ++i) b_rows |= (n0 + lrow + i * RPP < g.N ? 1u : 0u) << i;
  const unsigned a_dst = as_base + (lrow * GEMM_LDS + lch) * 8;
  const unsigned b_dst = bs_base + (lrow * GEMM_LDS + lch) * 8;
  // chunk q of a stage refill: q < A_CHUNKS from A, the rest from B
  auto load_chunk = [&](int q, int stage, long long k0, bool live) {
    if (q < A_CHUNKS) {
      if (BM % RPP == 0 || lrow + q * RPP < BM)      // a short tile ends inside the last pass
        cp_async16(a_dst + stage * (BM * GEMM_LDS * 8) + q * (RPP * GEMM_LDS * 8),
                   a_src + q * a_pass + k0, live && ((a_rows >> q) & 1u));
    } else {
      const int r = q - A_CHUNKS;
      cp_async16(b_dst + stage * (GEMM_BN * GEMM_LDS * 8) + r * (RPP * GEMM_LDS * 8),
                 b_src + r * b_pass + k0, live && ((b_rows >> r) & 1u));
    }
  };
  constexpr int N_CHUNKS = A_CHUNKS + B_CHUNKS;
  constexpr int N_STEPS = BK / 4;
  constexpr int CHUNKS_PER_STEP = (N_CHUNKS + N_STEPS - 1) / N_STEPS;

  double acc[MI][NJ][2];
#pragma unroll
  for (int i = 0; i < MI; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

#pragma unroll
  for (int s = 0; s < GEMM_STAGES - 1; ++s) {
    const long long k0 = (long long)(kt_begin + s) * BK;
    const bool live = s < nkt && k0 + lch < g.K;
#pragma unroll
    for (int q = 0; q < N_CHUNKS; ++q) load_chunk(q, s, k0, live);
    cp_async_commit();
  }

  const int frow = lane >> 2, fcol = lane & 3;
  const unsigned as_frag = as_base + ((wm * MI * 8 + frow) * GEMM_LDS + fcol) * 8;
  const unsigned bs_frag = bs_base + ((wn * NJ * 8 + frow) * GEMM_LDS + fcol) * 8;
  int cur = 0, nxt = GEMM_STAGES - 1;      // stage consumed / stage refilled at this iteration
  for (int kt = 0; kt < nkt; ++kt) {
    if (sym && kt > 0 && kt_begin + kt == kt_diag_end) {
      // leaving the diagonal block: what follows counts twice, i.e. what came before counts half
      // (the epilogue doubles the sum; both scalings are exact)
#pragma unroll
      for (int i = 0; i < MI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) { acc[i][j][0] *= 0.5; acc[i][j][1] *= 0.5; }
    }
    cp_async_wait<GEMM_STAGES - 2>();
    __syncthreads();
    // tile kt+STAGES-1 goes into the slot freed at iteration kt-1; its cp.async are issued
    // between the DMMA steps below (in the issue slots the 16-cycle DMMA cadence leaves
    // free) instead of in a block after the barrier, where both warps of a scheduler would
    // leave the DMMA pipe idle while they compute addresses
    const long long k_next = (long long)(kt_begin + kt + GEMM_STAGES - 1) * BK;
    const bool live = kt + GEMM_STAGES - 1 < nkt && k_next + lch < g.K;
    const unsigned as = as_frag + cur * (BM * GEMM_LDS * 8);
    const unsigned bs = bs_frag + cur * (GEMM_BN * GEMM_LDS * 8);
    double a[2][MI], b[2][NJ];
#pragma unroll
    for (int i = 0; i < MI; ++i) a[0][i] = lds_f64(as + i * (8 * GEMM_LDS * 8));
#pragma unroll
    for (int j = 0; j < NJ; ++j) b[0][j] = lds_f64(bs + j * (8 * GEMM_LDS * 8));
#pragma unroll
    for (int st = 0; st < N_STEPS; ++st) {
      const int c = st & 1;
      if (st + 1 < N_STEPS) {
#pragma unroll
        for (int i = 0; i < MI; ++i) a[c ^ 1][i] = lds_f64(as + i * (8 * GEMM_LDS * 8) + (st + 1) * 32);
#pragma unroll
        for (int j = 0; j < NJ; ++j) b[c ^ 1][j] = lds_f64(bs + j * (8 * GEMM_LDS * 8) + (st + 1) * 32);
      }
#pragma unroll
      for (int i = 0; i < MI; ++i) {
#pragma unroll
        for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[c][i], b[c][j]);
        // the refill chunks of this step, spread over its fragment rows
        constexpr int EVERY = MI >= CHUNKS_PER_STEP ? MI / CHUNKS_PER_STEP : 1;
        if ((i + 1) % EVERY == 0 && (i + 1) / EVERY <= CHUNKS_PER_STEP) {
          const int q = st * CHUNKS_PER_STEP + (i + 1) / EVERY - 1;
          if (q < N_CHUNKS) load_chunk(q, nxt, k_next, live);
        }
      }
      if (MI < CHUNKS_PER_STEP) {   // fewer fragment rows than chunks: the rest after the step
#pragma unroll
        for (int r = MI; r < CHUNKS_PER_STEP; ++r) {
          const int q = st * CHUNKS_PER_STEP + r;
          if (q < N_CHUNKS) load_chunk(q, nxt, k_next, live);
        }
      }
    }
    cp_async_commit();
    cur = cur + 1 == GEMM_STAGES ? 0 : cur + 1;
    nxt = nxt + 1 == GEMM_STAGES ? 0 : nxt + 1;
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
        if (sym && kt_end > kt_diag_end) s *= 2.0;
        g.C[(long long)bz * g.strideC + (long long)m * g.ldc + blockIdx.y * g.splits + sp] = g.alpha * s;
      }
    }
  }
}

// dst[j][i] = src[i][j] for i < rows, j < cols (32 x 32 tiles through shared memory); block (32, 8),
// grid (ceil(cols / 32), ceil(rows / 32))
__global__ void transpose_kernel(const double* __restrict__ src, long long src_ld, int rows, int cols,
                                 double* __restrict__ dst, long long dst_ld) {
  __shared__ double tile[32][33];
  const int x = blockIdx.x * 32 + threadIdx.x;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int y = blockIdx.y * 32 + threadIdx.y + k;
    if (x < cols && y < rows) tile[threadIdx.y + k][threadIdx.x] = src[(long long)y * src_ld + x];
  }
  __syncthreads();
  const int xo = blockIdx.y * 32 + threadIdx.x;
#pragma unroll
  for (int k = 0; k < 32; k += 8) {
    const int yo = blockIdx.x * 32 + threadIdx.y + k;
    if (xo < rows && yo < cols) dst[(long long)yo * dst_ld + xo] = tile[threadIdx.x][threadIdx.y + k];
  }
}

// A <- (A + A^T) / 2 on the leading n x n block of A[n][ld], in place (each pair i < j belongs to one thread)
__global__ void symmetrize_kernel(double* __restrict__ A, int n, long long ld) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x, i = blockIdx.y;
  if (j >= n || j <= i) return;
  const double v = 0.5 * (A[(long long)i * ld + j] + A[(long long)j * ld + i]);
  A[(long long)i * ld + j] = v;
  A[(long long)j * ld + i] = v;
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

// rows of A up to which a batch-height tile (8..64 rows) is used; 65..96 rows take three
// 32-row tiles, anything taller the 128-row tile
constexpr int GEMM_SMALL_M = 96;
inline int gemm_tile_m(int M) {
  if (M <= 64) return 8 * ((M + 7) / 8);
  return M <= GEMM_SMALL_M ? 32 : GEMM_BM;
}

template <int EPI, int BM, int BK>
inline int launch_dgemm_tile(const GemmArgs& g, dim3 grid, cudaStream_t stream) {
  constexpr int smem = gemm_smem_bytes(BM, BK);
  static bool attr_set = false;
  if (!attr_set) {
    VG_CUDA_CHECK(cudaFuncSetAttribute(dgemm_tn_kernel<EPI, BM, BK>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  dgemm_tn_kernel<EPI, BM, BK><<<grid, GEMM_THREADS, smem, stream>>>(g);
  return VG_OK;
}
template <int EPI>
inline int launch_dgemm_epi(const GemmArgs& g, int bm, dim3 grid, cudaStream_t stream) {
  switch (bm) {
    case 8: return launch_dgemm_tile<EPI, 8, GEMM_BK>(g, grid, stream);
    case 16: return launch_dgemm_tile<EPI, 16, GEMM_BK>(g, grid, stream);
    case 24: return launch_dgemm_tile<EPI, 24, GEMM_BK>(g, grid, stream);
    case 32: return launch_dgemm_tile<EPI, 32, GEMM_BK>(g, grid, stream);
    case 40: return launch_dgemm_tile<EPI, 40, GEMM_BK>(g, grid, stream);
    case 48: return launch_dgemm_tile<EPI, 48, GEMM_BK>(g, grid, stream);
    case 56: return launch_dgemm_tile<EPI, 56, GEMM_BK>(g, grid, stream);
    case 64: return launch_dgemm_tile<EPI, 64, GEMM_BK>(g, grid, stream);
    default: return launch_dgemm_tile<EPI, GEMM_BM, GEMM_BK_BIG>(g, grid, stream);
  }
}

// kernels launched by launch_dgemm (bench bookkeeping)
inline long long& gemm_launches() { static long long n = 0; return n; }

inline int launch_dgemm(const GemmArgs& g, int epilogue, cudaStream_t stream) {
  if (g.K % GEMM_BK != 0 || (g.lda & 1) || (g.ldb & 1) || g.M <= 0 || g.N <= 0) {
    vg_set_error("dgemm: bad shape M=%d N=%d K=%d lda=%lld ldb=%lld", g.M, g.N, g.K, g.lda, g.ldb);
    return VG_ERR_ARG;
  }
  // a batch a few rows taller than a multiple of 128 (the 2F^2+1 vectors of a Hessian stencil:
  // 513 for F = 16) would spend a whole 128-row tile on them: those rows go to a second launch
  // with a batch-height tile, which only streams the operator once more
  const int rem = g.M % GEMM_BM;
  if (g.M > GEMM_BM && rem > 0 && rem <= 64) {
    GemmArgs head = g, tail = g;
    head.M = g.M - rem;
    tail.M = rem;
    tail.A += (long long)head.M * g.lda;
    tail.C += (long long)head.M * g.ldc;
    if (g.D) tail.D += (long long)head.M * g.ldd;
    const int rc = launch_dgemm(head, epilogue, stream);
    return rc != VG_OK ? rc : launch_dgemm(tail, epilogue, stream);
  }
  const int bm = gemm_tile_m(g.M);
  dim3 grid(vg_cdiv(g.M, bm), vg_cdiv(g.N, GEMM_BN), g.batch * g.splits);
  const int rc = epilogue == GEMM_STORE ? launch_dgemm_epi<GEMM_STORE>(g, bm, grid, stream)
                                       : launch_dgemm_epi<GEMM_ROWDOT>(g, bm, grid, stream);
  if (rc != VG_OK) return rc;
  VG_CUDA_CHECK(cudaGetLastError());
  ++gemm_launches();
  return VG_OK;
}

}  // namespace vg
