// FP64-accurate GEMM on the 5th-generation tensor cores (tcgen05, kind::i8) by exact integer
// slicing -- the "Ozaki scheme" -- for the one contraction that dominates the path: the metal
// matrices,  C[b][i] = sum_k xmet[b][k] * Mcat[i][k]  (vega/metals.py:338-367 applied to a batch;
// K = n_metal * Nm = 37 680 on the eBOSS-shaped auto-correlation).
//
// tcgen05.mma has no FP64 kind, so the DMMA kernel (vg_gemm.cuh) sits at the FP64 pipe's
// ceiling.  Here every row of both operands is scaled by a power of two (exact) to |x| <= 1 and
// cut into S signed radix-128 digits  x = sum_t q_t 2^-(6+7t),  q_t in [-64, 64]  (exact: each
// digit is rint() of an exactly representable remainder).  Digit products are accumulated
// EXACTLY in s32 by the tensor cores (|q q'| <= 2^12, K_split * S * 2^12 < 2^31 is checked on
// the host), one TMEM accumulator per diagonal d = t + u, and the epilogue recombines
//   C = 2^(eA_i + eB_j - 12) * sum_d 2^(-7d) P_d          (Horner in FP64, a few ulp).
// The only approximation is the truncation to S digits and to the diagonals d < S:
//   |dC_ij| <= (S + 1) * K * 2^(-7S) * max_k|A_ik| * max_k|B_jk|
// (S = 7: 5e-10 relative to the product of the row maxima for K = 37 680 -- DESIGN.md has the
// budget against the 1e-10 * max|xi| bar; the digits of the static operator are a plan-time cost).
//
// Kernel: one CTA per (128 x 64 output tile, K split).  TMEM holds the S diagonal accumulators
// (S * 64 of 512 columns).  A warp-specialised pipeline: warp 0 issues the TMA loads
// (cp.async.bulk.tensor.3d: ALL S digit planes of a 64-wide K block in one box per operand, so a
// digit plane is fetched once and used by up to S products), warp 1 issues the S(S+1)/2 x 2
// tcgen05.mma of a K block from one thread, warp 2 owns the TMEM allocation, warps 4-7 read the
// accumulators back (tcgen05.ld) and write FP64.  Operands are K-major with the 64-byte swizzle.
#pragma once
#include <cuda.h>

#include "vg_common.cuh"

namespace vg {

// OZ_BK int8 elements per pipeline stage = the swizzle span (32 or 64 bytes).  Measured on C2 with
// S = 7: 64 bytes (2 stages of 86 KB) 4.16 ms for the two metal GEMMs, 32 bytes (5 stages of 43 KB,
// SWIZZLE_32B) 5.26 ms -- the finer pipeline does not pay for the narrower rows
constexpr int OZ_BM = 128, OZ_BN = 64, OZ_BK = 64;
static_assert(OZ_BK == 32 || OZ_BK == 64, "the stage K extent is one 32- or 64-byte swizzle span");
constexpr int OZ_MAX_S = 8;
constexpr int OZ_THREADS = 256;
constexpr int OZ_TMEM_COLS = 512;
constexpr int OZ_NONFINITE = INT_MIN;                // row exponent of a row holding NaN / Inf

struct OzakiArgs {
  int M, N, S;
  int kb_total, kb_per_split;      // K blocks of OZ_BK: all, per split (grid.z)
  const int* expA;                 // [M] row exponents of A
  const int* expB;                 // [N]
  double* C; long long ldc, split_stride;
  int* error_flag;                 // set when an mbarrier wait times out (never in a healthy run)
};

// ---------------------------------------------------------------------------------------------
// digit planes
// ---------------------------------------------------------------------------------------------
// exps[r] = e with max_k |A[r][k]| in [2^(e-1), 2^e)  (0 for an all-zero row)
__global__ void __launch_bounds__(256) ozaki_rowexp_kernel(const double* __restrict__ A, long long lda, int K,
                                                          int* __restrict__ exps) {
  __shared__ double red[256];
  __shared__ int bad[1];
  const long long row = blockIdx.x;
  if (threadIdx.x == 0) bad[0] = 0;
  __syncthreads();
  double m = 0.0;
  int nonfinite = 0;
  for (int k = threadIdx.x; k < K; k += 256) {
    const double v = fabs(A[row * lda + k]);
    if (!isfinite(v)) nonfinite = 1;
    else if (v > m) m = v;
  }
  if (nonfinite) bad[0] = 1;
  red[threadIdx.x] = m;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] = fmax(red[threadIdx.x], red[threadIdx.x + o]);
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    int e = 0;
    if (red[0] > 0.0) frexp(red[0], &e);
    e = max(e, -900);                   // (rows below 2^-900 come out as zeros)
    exps[row] = bad[0] ? OZ_NONFINITE : e;
  }
}

// Row statistics that enter the a-posteriori truncation bound of the digit GEMM: out[0] = max over
// rows of max_k|x|, out[1] = max over rows of sum_k|x|, out[2] = max over rows of #{k: x != 0}
// (non-negative doubles order like their bit patterns, so atomicMax on the bits is exact).
__global__ void __launch_bounds__(256) ozaki_stats_kernel(const double* __restrict__ A, long long lda, int K,
                                                         unsigned long long* __restrict__ out) {
  __shared__ double rmax[256], rsum[256];
  __shared__ int rnz[256];
  const long long row = blockIdx.x;
  double m = 0.0, s = 0.0;
  int nz = 0;
  for (int k = threadIdx.x; k < K; k += 256) {
    const double v = fabs(A[row * lda + k]);
    if (isfinite(v)) { m = fmax(m, v); s += v; nz += v != 0.0; }
  }
  rmax[threadIdx.x] = m; rsum[threadIdx.x] = s; rnz[threadIdx.x] = nz;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
      rmax[threadIdx.x] = fmax(rmax[threadIdx.x], rmax[threadIdx.x + o]);
      rsum[threadIdx.x] += rsum[threadIdx.x + o];
      rnz[threadIdx.x] += rnz[threadIdx.x + o];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    atomicMax(out + 0, (unsigned long long)__double_as_longlong(rmax[0]));
    atomicMax(out + 1, (unsigned long long)__double_as_longlong(rsum[0]));
    atomicMax(out + 2, (unsigned long long)__double_as_longlong((double)rnz[0]));
  }
}

// planes[t][r][k] = digit t of A[r][k] / 2^exps[r]; one thread per 4 consecutive k: a warp reads 1 KB of
// the row in two 16-byte loads per thread and writes 128 contiguous bytes per plane.  Columns k >= K
// (up to the padded row length Kp, a multiple of 16) are zero.
template <int S>
__global__ void __launch_bounds__(256) ozaki_slice_kernel(const double* __restrict__ A, long long lda, int K, int Kp,
                                                         long long R, const int* __restrict__ exps,
                                                         int8_t* __restrict__ planes, long long plane_stride) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  const int chunks = Kp / 4;
  const long long row = idx / chunks;
  const int ch = (int)(idx % chunks);
  if (row >= R) return;
  const int e = exps[row];
  // 2^(6 - e): the first digit is rint(x * 64)
  const double scale = e == OZ_NONFINITE ? 0.0 : __longlong_as_double((long long)(1023 + 6 - e) << 52);
  const double* src = A + row * lda + ch * 4;
  double v[4] = {0.0, 0.0, 0.0, 0.0};
  if (e != OZ_NONFINITE) {
    if (ch * 4 + 3 < K && (reinterpret_cast<unsigned long long>(src) & 15ull) == 0) {
      const double2 lo = *reinterpret_cast<const double2*>(src), hi = *reinterpret_cast<const double2*>(src + 2);
      v[0] = lo.x; v[1] = lo.y; v[2] = hi.x; v[3] = hi.y;
    } else {
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (ch * 4 + j < K) v[j] = src[j];
    }
  }
  unsigned packed[S];
#pragma unroll
  for (int t = 0; t < S; ++t) packed[t] = 0u;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    // |x| <= 64: digit t is x rounded (to nearest, ties to even) to a multiple of 2^(-7t), in units of
    // 2^(-7t).  Adding 1.5 * 2^(52 - 7t) does that rounding in the adder -- the sum's unit in the last
    // place IS 2^(-7t) -- and leaves the digit in the low mantissa bits (two's complement); subtracting
    // it back gives the rounded value, and the remainder x - that is exact.  Same digits as
    // q = rint(x), x = (x - q) * 128 with 3 FP64 adds per digit and no conversion (that form was
    // issue-bound: 80 % of the issue slots at 46 % of the HBM bandwidth).
    double x = v[j] * scale;
#pragma unroll
    for (int t = 0; t < S; ++t) {
      const double magic = __longlong_as_double(((long long)(1023 + 52 - 7 * t) << 52) | (1LL << 51));   // 1.5 * 2^(52-7t)
      const double tt = x + magic;
      packed[t] |= ((unsigned)__double2loint(tt) & 0xffu) << (8 * j);
      x -= tt - magic;
    }
  }
#pragma unroll
  for (int t = 0; t < S; ++t)
    *reinterpret_cast<unsigned*>(planes + t * plane_stride + row * Kp + ch * 4) = packed[t];
}

// ---------------------------------------------------------------------------------------------
// PTX wrappers (sm_100a)
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned oz_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void oz_mbar_init(unsigned bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void oz_mbar_expect_tx(unsigned bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
// bounded wait: a healthy run never comes near the limit; a broken pipeline sets the error flag
// and lets the kernel run out instead of hanging the device
__device__ __forceinline__ bool oz_mbar_wait(unsigned bar, unsigned parity, int* error_flag) {
  const long long t0 = clock64();
  for (;;) {
    unsigned done;
    asm volatile(
        "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
        : "=r"(done) : "r"(bar), "r"(parity) : "memory");
    if (done) return true;
    if (clock64() - t0 > 4000000000LL) { atomicExch(error_flag, 1); return false; }
  }
}
__device__ __forceinline__ void oz_tma_load_3d(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1,
                                               int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];\n"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// the same load delivered to the same shared-memory offset (and signalled on the barrier at the same
// offset) of every CTA of the cluster whose bit is set in `mask`
__device__ __forceinline__ void oz_tma_load_3d_mc(unsigned dst, const CUtensorMap* map, unsigned bar, int c0, int c1,
                                                  int c2, unsigned short mask) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;\n"
      ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "h"(mask) : "memory");
}
__device__ __forceinline__ unsigned oz_cluster_rank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void oz_cluster_sync() {
  asm volatile("barrier.cluster.arrive.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void oz_tmem_alloc(unsigned slot_smem, unsigned cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(slot_smem), "r"(cols));
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n");
}
__device__ __forceinline__ void oz_tmem_dealloc(unsigned taddr, unsigned cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(cols));
}
__device__ __forceinline__ void oz_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void oz_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
// D[tmem] (+)= A[smem] * B[smem]^T, s32 accumulate of s8 x s8
__device__ __forceinline__ void oz_mma_i8(unsigned tmem_d, unsigned long long adesc, unsigned long long bdesc,
                                          unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}\n"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
// arrives on the mbarrier when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void oz_mma_commit(unsigned bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// ... and on the barrier at the same offset in every CTA of the cluster in `mask`
__device__ __forceinline__ void oz_mma_commit_mc(unsigned bar, unsigned short mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
               ::"r"(bar), "h"(mask) : "memory");
}
__device__ __forceinline__ void oz_tmem_ld16(unsigned taddr, unsigned (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr));
}
__device__ __forceinline__ void oz_tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// shared-memory matrix descriptor of a K-major operand tile with the OZ_BK-byte swizzle: rows of
// OZ_BK bytes, 8-row groups 8 * OZ_BK bytes apart (SBO), descriptor version 1 (sm_100), layout type
// 4 (SWIZZLE_64B) or 6 (SWIZZLE_32B) -- field layout as in cute::UMMA::SmemDescriptor
__device__ __forceinline__ unsigned long long oz_smem_desc(unsigned smem_addr) {
  unsigned long long d = 0;
  d |= (unsigned long long)((smem_addr >> 4) & 0x3fffu);        // start address  [0,14)
  d |= (unsigned long long)0u << 16;                            // leading byte offset [16,30): unused (one atom along K)
  d |= (unsigned long long)(((8u * OZ_BK) >> 4) & 0x3fffu) << 32;   // stride byte offset  [32,46)
  d |= (unsigned long long)1u << 46;                            // version             [46,48)
  d |= (unsigned long long)(OZ_BK == 64 ? 4u : 6u) << 61;       // layout type         [61,64)
  return d;
}
// instruction descriptor (cute::UMMA::InstrDescriptor): s32 accumulator, s8 x s8, both K-major;
// n = 64 * (digit planes of the batch operand one instruction covers)
__device__ __forceinline__ constexpr unsigned oz_instr_desc(int n = OZ_BN) {
  return (2u << 4)                      // c_format = S32
         | (1u << 7) | (1u << 10)       // a_format = b_format = INT8 (signed)
         | ((unsigned)(n >> 3) << 17)       // n_dim
         | ((unsigned)(OZ_BM >> 4) << 24);  // m_dim
}

// status[b] |= VG_STATUS_DEVICE when a pipeline wait of the GEMM timed out
__global__ void flag_status_kernel(const int* __restrict__ error_flag, int* __restrict__ status, long long B) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B && *error_flag) atomicOr(status + b, VG_STATUS_DEVICE);
}

template <int S> struct OzakiSmem {
  static constexpr int A_STAGE = S * OZ_BM * OZ_BK;
  static constexpr int B_STAGE = S * OZ_BN * OZ_BK;
  static constexpr int STAGE = A_STAGE + B_STAGE;
  static constexpr int FIT = (220 * 1024 - 2048) / STAGE;
  static constexpr int STAGES = FIT > 8 ? 8 : FIT;
  static_assert(STAGES >= 2, "two pipeline stages must fit");
  static constexpr int BYTES = STAGES * STAGE + 1024 /* alignment slack */ + 256 /* barriers */;
};

// CL > 1: clusters of CL CTAs along the column tiles (grid.y, a multiple of CL) work on the SAME row
// tile and K range, so they need the same A stage: each CTA fetches a share of its digit planes
// (ceil(S / CL) of them, the last one the rest: map_a / map_a_rest differ by the planes per box) and
// the load is multicast to every CTA of the cluster -- the batch operand's planes cross the L2 ->
// shared-memory path once per CL column tiles instead of once per tile (the kernel was bound there:
// 7 TB/s of L2 reads at 54-70 % tensor-pipe activity, profiles/r02_ncu_full_c2_raw.csv).  A stage is
// free again when the MMAs of ALL CL CTAs have read it: every CTA's commit arrives on the empty
// barrier of every CTA.
template <int S, int CL>
__global__ void __launch_bounds__(OZ_THREADS, 1)
ozaki_gemm_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_a_rest,
                  const __grid_constant__ CUtensorMap map_b, OzakiArgs g) {
  using L = OzakiSmem<S>;
  extern __shared__ uint8_t oz_raw[];
  const unsigned raw = oz_smem_u32(oz_raw);
  const unsigned base = (raw + 1023u) & ~1023u;                  // swizzled tiles want 1024-byte alignment
  uint8_t* aligned = oz_raw + (base - raw);
  unsigned long long* bars = reinterpret_cast<unsigned long long*>(aligned + L::STAGES * L::STAGE);
  const unsigned bar0 = base + L::STAGES * L::STAGE;
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (L::STAGES + s); };
  const unsigned tfull_bar = bar0 + 8u * (2 * L::STAGES);
  unsigned* tmem_slot = reinterpret_cast<unsigned*>(bars + 2 * L::STAGES + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // m tiles vary fastest: a wave of CTAs covers every row block of a few column blocks, so the
  // operator's digit planes (the larger operand) stream from HBM once
  const int tile_m = blockIdx.x, tile_n = blockIdx.y, split = blockIdx.z;
  const int kb0 = split * g.kb_per_split;
  const int nkb = min(g.kb_total, kb0 + g.kb_per_split) - kb0;

  // this CTA's share of the A planes (CL > 1)
  constexpr int A_PER = (S + CL - 1) / CL;
  const int crank = CL > 1 ? (int)oz_cluster_rank() : 0;
  const int a_p0 = crank * A_PER;
  const int a_cnt = min(A_PER, S - a_p0);           // <= 0: nothing to fetch
  constexpr unsigned short CL_MASK = (unsigned short)((1u << CL) - 1u);

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&map_a_rest) : "memory");
    asm volatile("prefetch.tensormap [%0];\n" ::"l"(&map_b) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < L::STAGES; ++s) { oz_mbar_init(full_bar(s), 1); oz_mbar_init(empty_bar(s), CL); }
    oz_mbar_init(tfull_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  if (warp == 2) oz_tmem_alloc(oz_smem_u32(tmem_slot), OZ_TMEM_COLS);
  oz_fence_before();
  __syncthreads();
  if (CL > 1) oz_cluster_sync();      // every CTA's barriers exist before a peer signals them
  oz_fence_after();
  const unsigned tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer: one box of S digit planes per operand per K block =====
    for (int i = 0; i < nkb; ++i) {
      const int s = i % L::STAGES;
      const unsigned ph = (unsigned)(i / L::STAGES) & 1u;
      if (!oz_mbar_wait(empty_bar(s), ph ^ 1u, g.error_flag)) break;
      oz_mbar_expect_tx(full_bar(s), (unsigned)L::STAGE);
      const unsigned dst = base + s * L::STAGE;
      if (CL == 1) {
        oz_tma_load_3d(dst, &map_a, full_bar(s), (kb0 + i) * OZ_BK, tile_m * OZ_BM, 0);
      } else if (a_cnt > 0) {
        oz_tma_load_3d_mc(dst + a_p0 * (OZ_BM * OZ_BK), a_cnt == A_PER ? &map_a : &map_a_rest, full_bar(s),
                          (kb0 + i) * OZ_BK, tile_m * OZ_BM, a_p0, CL_MASK);
      }
      oz_tma_load_3d(dst + L::A_STAGE, &map_b, full_bar(s), (kb0 + i) * OZ_BK, tile_n * OZ_BN, 0);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer: diagonal d accumulates sum_{t+u=d} A_t B_u^T in TMEM columns [d*64, d*64+64).
    // The planes B_u of a stage lie one after the other (64 rows of 64 bytes each), and plane u of
    // the product with A_t lands in accumulator t+u, so ONE instruction of N = 64 m columns does
    // A_t x [B_u .. B_u+m-1] into the m adjacent accumulators t+u .. t+u+m-1.  That reads the A
    // tile from shared memory once per <= 256 columns instead of once per 64: 10 instructions per
    // 32-byte K step for S = 7 instead of 28, ~110 B/clk of operand reads instead of 192 (the
    // shared-memory port moves 128 B/clk, a 128 x 64 x 32 instruction takes 32 clk).
    for (int i = 0; i < nkb; ++i) {
      const int s = i % L::STAGES;
      const unsigned ph = (unsigned)(i / L::STAGES) & 1u;
      if (!oz_mbar_wait(full_bar(s), ph, g.error_flag)) break;
      oz_fence_after();
      const unsigned a0 = base + s * L::STAGE, b0 = a0 + L::A_STAGE;
#pragma unroll
      for (int ks = 0; ks < OZ_BK / 32; ++ks) {
#pragma unroll
        for (int t = 0; t < S; ++t) {
          const int m = S - t;                              // planes u = 0 .. m-1 pair with A_t
          const int m0 = m <= 4 ? m : (m + 1) / 2;          // first instruction, then the rest
          const unsigned long long ad = oz_smem_desc(a0 + t * (OZ_BM * OZ_BK) + ks * 32);
          const unsigned acc = (i > 0 || ks > 0 || t > 0) ? 1u : 0u;
          oz_mma_i8(tmem_base + t * OZ_BN, ad, oz_smem_desc(b0 + ks * 32), oz_instr_desc(m0 * OZ_BN), acc);
          if (m > m0)
            oz_mma_i8(tmem_base + (t + m0) * OZ_BN, ad, oz_smem_desc(b0 + m0 * (OZ_BN * OZ_BK) + ks * 32),
                      oz_instr_desc((m - m0) * OZ_BN), acc);
        }
      }
      // frees the stage when its MMAs have read it (CL > 1: in every CTA that multicasts into it)
      if (CL == 1) oz_mma_commit(empty_bar(s));
      else oz_mma_commit_mc(empty_bar(s), CL_MASK);
    }
    oz_mma_commit(tfull_bar);               // accumulators complete
  } else if (warp >= 4) {
    // ===== epilogue: thread = one row (TMEM lane), 16 columns at a time =====
    const int q = warp & 3;                 // a warp may only touch TMEM lanes [32 (warp % 4), +32)
    const int row = tile_m * OZ_BM + q * 32 + lane;
    if (oz_mbar_wait(tfull_bar, 0u, g.error_flag)) {
      oz_fence_after();
      const int ea = row < g.M ? g.expA[row] : 0;
      const bool bad_row = ea == OZ_NONFINITE;
      const double rs = bad_row ? 0.0 : __longlong_as_double((long long)(1023 + ea - 12) << 52);
      double* out = g.C + (long long)split * g.split_stride + (long long)row * g.ldc;
#pragma unroll
      for (int c = 0; c < OZ_BN / 16; ++c) {
        unsigned P[S][16];
#pragma unroll
        for (int d = 0; d < S; ++d)
          oz_tmem_ld16(tmem_base + ((unsigned)(q * 32) << 16) + (unsigned)(d * OZ_BN + c * 16), P[d]);
        oz_tmem_ld_wait();
        if (row < g.M) {
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            double v = (double)(int)P[S - 1][j];
#pragma unroll
            for (int d = S - 2; d >= 0; --d) v = fma(v, 0.0078125, (double)(int)P[d][j]);
            const int col = tile_n * OZ_BN + c * 16 + j;
            if (col < g.N) {
              const int eb = g.expB[col];
              const double cs = eb == OZ_NONFINITE ? __longlong_as_double(0x7ff8000000000000LL)
                                                   : __longlong_as_double((long long)(1023 + eb) << 52);
              out[col] = bad_row ? __longlong_as_double(0x7ff8000000000000LL) : v * rs * cs;
            }
          }
        }
      }
    }
  }
  oz_fence_before();
  __syncthreads();
  if (CL > 1) oz_cluster_sync();      // no CTA leaves while a peer may still write to it or signal it
  if (warp == 2) oz_tmem_dealloc(tmem_base, OZ_TMEM_COLS);
}

}  // namespace vg
