// Elementwise stages around the K4 GEMMs: peak/smooth/metal combination, broadband
// polynomials, residual against the data and the chi2 reduction.
//
// Restates vega/model.py:119-149,176-186 (component assembly; because the distortion
// matrix is linear the peak and smooth components are combined BEFORE it:
//   xi = mul_post * DM[ mul_pre*(bao_amp*peak + smooth + metals + inst) + (1+bao_amp)*add_pre ]
//        + (1+bao_amp)*add_post,
// which is algebraically the reference's bao_amp*xi_peak + xi_smooth with every broadband
// term applied to both components), vega/broadband_poly.py:74-198 and the quadratic form
// of vega/vega_interface.py:294-319.
#pragma once
#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_gemm.cuh"
#include "vg_pk.cuh"

namespace vg {

struct BroadbandDev {
  int n_terms;
  const vg_bb_term* terms;
  const double* design_pre;  int rows_pre;    // [rows][n_model]
  const double* design_post; int rows_post;   // [rows][n_dist]
  const double* sky_pre;                      // [2][n_model]: rt, window
  const double* sky_post;                     // [2][n_dist]
};

// value of the broadband at bin i for position pos (0 pre-mul 1 pre-add 2 post-mul 3 post-add)
__device__ __forceinline__ double bb_eval(const BroadbandDev& bb, int pos, const double* p,
                                          int i, int n_bins) {
  const bool mul = (pos & 1) == 0;
  const double* design = pos < 2 ? bb.design_pre : bb.design_post;
  const double* sky = pos < 2 ? bb.sky_pre : bb.sky_post;
  double total = mul ? 1.0 : 0.0;
  for (int t = 0; t < bb.n_terms; ++t) {
    const vg_bb_term& term = bb.terms[t];
    if (term.position != pos) continue;
    double poly = 0.0;
    if (term.kind == 0) {
      for (int c = 0; c < term.n_coef; ++c)
        poly += p[term.idx_coef[c]] * design[(long long)(term.design_offset + c) * n_bins + i];
    } else {
      const double scale = p[term.idx_sky[0]], sigma = p[term.idx_sky[1]];
      const double rt = sky[i];
      poly = sky[n_bins + i] != 0.0
                 ? scale / (sigma * sqrt(2.0 * M_PI)) * exp(-0.5 * (rt / sigma) * (rt / sigma))
                 : 0.0;
    }
    if (mul) total *= (1.0 + poly); else total += poly;
  }
  return total;
}

// xmet_sum[b][i] = sum_c xmet[b][c*ld + i]  (identity metal matrices, vega/data.py:683-684)
__global__ void metal_sum_kernel(const double* __restrict__ xmet, int n_metal, int ld, int n,
                                 long long B, double* __restrict__ out) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * ld) return;
  long long b = idx / ld;
  int i = (int)(idx % ld);
  double s = 0.0;
  if (i < n)
    for (int c = 0; c < n_metal; ++c) s += xmet[(b * n_metal + c) * ld + i];
  out[idx] = s;
}

struct CombineArgs {
  const double* core;      // [2][B][ld]
  const double* met;       // [B][ld] or null
  const double* met_fold;  // [B][ld] or null: metal terms folded into static vectors
  const double* inst;      // [n] or null
  int idx_inst_amp; double inst_amp_default;
  int idx_bao_amp;
  BroadbandDev bb;
  const double* pfull; int npf;
  int n, ld; long long B;
  double* x;               // [B][ld]
  // fast_metals metal x metal terms: sum_f b1*b2(*extra)(theta) * frozen[f][i]
  const double* frozen; int n_frozen; const int* frozen_ids; const vg_eval_desc* evals;
  int direct;              // 1: x = full component alone (Model.compute_direct), no bao_amp
};

// weight of a metal term from its bias parameters (vega/metals.py:289-313)
__device__ __forceinline__ double metal_bias_weight(const vg_eval_desc& d, const double* p) {
  double b1, be1, b2, be2;
  tracer_bias_beta(d, p, 0, b1, be1);
  tracer_bias_beta(d, p, 1, b2, be2);
  double w = b1 * b2;
  if (d.idx_weight_extra >= 0) w *= p[d.idx_weight_extra];
  if (d.idx_weight_extra2 >= 0) w *= p[d.idx_weight_extra2];
  return w;
}

// rp-only metal matrices (new_metals with rp_only_metal_mats, vega/metals.py:354-358):
//   met[b][rp][rt] = sum_c sum_rp' A_c[rp][rp'] xmet[b][c][rp'][rt]
// -- Np^2 numbers per metal correlation instead of (Np Nt)^2, and 2 Np^2 Nt flops per vector
// instead of 2 (Np Nt)^2.  A small batched GEMM (M = Np, N = Nt, K = n_metal*Np per vector):
// CTA tile 40 rp x 32 rt, K in chunks of 32 staged in shared memory (A chunk [40][32+1],
// X chunk [32][32]); a thread owns one rt column and 5 rp rows; the A entries are
// warp-uniform (broadcast) shared loads.  grid = (rt tiles, rp chunks, B), block 256.
constexpr int MRP_TM = 40, MRP_TN = 32, MRP_TK = 32, MRP_R = 5;
constexpr int MRP_MAX_NP = 4096;
__global__ void __launch_bounds__(256)
metal_rp_kernel(const double* __restrict__ xmet, int n_metal, int ld, const double* __restrict__ A,
                int np, int nt, double* __restrict__ met) {
  __shared__ double As[MRP_TM][MRP_TK + 1];
  __shared__ double Xs[MRP_TK][MRP_TN];
  const long long b = blockIdx.z;
  const int rt0 = blockIdx.x * MRP_TN, rp0 = blockIdx.y * MRP_TM;
  const int tid = threadIdx.x, lane = tid & 31, g = tid >> 5;      // g: 8 groups of 5 rows
  double acc[MRP_R];
#pragma unroll
  for (int r = 0; r < MRP_R; ++r) acc[r] = 0.0;
  for (int c = 0; c < n_metal; ++c) {
    const double* Ac = A + (long long)c * np * np;
    const double* x = xmet + (b * n_metal + c) * ld;
    for (int k0 = 0; k0 < np; k0 += MRP_TK) {
      __syncthreads();
      for (int i = tid; i < MRP_TM * MRP_TK; i += 256) {
        const int r = i / MRP_TK, k = i % MRP_TK;
        As[r][k] = (rp0 + r < np && k0 + k < np) ? __ldg(Ac + (long long)(rp0 + r) * np + k0 + k) : 0.0;
      }
      for (int i = tid; i < MRP_TK * MRP_TN; i += 256) {
        const int k = i / MRP_TN, t = i % MRP_TN;
        Xs[k][t] = (k0 + k < np && rt0 + t < nt) ? x[(k0 + k) * nt + rt0 + t] : 0.0;
      }
      __syncthreads();
#pragma unroll 8
      for (int k = 0; k < MRP_TK; ++k) {
        const double xv = Xs[k][lane];
#pragma unroll
        for (int r = 0; r < MRP_R; ++r) acc[r] = fma(As[g * MRP_R + r][k], xv, acc[r]);
      }
    }
  }
  const int rt = rt0 + lane;
  if (rt < nt) {
#pragma unroll
    for (int r = 0; r < MRP_R; ++r) {
      const int rp = rp0 + g * MRP_R + r;
      if (rp < np) met[b * ld + rp * nt + rt] = acc[r];
    }
  }
  // pad columns stay finite (written by the first CTA of each vector)
  if (blockIdx.x == 0 && blockIdx.y == 0)
    for (int i = np * nt + tid; i < ld; i += 256) met[b * ld + i] = 0.0;
}

// The same contraction on the FP64 tensor pipe (DMMA m8n8k4, tcgen05 has no FP64 kind).  Per
// vector it is a small GEMM  met[rp][rt] = [A_0 | A_1 | ...][rp][c rp'] . X[c rp'][rt]  with
// M = Np (75 / 150), N = Nt (75), K = n_metal Np.  CTA tile 80 rp x 40 rt x 32 k, 5 warps, a warp
// owns 16 rp rows x the 40 rt columns (2 x 5 DMMA fragments: one B fragment feeds two DMMAs);
// shared rows padded to 36 / 44 doubles so the 64-bit fragment loads of a half-warp fall in 16
// distinct bank pairs.  34 KB of shared memory and ~80 registers per CTA thread: 5-6 CTAs per SM
// hide the refill (next chunk prefetched into registers during the DMMAs, two barriers per
// 32 k), and the half-width tile keeps the
// wave tail short (a 451-vector Hessian stencil is 902 / 1804 CTAs).  Summation order over k is the FMA kernel's (ascending), per
// output element, so results are independent of the batch composition.
constexpr int MRD_TM = 80, MRD_TN = 40, MRD_TK = 32, MRD_SA = MRD_TK + 4, MRD_SX = MRD_TN + 4;
constexpr int MRD_THREADS = 160;
__global__ void __launch_bounds__(MRD_THREADS, 3)
metal_rp_dmma_kernel(const double* __restrict__ xmet, int n_metal, int ld, const double* __restrict__ A,
                     int np, int nt, double* __restrict__ met) {
  __shared__ double As[MRD_TM * MRD_SA];
  __shared__ double Xs[MRD_TK * MRD_SX];
  const long long b = blockIdx.z;
  const int rt0 = blockIdx.x * MRD_TN, rp0 = blockIdx.y * MRD_TM;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int gid = lane >> 2, tig = lane & 3;
  double acc[2][MRD_TN / 8][2];
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int n = 0; n < MRD_TN / 8; ++n) acc[m][n][0] = acc[m][n][1] = 0.0;
  const double* a_lo = As + (warp * 16 + gid) * MRD_SA + tig;
  const double* b_lo = Xs + tig * MRD_SX + gid;
  // chunks q = (matrix c, 32 k) in order; the next chunk's operands are fetched into registers
  // while the DMMAs of the current one run, and stored to shared memory after the barrier
  constexpr int NA = MRD_TM * MRD_TK / MRD_THREADS, NX = MRD_TK * MRD_TN / MRD_THREADS;
  static_assert(NA * MRD_THREADS == MRD_TM * MRD_TK && NX * MRD_THREADS == MRD_TK * MRD_TN, "fill mapping");
  const int nkc = (np + MRD_TK - 1) / MRD_TK, n_chunks = n_metal * nkc;
  double ra[NA], rx[NX];
  auto fetch = [&](int q) {
    const int c = q / nkc, k0 = (q - c * nkc) * MRD_TK;
    const double* Ac = A + (long long)c * np * np;
    const double* x = xmet + (b * n_metal + c) * ld;
#pragma unroll
    for (int j = 0; j < NA; ++j) {
      const int i = tid + j * MRD_THREADS, r = i / MRD_TK, k = i % MRD_TK;
      ra[j] = (rp0 + r < np && k0 + k < np) ? __ldg(Ac + (long long)(rp0 + r) * np + k0 + k) : 0.0;
    }
#pragma unroll
    for (int j = 0; j < NX; ++j) {
      const int i = tid + j * MRD_THREADS, k = i / MRD_TN, t = i % MRD_TN;
      rx[j] = (k0 + k < np && rt0 + t < nt) ? __ldg(x + (k0 + k) * nt + rt0 + t) : 0.0;
    }
  };
  fetch(0);
  for (int q = 0; q < n_chunks; ++q) {
    const int k0 = (q % nkc) * MRD_TK;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NA; ++j) {
      const int i = tid + j * MRD_THREADS;
      As[(i / MRD_TK) * MRD_SA + i % MRD_TK] = ra[j];
    }
#pragma unroll
    for (int j = 0; j < NX; ++j) {
      const int i = tid + j * MRD_THREADS;
      Xs[(i / MRD_TN) * MRD_SX + i % MRD_TN] = rx[j];
    }
    __syncthreads();
    if (q + 1 < n_chunks) fetch(q + 1);
#pragma unroll
    for (int kk = 0; kk < MRD_TK; kk += 4) {
      if (k0 + kk >= np) break;            // tail chunk of a matrix (75 = 32 + 32 + 11): CTA-uniform
      const double a0 = a_lo[kk], a1 = a_lo[8 * MRD_SA + kk];
#pragma unroll
      for (int n = 0; n < MRD_TN / 8; ++n) {
        const double bf = b_lo[kk * MRD_SX + n * 8];
        dmma884(acc[0][n][0], acc[0][n][1], a0, bf);
        dmma884(acc[1][n][0], acc[1][n][1], a1, bf);
      }
    }
  }
#pragma unroll
  for (int m = 0; m < 2; ++m) {
    const int rp = rp0 + warp * 16 + m * 8 + gid;
    if (rp >= np) continue;
#pragma unroll
    for (int n = 0; n < MRD_TN / 8; ++n)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int rt = rt0 + n * 8 + tig * 2 + j;
        if (rt < nt) met[b * ld + rp * nt + rt] = acc[m][n][j];
      }
  }
  if (blockIdx.x == 0 && blockIdx.y == 0)
    for (int i = np * nt + tid; i < ld; i += MRD_THREADS) met[b * ld + i] = 0.0;
}

__global__ void combine_pre_kernel(CombineArgs a) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.B * a.ld) return;
  long long b = idx / a.ld;
  int i = (int)(idx % a.ld);
  if (i >= a.n) { a.x[idx] = 0.0; return; }
  const double* p = a.pfull + b * a.npf;
  const double amp = a.direct ? 0.0 : par(p, a.idx_bao_amp, 1.0);
  double v = a.direct ? a.core[a.B * a.ld + idx] : amp * a.core[idx] + a.core[a.B * a.ld + idx];
  if (a.met) v += a.met[idx];
  if (a.met_fold) v += a.met_fold[idx];
  for (int f = 0; f < a.n_frozen; ++f)
    v = fma(metal_bias_weight(a.evals[a.frozen_ids[f]], p), a.frozen[(long long)f * a.ld + i], v);
  if (a.inst) v += par(p, a.idx_inst_amp, a.inst_amp_default) * a.inst[i];
  if (a.bb.n_terms) {
    v *= bb_eval(a.bb, 0, p, i, a.n);
    v += (1.0 + amp) * bb_eval(a.bb, 1, p, i, a.n);
  }
  a.x[idx] = v;
}

struct PostArgs {
  const double* y; int ldy;   // distorted model rows; y[b][row]
  const double* y_add;        // nullable, same layout: a term already on the distorted bins (fused metal operator)
  const int* mask_idx;        // [n_out] distorted-grid bin of output j (null: j)
  int y_is_masked;            // 1: y row j already is bin mask_idx[j]
  BroadbandDev bb; int n_dist;
  const double* pfull; int npf; int idx_bao_amp;
  const double* data;         // [n_out] or null (model output)
  const int* mock_idx;        // per vector: row of `data` ([n_mock][data_stride]), null: row 0
  int data_stride;
  int direct;                 // 1: additive broadband enters once (no peak component)
  int n_out, ncols, ldo; long long B;   // columns j < ncols are written (zero for j >= n_out)
  double* out;                // [B][ldo]: model (data == null) or residual data - model
};

__global__ void post_kernel(PostArgs a) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= a.B * a.ncols) return;
  long long b = idx / a.ncols;
  int j = (int)(idx % a.ncols);
  double* o = a.out + b * a.ldo + j;
  if (j >= a.n_out) { *o = 0.0; return; }
  const double* p = a.pfull + b * a.npf;
  const int bin = a.mask_idx ? a.mask_idx[j] : j;
  double v = a.y[b * a.ldy + (a.y_is_masked ? j : bin)];
  if (a.y_add) v += a.y_add[b * a.ldy + (a.y_is_masked ? j : bin)];
  if (a.bb.n_terms) {
    const double amp = a.direct ? 0.0 : par(p, a.idx_bao_amp, 1.0);
    v *= bb_eval(a.bb, 2, p, bin, a.n_dist);
    v += (1.0 + amp) * bb_eval(a.bb, 3, p, bin, a.n_dist);
  }
  if (a.data) {
    const double* dvec = a.mock_idx ? a.data + (long long)a.mock_idx[b] * a.data_stride : a.data;
    *o = dvec[j] - v;
  } else {
    *o = v;
  }
}

// chi2[b] (+)= sum_j diff[b][j]^2   (identity covariance)
__global__ void __launch_bounds__(256) chi2_identity_kernel(const double* __restrict__ diff, int n, int ld,
                                                           double* __restrict__ chi2, int accumulate) {
  __shared__ double red[256];
  const long long b = blockIdx.x;
  double s = 0.0;
  for (int j = threadIdx.x; j < n; j += 256) { double v = diff[b * ld + j]; s = fma(v, v, s); }
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) chi2[b] = (accumulate ? chi2[b] : 0.0) + red[0];
}

// chi2[b] (+)= sum_t partial[b][t]  (row-dot partials of the C^-1 GEMM)
__global__ void chi2_partials_kernel(const double* __restrict__ partial, int n_tiles, long long B,
                                     double* __restrict__ chi2, int accumulate) {
  long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s = 0.0;
  for (int t = 0; t < n_tiles; ++t) s += partial[b * n_tiles + t];
  chi2[b] = (accumulate ? chi2[b] : 0.0) + s;
}

// chi2[b] += part[b]   (the correlations that ran on the second lane of a thin batch)
__global__ void add_vector_kernel(const double* __restrict__ part, long long B, double* __restrict__ chi2) {
  long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) chi2[b] += part[b];
}

// chi2[b] += sum_priors (theta - mean)^2 / sigma^2   (vega/vega_interface.py:423-446)
__global__ void prior_kernel(const double* __restrict__ pfull, int npf, const int* __restrict__ cols,
                             const double* __restrict__ mean, const double* __restrict__ sigma,
                             int n_prior, long long B, double* __restrict__ chi2) {
  long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double s = 0.0;
  for (int i = 0; i < n_prior; ++i) {
    const double dv = pfull[b * npf + cols[i]] - mean[i];
    s += dv * dv / (sigma[i] * sigma[i]);
  }
  chi2[b] += s;
}

// copy a [B][ld_src] block into a column window of a wider [B][ld_dst] buffer
__global__ void copy_cols_kernel(const double* __restrict__ src, int ld_src, int n, long long B,
                                 double* __restrict__ dst, int ld_dst, int col0) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * n) return;
  long long b = idx / n;
  int j = (int)(idx % n);
  dst[b * ld_dst + col0 + j] = src[b * ld_src + j];
}

}  // namespace vg
