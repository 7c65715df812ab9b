// Moment form of K1+K2a+K2b for evaluations whose only sampled P(k,mu) parameters are the
// Kaiser / HCD amplitudes.
//
// With F = F_hcd(k_par) and S = the product of every other factor both fixed for the batch,
// the reference's P(k,mu) (vega/power_spectrum.py:87-196 with :198-222 and :263-309) is
//   P = pk_lin(k) S(mu,k) [(a0 + h0 F) + (a1 + h1 F) mu^2]_t0 [(a0 + h0 F) + (a1 + h1 F) mu^2]_t1
// i.e. a polynomial in (F, mu^2) whose 9 coefficients c[a][j] (powers F^a mu^2j, a,j = 0..2)
// depend only on the parameter vector.  The mu projection (vega/pktoxi.py:79-97), the
// FFTLog (mcfit.P2xi / Pk2Mp) and the cubic spline are linear, so
//   knots_l[b] = sum_{a,j} c[a][j](theta_b) * K_l[a][j],
//   K_l[a][j]  = (Hankel+spline operator)_l  applied to  pk_lin(k) sum_mu w_l(mu) S F^a mu^2j
// with K built once per plan by the same DGEMM kernel the per-vector path uses.  The hot
// path of such an evaluation is then 3 or 9 FMAs per knot instead of a pass over the
// 814 000-point grid and a Hankel GEMM row.  Evaluations with velocity dispersion,
// sampled L0_hcd, sampled non-linear parameters, UV-dependent bias or the effective-bias
// division of fast_metals keep the full per-point pass (vg_pk.cuh).
#pragma once
#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_pk.cuh"

namespace vg {

constexpr int MOM_MAX = 9;   // F^a mu^2j, index a*3 + j

// Host-side eligibility test (mirrors the kernel-class selection in vg_plan_finalize)
inline bool moment_eligible(const vg_eval_desc& d) {
  if (d.dyn_mask != 0) return false;
  const bool uv = (d.uvb || d.heii) && (d.uv_on[0] || d.uv_on[1]);
  if (uv) return false;
  const bool hcd = d.hcd_model && (d.hcd_on[0] || d.hcd_on[1]);
  if (hcd && d.hcd_grid < 0) return false;           // L0_hcd sampled: F differs per vector
  if (hcd && d.fast_metals) return false;            // Kaiser / effective bias is rational in F
  return true;
}
inline int moment_count(const vg_eval_desc& d) {
  return (d.hcd_model && (d.hcd_on[0] || d.hcd_on[1])) ? 9 : 3;
}

// Plan time: mom[j][m][l][nk_pad] = pk_lin(k) * sum_i wleg[l][i] S(mu_i,k) F(mu_i,k)^a mu_i^2j
// for the moment evaluations ids[j]; m = a*3 + j2.  grid = (ceil(nk/128), n_ids), block 128.
__global__ void __launch_bounds__(128)
moment_grid_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                   double* __restrict__ mom) {
  const int ik = blockIdx.x * 128 + threadIdx.x;
  if (ik >= g.nk) return;
  const vg_eval_desc& d = evals[ids[blockIdx.y]];
  const bool hcd = d.hcd_model && (d.hcd_on[0] || d.hcd_on[1]);
  const int na = hcd ? 3 : 1;
  const double* S = d.static_grid >= 0 ? g.grids[d.static_grid] + ik : nullptr;
  const double* F = hcd ? g.grids[d.hcd_grid] + ik : nullptr;
  double acc[MOM_MAX][PK_MAX_ELL];
#pragma unroll
  for (int m = 0; m < MOM_MAX; ++m)
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) acc[m][l] = 0.0;
  for (int i = 0; i < g.nmu; ++i) {
    const double mu = g.mu[i], mu2 = mu * mu;
    const double s = S ? S[(long long)i * g.nk] : 1.0;
    const double f = F ? F[(long long)i * g.nk] : 0.0;
    double w[PK_MAX_ELL];
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) w[l] = l < g.n_ell ? g.wleg[l * g.nmu + i] : 0.0;
    double fa = s;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      if (a < na) {
        double v = fa;
#pragma unroll
        for (int j = 0; j < 3; ++j) {
#pragma unroll
          for (int l = 0; l < PK_MAX_ELL; ++l) acc[a * 3 + j][l] = fma(w[l], v, acc[a * 3 + j][l]);
          v *= mu2;
        }
        fa *= f;
      }
    }
  }
  const double pk_lin = g.pk_lin[d.pk_lin_sel][ik];
  for (int m = 0; m < MOM_MAX; ++m)
    for (int l = 0; l < g.n_ell; ++l)
      mom[(((long long)blockIdx.y * MOM_MAX + m) * g.n_ell + l) * g.nk_pad + ik] = acc[m][l] * pk_lin;
}

// Coefficients of F^a mu^2j in the product of the two tracers' Kaiser terms
__device__ __forceinline__ void moment_coefs(const KaiserCoef& c, double* q) {
  const double A0 = c.a0[0], H0 = c.h0[0], A1 = c.a1[0], H1 = c.h1[0];
  const double B0 = c.a0[1], G0 = c.h0[1], B1 = c.a1[1], G1 = c.h1[1];
  // mu^0: (A0 + H0 F)(B0 + G0 F)
  q[0 * 3 + 0] = A0 * B0;
  q[1 * 3 + 0] = A0 * G0 + H0 * B0;
  q[2 * 3 + 0] = H0 * G0;
  // mu^2: (A0 + H0 F)(B1 + G1 F) + (A1 + H1 F)(B0 + G0 F)
  q[0 * 3 + 1] = A0 * B1 + A1 * B0;
  q[1 * 3 + 1] = A0 * G1 + H0 * B1 + A1 * G0 + H1 * B0;
  q[2 * 3 + 1] = H0 * G1 + H1 * G0;
  // mu^4: (A1 + H1 F)(B1 + G1 F)
  q[0 * 3 + 2] = A1 * B1;
  q[1 * 3 + 2] = A1 * G1 + H1 * B1;
  q[2 * 3 + 2] = H1 * G1;
}

// Hot path: tables[b][j][x] = sum_m c_m(theta_b, eval ids[j]) * ktab[j][m][x],  x over the
// n_ell*2*n_out_pad knot values and second derivatives of the evaluation.
// grid = (ceil(row_len / (256*MOM_XPT)), n_ids, ceil(B / MOM_NB)), block 256.
constexpr int MOM_NB = 4, MOM_XPT = 2;
__global__ void __launch_bounds__(256)
moment_tables_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids, int n_ids,
                     const double* __restrict__ pfull, int npf, int B, const double* __restrict__ ktab,
                     int row_len, double* __restrict__ tables) {
  __shared__ double cq[MOM_NB][MOM_MAX];
  __shared__ int s_nm;
  const int j = blockIdx.y, b0 = blockIdx.z * MOM_NB;
  if (threadIdx.x < MOM_NB) {
    const vg_eval_desc& d = evals[ids[j]];
    const double* p = pfull + (long long)min(b0 + (int)threadIdx.x, B - 1) * npf;
    KaiserCoef kc = kaiser_setup(d, p, 1.0);   // k only enters through the UV terms (excluded)
    if (d.same_tracer) { kc.a0[1] = kc.a0[0]; kc.a1[1] = kc.a1[0]; kc.h0[1] = kc.h0[0]; kc.h1[1] = kc.h1[0]; }
    double q[MOM_MAX];
    moment_coefs(kc, q);
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[threadIdx.x][m] = q[m];
    if (threadIdx.x == 0) s_nm = (d.hcd_model && (d.hcd_on[0] || d.hcd_on[1])) ? 3 : 1;
  }
  __syncthreads();
  const int na = s_nm;
  const double* kt = ktab + (long long)j * MOM_MAX * row_len;
#pragma unroll
  for (int u = 0; u < MOM_XPT; ++u) {
    const int x = (blockIdx.x * MOM_XPT + u) * 256 + threadIdx.x;
    if (x >= row_len) break;
    double acc[MOM_NB];
#pragma unroll
    for (int e = 0; e < MOM_NB; ++e) acc[e] = 0.0;
    for (int a = 0; a < na; ++a)
#pragma unroll
      for (int jj = 0; jj < 3; ++jj) {
        const int m = a * 3 + jj;
        const double v = __ldg(kt + (long long)m * row_len + x);
#pragma unroll
        for (int e = 0; e < MOM_NB; ++e) acc[e] = fma(cq[e][m], v, acc[e]);
      }
#pragma unroll
    for (int e = 0; e < MOM_NB; ++e)
      if (b0 + e < B) tables[((long long)(b0 + e) * n_ids + j) * row_len + x] = acc[e];
  }
}

// Moment form of the xi stage for evaluations whose coordinates and per-bin factors are the
// same for every vector (vg_eval_desc::static_xi): the rescale / spline read-out / Legendre
// sum / evolution factors (vega/correlation_func.py:117-236, vega/pktoxi.py:141-163) are
// linear in the knot tables, so they are applied ONCE, in vg_plan_finalize, to the static
// moment tables:  X[j][m][bin] = xi stage of evaluation ids[j] on K[owner(j)][m].  Hot path:
//   out[b][slot(j)][bin] = w_j(theta_b) * sum_m c_m(theta_b; owner(j)) * X[j][m][bin]
// -- 3 or 9 FMAs per bin instead of a knot-table row and a gather.  `owners[j]` is the
// evaluation whose P(k,mu) the table belongs to (j itself unless it shares a table).
// grid = (ceil(ld / 256), n_ids, ceil(B / MXI_NB)), block 256.
constexpr int MXI_NB = 8;
__global__ void __launch_bounds__(256)
moment_xi_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                 const int* __restrict__ owners, const double* __restrict__ pfull, int npf, int B,
                 const double* __restrict__ X, const int* __restrict__ rows, int ld,
                 double* __restrict__ core, double* __restrict__ xmet,
                 const int* __restrict__ metal_map, int n_metal) {
  __shared__ double cq[MXI_NB][MOM_MAX];
  __shared__ int s_na, s_slot;
  const int j = blockIdx.y, b0 = blockIdx.z * MXI_NB;
  if (threadIdx.x < MXI_NB) {
    const vg_eval_desc& d = evals[ids[j]];
    const vg_eval_desc& o = evals[owners[j]];
    const double* p = pfull + (long long)min(b0 + (int)threadIdx.x, B - 1) * npf;
    KaiserCoef kc = kaiser_setup(o, p, 1.0);
    if (o.same_tracer) { kc.a0[1] = kc.a0[0]; kc.a1[1] = kc.a1[0]; kc.h0[1] = kc.h0[0]; kc.h1[1] = kc.h1[0]; }
    double q[MOM_MAX];
    moment_coefs(kc, q);
    double w = 1.0;                       // weight_const is part of X
    if (d.weight_bias) {
      double b1, be1, b2, be2;
      tracer_bias_beta(d, p, 0, b1, be1);
      tracer_bias_beta(d, p, 1, b2, be2);
      w = b1 * b2;
    }
    if (d.idx_weight_extra >= 0) w *= p[d.idx_weight_extra];
    if (d.idx_weight_extra2 >= 0) w *= p[d.idx_weight_extra2];
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[threadIdx.x][m] = w * q[m];
    if (threadIdx.x == 0) {
      s_na = (o.hcd_model && (o.hcd_on[0] || o.hcd_on[1])) ? 3 : 1;
      s_slot = d.out_slot;
    }
  }
  __syncthreads();
  const int x = blockIdx.x * 256 + threadIdx.x;
  if (x >= ld) return;
  const int na = s_na;
  const double* xt = X + (long long)(rows ? rows[j] : j) * MOM_MAX * ld + x;
  double acc[MXI_NB];
#pragma unroll
  for (int e = 0; e < MXI_NB; ++e) acc[e] = 0.0;
  for (int a = 0; a < na; ++a)
#pragma unroll
    for (int jj = 0; jj < 3; ++jj) {
      const int m = a * 3 + jj;
      const double v = __ldg(xt + (long long)m * ld);
#pragma unroll
      for (int e = 0; e < MXI_NB; ++e) acc[e] = fma(cq[e][m], v, acc[e]);
    }
  const int slot = s_slot;
#pragma unroll
  for (int e = 0; e < MXI_NB; ++e) {
    const int b = b0 + e;
    if (b >= B) break;
    double* out = slot < 2 ? core + ((long long)slot * B + b) * ld
                           : xmet + ((long long)b * n_metal + metal_map[slot - 2]) * ld;
    out[x] = acc[e];
  }
}

// Folded metal terms (VG_OPT_FOLD_METALS): a metal evaluation whose xi stage is in moment
// form is a 3/9-term combination of static vectors X[j][m], so its matrix product is the same
// combination of Y[j][m] = M_c X[j][m], built once at finalize.  One pass sums them all:
//   out[b][bin] = sum_j w_j(theta_b) sum_m c_m(theta_b; owner(j)) Y[j][m][bin]
// grid = (ceil(ld / 256), ceil(B / MXI_NB)), block 256, dynamic smem n_ids*MXI_NB*MOM_MAX doubles
// + n_ids ints.
__global__ void __launch_bounds__(256)
moment_fold_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                   const int* __restrict__ owners, int n_ids, const double* __restrict__ pfull, int npf,
                   int B, const double* __restrict__ Y, int ld, double* __restrict__ out) {
  extern __shared__ double fold_smem[];
  double* cq = fold_smem;                                   // [n_ids][MOM_MAX][MXI_NB]
  int* s_na = reinterpret_cast<int*>(fold_smem + (size_t)n_ids * MOM_MAX * MXI_NB);
  const int b0 = blockIdx.y * MXI_NB;
  for (int t = threadIdx.x; t < n_ids * MXI_NB; t += 256) {
    const int j = t / MXI_NB, e = t % MXI_NB;
    const vg_eval_desc& d = evals[ids[j]];
    const vg_eval_desc& o = evals[owners[j]];
    const double* p = pfull + (long long)min(b0 + e, B - 1) * npf;
    KaiserCoef kc = kaiser_setup(o, p, 1.0);
    if (o.same_tracer) { kc.a0[1] = kc.a0[0]; kc.a1[1] = kc.a1[0]; kc.h0[1] = kc.h0[0]; kc.h1[1] = kc.h1[0]; }
    double q[MOM_MAX];
    moment_coefs(kc, q);
    double w = 1.0;                       // weight_const is part of Y
    if (d.weight_bias) {
      double b1, be1, b2, be2;
      tracer_bias_beta(d, p, 0, b1, be1);
      tracer_bias_beta(d, p, 1, b2, be2);
      w = b1 * b2;
    }
    if (d.idx_weight_extra >= 0) w *= p[d.idx_weight_extra];
    if (d.idx_weight_extra2 >= 0) w *= p[d.idx_weight_extra2];
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[((size_t)j * MOM_MAX + m) * MXI_NB + e] = w * q[m];
    if (e == 0) s_na[j] = (o.hcd_model && (o.hcd_on[0] || o.hcd_on[1])) ? 3 : 1;
  }
  __syncthreads();
  const int x = blockIdx.x * 256 + threadIdx.x;
  if (x >= ld) return;
  double acc[MXI_NB];
#pragma unroll
  for (int e = 0; e < MXI_NB; ++e) acc[e] = 0.0;
  for (int j = 0; j < n_ids; ++j) {
    const double* yt = Y + (long long)j * MOM_MAX * ld + x;
    const int nm = 3 * s_na[j];
    for (int m = 0; m < nm; ++m) {
      const double v = __ldg(yt + (long long)m * ld);
      const double* cj = cq + ((size_t)j * MOM_MAX + m) * MXI_NB;
#pragma unroll
      for (int e = 0; e < MXI_NB; ++e) acc[e] = fma(cj[e], v, acc[e]);
    }
  }
#pragma unroll
  for (int e = 0; e < MXI_NB; ++e) {
    const int b = b0 + e;
    if (b >= B) break;
    out[(long long)b * ld + x] = acc[e];
  }
}

}  // namespace vg
