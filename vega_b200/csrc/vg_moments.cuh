// Moment form of K1+K2a+K2b for evaluations whose only sampled P(k,mu) parameters are the
// Kaiser / HCD amplitudes.
//
// With F = F_hcd(k_par), G = G_uv(k) the scale-dependent UV / HeII bias term at FIXED UV parameters
// and S = the product of every other factor, all fixed for the batch, the reference's P(k,mu)
// (vega/power_spectrum.py:87-196 with :198-222, :224-261 and :263-309) is
//   P = pk_lin(k) S(mu,k) [(a0 + g G + h0 F) + (a1 + h1 F) mu^2]_t0 [(a0 + g G + h0 F) + (a1 + h1 F) mu^2]_t1
// (g = 1 for a tracer with the UV term, else 0: power_spectrum.py:224-261 adds G to the bias and
// rescales beta so that bias * beta is unchanged), i.e. a polynomial in (G, F, mu^2) whose 14
// coefficients depend only on the parameter vector: c[a][j] for F^a mu^2j (a, j = 0..2) and five
// more for G, G F, G^2 (mu^0) and G, G F (mu^2).  The mu projection (vega/pktoxi.py:79-97), the
// FFTLog (mcfit.P2xi / Pk2Mp) and the cubic spline are linear, so
//   knots_l[b] = sum_m c_m(theta_b) * K_l[m],
//   K_l[m]     = (Hankel+spline operator)_l  applied to  pk_lin(k) G(k)^p sum_mu w_l(mu) S F^a mu^2j
// with K built once per plan by the same DGEMM kernel the per-vector path uses.  The hot
// path of such an evaluation is then 3, 9 or 14 FMAs per knot instead of a pass over the
// 814 000-point grid and a Hankel GEMM row.  Evaluations with sampled L0_hcd, sampled non-linear
// or UV parameters, or the effective-bias division of fast_metals keep the full per-point pass
// (vg_pk.cuh); a sampled velocity dispersion goes through the tables at the end of this file.
#pragma once
#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_pk.cuh"

namespace vg {

// F^a mu^2j at index a*3 + j (0..8); 9: G, 10: G F, 11: G^2 (mu^0); 12: G mu^2, 13: G F mu^2
constexpr int MOM_MAX = 14;

__host__ __device__ inline bool moment_has_uv(const vg_eval_desc& d) {
  return (d.uvb || d.heii) && (d.uv_on[0] || d.uv_on[1]);
}
__host__ __device__ inline bool moment_has_hcd(const vg_eval_desc& d) {
  return d.hcd_model && (d.hcd_on[0] || d.hcd_on[1]);
}
// Host-side eligibility test (mirrors the kernel-class selection in vg_plan_finalize)
inline bool moment_eligible(const vg_eval_desc& d) {
  if (d.dyn_mask != 0) return false;
  if (moment_has_uv(d) && (!d.uv_static || d.fast_metals)) return false;   // sampled UV parameters: G differs
                                                     // per vector; fast_metals: beta alone is rational in G
  const bool hcd = moment_has_hcd(d);
  if (hcd && d.hcd_grid < 0) return false;           // L0_hcd sampled: F differs per vector
  if (hcd && d.fast_metals) return false;            // Kaiser / effective bias is rational in F
  return true;
}
// terms an evaluation's loops run over (the tables of the others are zero)
__host__ __device__ inline int moment_count(const vg_eval_desc& d) {
  return moment_has_uv(d) ? MOM_MAX : (moment_has_hcd(d) ? 9 : 3);
}

// G_uv(k) at the parameters p: what uv_bias_beta adds to a unit bias (vega/power_spectrum.py:224-261)
__device__ __forceinline__ double moment_uv_term(const vg_eval_desc& d, const double* p, double k) {
  double b = 1.0, beta = 1.0;
  uv_bias_beta(d, p, k, b, beta);
  return b - 1.0;
}

// One grid point's contribution to the F^a mu^2j accumulators: v = S (x any extra static factor),
// f = F, na = powers of F carried (1 or 3)
__device__ __forceinline__ void moment_accumulate(double (&acc)[9][PK_MAX_ELL], const double (&w)[PK_MAX_ELL],
                                                  double v, double f, double mu2, int na) {
  double fa = v;
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    if (a < na) {
      double u = fa;
#pragma unroll
      for (int j = 0; j < 3; ++j) {
#pragma unroll
        for (int l = 0; l < PK_MAX_ELL; ++l) acc[a * 3 + j][l] = fma(w[l], u, acc[a * 3 + j][l]);
        u *= mu2;
      }
      fa *= f;
    }
  }
}
// out[m][l][nk_pad] at column ik; the five UV tables are F^a mu^2j ones times G(k)^p (G does not
// depend on mu)
__device__ __forceinline__ void moment_store(const double (&acc)[9][PK_MAX_ELL], double pk_lin, double gk, bool uv,
                                             int n_ell, int nk_pad, int ik, double* __restrict__ out) {
  for (int m = 0; m < MOM_MAX; ++m)
    for (int l = 0; l < n_ell; ++l) {
      double v;
      if (m < 9) v = acc[m][l];
      else if (!uv) v = 0.0;
      else if (m == 9) v = gk * acc[0][l];            // G
      else if (m == 10) v = gk * acc[3][l];           // G F
      else if (m == 11) v = gk * gk * acc[0][l];      // G^2
      else if (m == 12) v = gk * acc[1][l];           // G mu^2
      else v = gk * acc[4][l];                        // G F mu^2
      out[((long long)m * n_ell + l) * nk_pad + ik] = v * pk_lin;
    }
}

// Plan time: mom[j][m][l][nk_pad] = pk_lin(k) G(k)^p * sum_i wleg[l][i] S(mu_i,k) F(mu_i,k)^a mu_i^2j
// for the moment evaluations ids[j].  grid = (ceil(nk/128), n_ids), block 128.
__global__ void __launch_bounds__(128)
moment_grid_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                   const double* __restrict__ base, double* __restrict__ mom) {
  const int ik = blockIdx.x * 128 + threadIdx.x;
  if (ik >= g.nk) return;
  const vg_eval_desc& d = evals[ids[blockIdx.y]];
  const bool hcd = moment_has_hcd(d), uv = moment_has_uv(d);
  const int na = hcd ? 3 : 1;
  const double* S = d.static_grid >= 0 ? g.grids[d.static_grid] + ik : nullptr;
  const double* F = hcd ? g.grids[d.hcd_grid] + ik : nullptr;
  double acc[9][PK_MAX_ELL];
#pragma unroll
  for (int m = 0; m < 9; ++m)
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) acc[m][l] = 0.0;
  for (int i = 0; i < g.nmu; ++i) {
    const double mu = g.mu[i], mu2 = mu * mu;
    const double s = S ? S[(long long)i * g.nk] : 1.0;
    const double f = F ? F[(long long)i * g.nk] : 0.0;
    double w[PK_MAX_ELL];
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) w[l] = l < g.n_ell ? g.wleg[l * g.nmu + i] : 0.0;
    moment_accumulate(acc, w, s, f, mu2, na);
  }
  const double gk = uv ? moment_uv_term(d, base, g.k[ik]) : 0.0;
  moment_store(acc, g.pk_lin[d.pk_lin_sel][ik], gk, uv, g.n_ell, g.nk_pad, ik,
               mom + (long long)blockIdx.y * MOM_MAX * g.n_ell * g.nk_pad);
}

// Coefficients of the MOM_MAX terms in the product of the two tracers' Kaiser terms; g0, g1 = 1 for
// a tracer that carries the UV term (c holds the coefficients WITHOUT it: kaiser_setup(.., false))
__host__ __device__ __forceinline__ void moment_coefs(const KaiserCoef& c, double g0, double g1, double* q) {
  const double A0 = c.a0[0], H0 = c.h0[0], A1 = c.a1[0], H1 = c.h1[0];
  const double B0 = c.a0[1], G0 = c.h0[1], B1 = c.a1[1], G1 = c.h1[1];
  // mu^0: (A0 + g0 G + H0 F)(B0 + g1 G + G0 F)
  q[0 * 3 + 0] = A0 * B0;
  q[1 * 3 + 0] = A0 * G0 + H0 * B0;
  q[2 * 3 + 0] = H0 * G0;
  // mu^2: (A0 + g0 G + H0 F)(B1 + G1 F) + (A1 + H1 F)(B0 + g1 G + G0 F)
  q[0 * 3 + 1] = A0 * B1 + A1 * B0;
  q[1 * 3 + 1] = A0 * G1 + H0 * B1 + A1 * G0 + H1 * B0;
  q[2 * 3 + 1] = H0 * G1 + H1 * G0;
  // mu^4: (A1 + H1 F)(B1 + G1 F)
  q[0 * 3 + 2] = A1 * B1;
  q[1 * 3 + 2] = A1 * G1 + H1 * B1;
  q[2 * 3 + 2] = H1 * G1;
  // the UV terms
  q[9] = A0 * g1 + g0 * B0;       // G
  q[10] = g0 * G0 + H0 * g1;      // G F
  q[11] = g0 * g1;                // G^2
  q[12] = g0 * B1 + A1 * g1;      // G mu^2
  q[13] = g0 * G1 + H1 * g1;      // G F mu^2
}

// The coefficients of evaluation d at the parameters p (same_tracer: the second tracer is the first)
__host__ __device__ __forceinline__ void moment_coefs_of(const vg_eval_desc& d, const double* p, double* q) {
  KaiserCoef kc = kaiser_setup(d, p, 1.0, false);
  double g0 = 0.0, g1 = 0.0;
  if (d.uvb || d.heii) { g0 = d.uv_on[0] ? 1.0 : 0.0; g1 = d.uv_on[1] ? 1.0 : 0.0; }
  if (d.same_tracer) { kc.a0[1] = kc.a0[0]; kc.a1[1] = kc.a1[0]; kc.h0[1] = kc.h0[0]; kc.h1[1] = kc.h1[0]; g1 = g0; }
  moment_coefs(kc, g0, g1, q);
}

// Hot path: tables[b][j][x] = sum_m c_m(theta_b, eval ids[j]) * ktab[j][m][x],  x over the
// n_ell*2*n_out_pad knot values and second derivatives of the evaluation.
// n_slots >= n_ids tables per vector (the velocity-dispersion tables follow the n_ids of this kernel).
// grid = (ceil(row_len / (256*MOM_XPT*x_iters)), n_ids, ceil(B / MOM_NB)), block 256.  A block covers
// x_iters stretches of 256*MOM_XPT values: the coefficients of its MOM_NB vectors (a serial chain of
// parameter loads on MOM_NB threads while the others wait) are paid once per block, so large batches
// take long blocks (moment_x_iters) and a thin batch keeps many short ones.
constexpr int MOM_NB = 4, MOM_XPT = 2;
// stretches per block: all of them once the grid still holds `enough` blocks, fewer for thin batches
inline int moment_x_iters(long long stretches, long long blocks_per_stretch, long long enough = 1184) {
  long long it = stretches * blocks_per_stretch / enough;
  if (it < 1) it = 1;
  if (it > stretches) it = stretches;
  return (int)it;
}
__global__ void __launch_bounds__(256)
moment_tables_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids, int n_ids,
                     const double* __restrict__ pfull, int npf, int B, const double* __restrict__ ktab,
                     int row_len, int n_slots, double* __restrict__ tables, int x_iters) {
  __shared__ double cq[MOM_NB][MOM_MAX];
  __shared__ int s_nm;
  const int j = blockIdx.y, b0 = blockIdx.z * MOM_NB;
  if (threadIdx.x < MOM_NB) {
    const vg_eval_desc& d = evals[ids[j]];
    const double* p = pfull + (long long)min(b0 + (int)threadIdx.x, B - 1) * npf;
    double q[MOM_MAX];
    moment_coefs_of(d, p, q);     // k only enters through the UV term, which the tables hold
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[threadIdx.x][m] = q[m];
    if (threadIdx.x == 0) s_nm = moment_count(d);
  }
  __syncthreads();
  const int nm = s_nm;
  const double* kt = ktab + (long long)j * MOM_MAX * row_len;
  for (int it = 0; it < x_iters; ++it)
#pragma unroll
  for (int u = 0; u < MOM_XPT; ++u) {
    const int x = ((blockIdx.x * x_iters + it) * MOM_XPT + u) * 256 + threadIdx.x;
    if (x >= row_len) break;
    double acc[MOM_NB];
#pragma unroll
    for (int e = 0; e < MOM_NB; ++e) acc[e] = 0.0;
    for (int m = 0; m < nm; ++m) {
      const double v = __ldg(kt + (long long)m * row_len + x);
#pragma unroll
      for (int e = 0; e < MOM_NB; ++e) acc[e] = fma(cq[e][m], v, acc[e]);
    }
#pragma unroll
    for (int e = 0; e < MOM_NB; ++e)
      if (b0 + e < B) tables[((long long)(b0 + e) * n_slots + j) * row_len + x] = acc[e];
  }
}

// weight of an evaluation's xi in its output slot (vega/metals.py:316-334); weight_const is part of
// the static tables
__device__ __forceinline__ double moment_weight(const vg_eval_desc& d, const double* p) {
  double w = 1.0;
  if (d.weight_bias) {
    double b1, be1, b2, be2;
    tracer_bias_beta(d, p, 0, b1, be1);
    tracer_bias_beta(d, p, 1, b2, be2);
    w = b1 * b2;
  }
  if (d.idx_weight_extra >= 0) w *= p[d.idx_weight_extra];
  if (d.idx_weight_extra2 >= 0) w *= p[d.idx_weight_extra2];
  return w;
}

// Moment form of the xi stage for evaluations whose coordinates and per-bin factors are the
// same for every vector (vg_eval_desc::static_xi): the rescale / spline read-out / Legendre
// sum / evolution factors (vega/correlation_func.py:117-236, vega/pktoxi.py:141-163) are
// linear in the knot tables, so they are applied ONCE, in vg_plan_finalize, to the static
// moment tables:  X[j][m][bin] = xi stage of evaluation ids[j] on K[owner(j)][m].  Hot path:
//   out[b][slot(j)][bin] = w_j(theta_b) * sum_m c_m(theta_b; owner(j)) * X[j][m][bin]
// -- 3, 9 or 14 FMAs per bin instead of a knot-table row and a gather.  `owners[j]` is the
// evaluation whose P(k,mu) the table belongs to (j itself unless it shares a table).
// grid = (ceil(ld / (256*x_iters)), n_ids, ceil(B / MXI_NB)), block 256 (x_iters: as in moment_tables_kernel).
constexpr int MXI_NB = 8;
__global__ void __launch_bounds__(256)
moment_xi_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                 const int* __restrict__ owners, const double* __restrict__ pfull, int npf, int B,
                 const double* __restrict__ X, const int* __restrict__ rows, int ld,
                 double* __restrict__ core, double* __restrict__ xmet,
                 const int* __restrict__ metal_map, int n_metal, int x_iters) {
  __shared__ double cq[MXI_NB][MOM_MAX];
  __shared__ int s_nm, s_slot;
  const int j = blockIdx.y, b0 = blockIdx.z * MXI_NB;
  if (threadIdx.x < MXI_NB) {
    const vg_eval_desc& d = evals[ids[j]];
    const vg_eval_desc& o = evals[owners[j]];
    const double* p = pfull + (long long)min(b0 + (int)threadIdx.x, B - 1) * npf;
    double q[MOM_MAX];
    moment_coefs_of(o, p, q);
    const double w = moment_weight(d, p);
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[threadIdx.x][m] = w * q[m];
    if (threadIdx.x == 0) {
      s_nm = moment_count(o);
      s_slot = d.out_slot;
    }
  }
  __syncthreads();
  const int nm = s_nm;
  const int slot = s_slot;
  for (int it = 0; it < x_iters; ++it) {
    const int x = (blockIdx.x * x_iters + it) * 256 + threadIdx.x;
    if (x >= ld) return;
    const double* xt = X + (long long)(rows ? rows[j] : j) * MOM_MAX * ld + x;
    double acc[MXI_NB];
#pragma unroll
    for (int e = 0; e < MXI_NB; ++e) acc[e] = 0.0;
    for (int m = 0; m < nm; ++m) {
      const double v = __ldg(xt + (long long)m * ld);
#pragma unroll
      for (int e = 0; e < MXI_NB; ++e) acc[e] = fma(cq[e][m], v, acc[e]);
    }
#pragma unroll
    for (int e = 0; e < MXI_NB; ++e) {
      const int b = b0 + e;
      if (b >= B) break;
      double* out = slot < 2 ? core + ((long long)slot * B + b) * ld
                             : xmet + ((long long)b * n_metal + metal_map[slot - 2]) * ld;
      out[x] = acc[e];
    }
  }
}

// Folded metal terms (VG_OPT_FOLD_METALS): a metal evaluation whose xi stage is in moment
// form is a 3/9/14-term combination of static vectors X[j][m], so its matrix product is the same
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
  int* s_nm = reinterpret_cast<int*>(fold_smem + (size_t)n_ids * MOM_MAX * MXI_NB);
  const int b0 = blockIdx.y * MXI_NB;
  for (int t = threadIdx.x; t < n_ids * MXI_NB; t += 256) {
    const int j = t / MXI_NB, e = t % MXI_NB;
    const vg_eval_desc& d = evals[ids[j]];
    const vg_eval_desc& o = evals[owners[j]];
    const double* p = pfull + (long long)min(b0 + e, B - 1) * npf;
    double q[MOM_MAX];
    moment_coefs_of(o, p, q);
    const double w = moment_weight(d, p);
#pragma unroll
    for (int m = 0; m < MOM_MAX; ++m) cq[((size_t)j * MOM_MAX + m) * MXI_NB + e] = w * q[m];
    if (e == 0) s_nm[j] = moment_count(o);
  }
  __syncthreads();
  const int x = blockIdx.x * 256 + threadIdx.x;
  if (x >= ld) return;
  double acc[MXI_NB];
#pragma unroll
  for (int e = 0; e < MXI_NB; ++e) acc[e] = 0.0;
  for (int j = 0; j < n_ids; ++j) {
    const double* yt = Y + (long long)j * MOM_MAX * ld + x;
    const int nm = s_nm[j];
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

// ---------------------------------------------------------------------------------------------
// Velocity-dispersion tables: the moment form extended to evaluations whose ONLY sampled
// non-polynomial P(k,mu) factor is one velocity dispersion sigma_v (vega/power_spectrum.py:588-636:
// a Lorentzian 1/sqrt(1 + (k mu sigma)^2) or a Gaussian exp(-(k mu sigma)^2 / 4) per discrete
// tracer).  For fixed sigma the factor is one more static grid, so the knot tables are
//   knots_l[b] = sum_m c_m(theta_b) * K_l[m](sigma_b),
// and K(sigma) -- smooth (analytic) in sigma -- is represented over the parameter's [sample] limits
// [lo, hi] by its Chebyshev series in  tau = log(sigma^2 + eps^2) / 2  (in tau the Lorentzian's
// singularities sigma = +- i / (k mu) stay a distance >= pi/4 from the real axis for every k mu
// >= 1/eps; eps only flattens the map near sigma = 0):
//   K(sigma) = sum_{n < N} T_n(t(sigma)) C_n,      t = (2 tau - tau_lo - tau_hi) / (tau_hi - tau_lo).
// C_n come from VT_NODES Chebyshev-Gauss nodes (per-point pass with the velocity factor + the
// Hankel/spline operator, at finalize); N is the smallest order count whose tables reproduce the
// per-point pass on VT_NTEST test values of sigma to VT_TOL * max|xi| (checked at finalize on the
// xi stage's output; no N <= VT_NODES passing = the evaluations keep the per-point pass).
// Hot path: W[b][n*nm + i] = T_n(t_b) c_{term i}(theta_b) over the nm terms whose coefficient is not
// structurally zero (6 for HCD on one tracer, 8 with its UV term), then ONE GEMM per evaluation
//   tables[b][x] = sum_k W[b][k] KT[x][k]        (K = N * nm, the DMMA kernel of vg_gemm.cuh)
// instead of the 814 000-point pass and the Hankel GEMM rows of that evaluation.
// ---------------------------------------------------------------------------------------------
constexpr int VT_NODES = 64;          // Chebyshev-Gauss nodes the series is built from
constexpr int VT_NTEST = 8;           // test values of sigma at finalize
constexpr double VT_TOL = 2e-12;      // accepted |xi(tables) - xi(per point)| / max|xi|, per output slot
constexpr double VT_EPS = 0.05;       // eps of the tau map [Mpc/h]

// the parameter column of the single sampled velocity dispersion, or -1
inline int vt_sigma_column(const vg_eval_desc& d) {
  if (d.dyn_mask != VG_F_VELO) return -1;
  const int i0 = d.idx_velo[0], i1 = d.idx_velo[1];
  if (i0 >= 0 && i1 >= 0 && i0 != i1) return -1;     // two dispersions: a 2-D family
  return i0 >= 0 ? i0 : i1;
}
// everything but the velocity factor must qualify for the moment form
inline bool vt_eligible(const vg_eval_desc& d) {
  if (vt_sigma_column(d) < 0) return false;
  vg_eval_desc e = d;
  e.dyn_mask = 0;
  return moment_eligible(e);
}
// The terms of an evaluation whose coefficient is not structurally zero (unit amplitudes where a
// tracer carries the term, zero where it does not): idx[0..n) in increasing order.
struct VtTerms { int n; int idx[MOM_MAX]; };
inline VtTerms vt_terms(const vg_eval_desc& d) {
  KaiserCoef kc;
  for (int t = 0; t < 2; ++t) {
    const int s = d.same_tracer ? 0 : t;
    const double h = (d.hcd_model && d.hcd_on[s]) ? 1.0 : 0.0;
    kc.a0[t] = kc.a1[t] = 1.0; kc.h0[t] = kc.h1[t] = h;
  }
  kc.L0 = 1.0;
  const bool uv = d.uvb || d.heii;
  const double g0 = uv && d.uv_on[0] ? 1.0 : 0.0;
  const double g1 = uv && d.uv_on[d.same_tracer ? 0 : 1] ? 1.0 : 0.0;
  double q[MOM_MAX];
  moment_coefs(kc, g0, g1, q);
  VtTerms out{};
  for (int m = 0; m < MOM_MAX; ++m)
    if (q[m] != 0.0) out.idx[out.n++] = m;
  return out;
}

struct VtMap { double lo, hi, tau_lo, tau_hi, eps; };
__host__ __device__ inline double vt_tau(double sigma, double eps) { return 0.5 * log(sigma * sigma + eps * eps); }
__host__ __device__ inline double vt_sigma_of(double t, const VtMap& m) {
  const double tau = 0.5 * (m.tau_lo + m.tau_hi) + 0.5 * (m.tau_hi - m.tau_lo) * t;
  const double s2 = exp(2.0 * tau) - m.eps * m.eps;
  return s2 > 0.0 ? sqrt(s2) : 0.0;
}

// Plan time: mom[(j*nq + q)][m][l][nk_pad] = moment_grid_kernel's tables of evaluation ids[j] with the
// velocity factor V(k mu_i; sigma[q]) inside the mu sum.  grid = (ceil(nk/128), nq, n_ids), block 128.
__global__ void __launch_bounds__(128)
vtab_grid_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                 const double* __restrict__ base, const double* __restrict__ sigma, int nq,
                 double* __restrict__ mom) {
  const int ik = blockIdx.x * 128 + threadIdx.x;
  if (ik >= g.nk) return;
  const vg_eval_desc& d = evals[ids[blockIdx.z]];
  const bool hcd = moment_has_hcd(d), uv = moment_has_uv(d);
  const int na = hcd ? 3 : 1;
  const double* S = d.static_grid >= 0 ? g.grids[d.static_grid] + ik : nullptr;
  const double* F = hcd ? g.grids[d.hcd_grid] + ik : nullptr;
  const double k = g.k[ik];
  const double sg = sigma[blockIdx.y];
  const int n_velo = (d.idx_velo[0] >= 0) + (d.idx_velo[1] >= 0);
  const int gauss = d.velo_kind == 1;
  double acc[9][PK_MAX_ELL];
#pragma unroll
  for (int m = 0; m < 9; ++m)
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) acc[m][l] = 0.0;
  for (int i = 0; i < g.nmu; ++i) {
    const double mu = g.mu[i], mu2 = mu * mu;
    const double x = k * mu * sg;
    // pk_factors(VG_F_VELO): one factor per discrete tracer
    const double v1 = gauss ? exp(-0.25 * x * x) : 1.0 / sqrt(1.0 + x * x);
    const double v = n_velo == 2 ? v1 * v1 : v1;
    const double s = (S ? S[(long long)i * g.nk] : 1.0) * v;
    const double f = F ? F[(long long)i * g.nk] : 0.0;
    double w[PK_MAX_ELL];
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) w[l] = l < g.n_ell ? g.wleg[l * g.nmu + i] : 0.0;
    moment_accumulate(acc, w, s, f, mu2, na);
  }
  const double gk = uv ? moment_uv_term(d, base, k) : 0.0;
  const long long row = (long long)blockIdx.z * nq + blockIdx.y;
  moment_store(acc, g.pk_lin[d.pk_lin_sel][ik], gk, uv, g.n_ell, g.nk_pad, ik,
               mom + row * MOM_MAX * g.n_ell * g.nk_pad);
}

// Plan time: node values -> Chebyshev coefficients of the listed terms, K-major for the GEMM:
//   KT[x][n*nm + i] = sum_q D[n][q] node[q][terms[i]][x],   D[n][q] = (2 - [n = 0]) / nq * cos(pi n (q + 1/2) / nq)
// node: [nq][MOM_MAX][row_len] of ONE evaluation.  grid = (ceil(row_len/128), nq), block 128.
__global__ void __launch_bounds__(128)
vtab_cheb_kernel(const double* __restrict__ node, int nq, VtTerms terms, int row_len, double* __restrict__ KT) {
  const int x = blockIdx.x * 128 + threadIdx.x, n = blockIdx.y;
  if (x >= row_len) return;
  const int nm = terms.n;
  for (int i = 0; i < nm; ++i) {
    const int m = terms.idx[i];
    double acc = 0.0;
    for (int q = 0; q < nq; ++q)
      acc = fma(cospi((double)n * ((double)q + 0.5) / (double)nq), node[((long long)q * MOM_MAX + m) * row_len + x], acc);
    KT[(long long)x * ((long long)nq * nm) + (long long)n * nm + i] = acc * (n == 0 ? 1.0 : 2.0) / (double)nq;
  }
}

// Hot path: W[j][b][n*nm_j + i] = T_n(t(sigma_b)) * c_{terms_j[i]}(theta_b; evaluation ids[j]), zero
// beyond n_keep*nm_j up to ldw.  terms: [n_ids] term lists.  grid = (B, n_ids), block 128: one block per
// row of W -- thread 0 evaluates the coefficients and t, every thread runs the Chebyshev recurrence up
// to its own order (at most 64 steps) and the row is written coalesced.  (One thread per row wrote its
// ~300 doubles one after the other, ldw doubles apart from its neighbour's: 47 us whatever the batch,
// on the critical path of a single vector.)
__global__ void __launch_bounds__(128)
vtab_weights_kernel(const vg_eval_desc* __restrict__ evals, const int* __restrict__ ids,
                    const VtTerms* __restrict__ terms, const double* __restrict__ pfull, int npf, int B,
                    VtMap map, int n_keep, int ldw, double* __restrict__ W) {
  const int b = blockIdx.x, j = blockIdx.y;
  if (b >= B) return;
  const VtTerms& tm = terms[j];
  const int nm = tm.n;
  __shared__ double qs[MOM_MAX];
  __shared__ double ts;
  if (threadIdx.x == 0) {
    const vg_eval_desc& d = evals[ids[j]];
    const double* p = pfull + (long long)b * npf;
    double q[MOM_MAX];
    moment_coefs_of(d, p, q);
#pragma unroll
    for (int i = 0; i < MOM_MAX; ++i) qs[i] = q[i];
    const int col = d.idx_velo[0] >= 0 ? d.idx_velo[0] : d.idx_velo[1];
    const double sg = fmin(fmax(fabs(p[col]), map.lo), map.hi);
    const double t = (2.0 * vt_tau(sg, map.eps) - map.tau_lo - map.tau_hi) / (map.tau_hi - map.tau_lo);
    ts = fmin(fmax(t, -1.0), 1.0);
  }
  __syncthreads();
  const double t = ts;
  double* w = W + ((long long)j * B + b) * ldw;
  const int n_used = n_keep * nm;
  for (int k = threadIdx.x; k < ldw; k += 128) {
    double v = 0.0;
    if (k < n_used) {
      const int n = k / nm, i = k - n * nm;
      double t0 = 1.0, t1 = t;
      for (int m = 0; m < n; ++m) {
        const double t2 = 2.0 * t * t1 - t0;
        t0 = t1; t1 = t2;
      }
      v = t0 * qs[tm.idx[i]];
    }
    w[k] = v;
  }
}

// flag[0] |= 1 when a vector's sigma (column col of theta) lies outside [lo, hi] or is not finite
__global__ void vtab_range_kernel(const double* __restrict__ theta, int n_theta, int tcol, long long B, double lo,
                                  double hi, int* __restrict__ flag) {
  const long long b = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const double v = theta[b * n_theta + tcol];
  if (!(v >= lo && v <= hi)) atomicOr(flag, 1);
}

// out[0] = max |a - b| (as bits), out[1] = max |b| over rows blockIdx.y (row stride `stride`) and
// the first n columns (plan-time check of the tables against the per-point pass)
__global__ void vtab_compare_kernel(const double* __restrict__ a, const double* __restrict__ b, int n,
                                    long long stride, unsigned long long* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const long long o = (long long)blockIdx.y * stride + i;
  const double d = fabs(a[o] - b[o]), s = fabs(b[o]);
  if (!(d == d) || !(s == s)) { atomicMax(out, (unsigned long long)__double_as_longlong(1e300)); return; }
  atomicMax(out, (unsigned long long)__double_as_longlong(d));
  atomicMax(out + 1, (unsigned long long)__double_as_longlong(s));
}

}  // namespace vg
