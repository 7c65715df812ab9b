// K1 + K2a: anisotropic power spectrum P(k,mu) on the (mu,k) grid and its projection onto
// Legendre multipoles, for a batch of parameter vectors.
//
// Restates vega/power_spectrum.py:87-636 (PowerSpectrum.compute and its factors), the
// helpers vega/utils.py:28-108,396-441 and the projection of vega/pktoxi.py:79-97,138
// (midpoint rule on nmu uniform mu bins, weights dmu*(2l+1)*L_l(mu)).
//
// Layout: every grid is [nmu][nk] with k fastest, so a warp reads 32 consecutive k of one
// mu row (256 B coalesced).  Factors whose parameters are fixed for the whole batch are
// folded once into a static grid S(mu,k) (the reference caches the same factors between
// calls: power_spectrum.py:139-141,311-324,407-417,459-479); factors with sampled
// parameters are evaluated per vector.  The fused kernel never materialises P(k,mu): a
// thread owns one k and a strided slice of mu for NB parameter vectors at a time, so each
// static-grid value is read once per NB vectors and P_l(k) leaves the SM as 4 numbers.
#pragma once
#include "../../include/vega_b200.h"
#include "vg_common.cuh"

namespace vg {

constexpr double DEFAULT_GROWTH_RATE = 0.970386;   // vega/utils.py:63
constexpr int PK_MAX_ELL = 4;

struct PkGridDev {
  int nk, nk_pad, nmu, n_ell;
  int std_ells;                // multipoles are exactly (0, 2, 4, 6): Legendre weights in registers
  const double* k;
  const double* pk_lin[4];     // 0 full-smooth, 1 smooth, 2 full, 3 direct P(k)
  const double* pk_fid;
  const double* mu;
  const double* wleg;          // [n_ell][nmu]
  const double* fv_x; const double* fv_f; int fv_n;
  const double* const* grids;  // static grids, each [nmu][nk]
};

__host__ __device__ __forceinline__ double par(const double* p, int idx, double dflt) {
  return idx >= 0 ? p[idx] : dflt;
}

// (bias, beta) of tracer t from any two of (bias, bias_eta, beta): vega/utils.py:45-81
__host__ __device__ __forceinline__ void tracer_bias_beta(const vg_eval_desc& d, const double* p, int t,
                                                 double& bias, double& beta) {
  const double f = par(p, d.idx_growth_rate, DEFAULT_GROWTH_RATE);
  const double eta = par(p, d.idx_bias_eta[t], 0.0);
  bias = par(p, d.idx_bias[t], 0.0);
  beta = par(p, d.idx_beta[t], 0.0);
  if (d.idx_bias[t] < 0) bias = eta * f / beta;
  if (d.idx_beta[t] < 0) beta = eta * f / bias;
}

// vega/power_spectrum.py:224-261: scale-dependent bias from UV / HeII fluctuations
__host__ __device__ __forceinline__ void uv_bias_beta(const vg_eval_desc& d, const double* p, double k,
                                             double& bias, double& beta) {
  double eff = bias;
  if (d.uvb) {
    double x = k * p[d.idx_lambda_uv];
    double w = atan(x) / x;
    eff += p[d.idx_bias_gamma] * w / (1.0 + p[d.idx_bias_prim] * w);
  }
  if (d.heii) {
    double x = k * p[d.idx_lambda_heii];
    double w = atan(x) / x;
    eff += p[d.idx_bias_gamma_e] * w / (1.0 + p[d.idx_bias_prim] * w);
  }
  beta = beta * bias / eff;
  bias = eff;
}

// np.interp(x, xp, fp, left=1, right=0) on the Fvoigt table (power_spectrum.py:376-378)
__device__ __forceinline__ double fvoigt_interp(const PkGridDev& g, double x) {
  if (x < g.fv_x[0]) return 1.0;
  if (x > g.fv_x[g.fv_n - 1]) return 0.0;
  int lo = 0, hi = g.fv_n - 1;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (g.fv_x[mid] <= x) lo = mid; else hi = mid;
  }
  double x0 = g.fv_x[lo], x1 = g.fv_x[hi];
  double slope = (g.fv_f[hi] - g.fv_f[lo]) / (x1 - x0);
  return slope * (x - x0) + g.fv_f[lo];
}

// F_hcd(k_par): Rogers2018 exp(-L0 k_par), Fvoigt table, sinc (power_spectrum.py:326-380)
__device__ __forceinline__ double hcd_profile(const vg_eval_desc& d, const PkGridDev& g,
                                              double L0, double kpar) {
  if (d.hcd_model == 1) return exp(-L0 * kpar);
  if (d.hcd_model == 2) return fvoigt_interp(g, L0 * kpar);
  double x = kpar * L0;
  return sin(x) / x;
}

// Product of the factors selected by `mask` at one grid point.
__device__ double pk_factors(uint32_t mask, const vg_eval_desc& d, const double* p,
                             double k, double mu, double pk_fid_k, int* bad) {
  const double kpar = k * mu;
  const double kperp = k * sqrt(1.0 - mu * mu);
  double f = 1.0;
  if (mask & VG_F_ARINYO) {          // power_spectrum.py:435-479
    const double q1 = p[d.idx_arinyo[0]], q2 = par(p, d.idx_arinyo[1], 0.0);
    const double kv = p[d.idx_arinyo[2]], av = p[d.idx_arinyo[3]];
    const double bv = p[d.idx_arinyo[4]], kp = p[d.idx_arinyo[5]];
    const double d2 = k * k * k * pk_fid_k / (2.0 * M_PI * M_PI);
    const double growth = q1 * d2 + q2 * d2 * d2;
    const double vel = pow(k / kv, av) * pow(fabs(mu), bv);
    double dnl = exp(growth * (1.0 - vel) - (k / kp) * (k / kp));
    if (!isfinite(dnl)) *bad |= VG_STATUS_ARINYO;
    if (d.arinyo_mode == 1) dnl = sqrt(dnl);
    else if (d.arinyo_mode == 0) dnl = 1.0;
    f *= dnl;
  }
  if (mask & VG_F_MCDONALD) {        // power_spectrum.py:419-433
    const double kvel = 1.22 * pow(1.0 + k / 0.923, 0.451);
    f *= exp(pow(k / 6.4, 0.569) - pow(k / 15.3, 2.01) - pow(k * mu / kvel, 1.5));
  }
  if (mask & VG_F_GK) {              // power_spectrum.py:481-502
    if (d.gk_drp != 0.0) { double x = kpar * d.gk_drp / 2.0; f *= sin(x) / x; }
    if (d.gk_drt != 0.0) { double x = kperp * d.gk_drt / 2.0; f *= sin(x) / x; }
  }
  if (mask & VG_F_MOCKGK) {          // power_spectrum.py:143-160
    double drp = d.mock_bin_size, drt = d.mock_bin_size;
    if (d.mock_los_mode == 1) drp *= 1.0 + p[d.idx_growth_rate];
    else if (d.mock_los_mode == 2) drp *= 1.0 + p[d.idx_los_smooth_amp];
    else if (d.mock_los_mode == 3) drt = 0.0;
    if (drp != 0.0) { double x = kpar * drp / 2.0; f *= sin(x) / x; }
    if (drt != 0.0) { double x = kperp * drt / 2.0; f *= sin(x) / x; }
  }
  if (mask & VG_F_PEAKNL) {          // power_spectrum.py:382-417
    const double gr = par(p, d.idx_growth_rate, 0.0);
    double sp = par(p, d.idx_sigmaNL[0], 0.0), st = par(p, d.idx_sigmaNL[1], 0.0);
    if (d.idx_sigmaNL[0] < 0) sp = st * (1.0 + gr);
    if (d.idx_sigmaNL[1] < 0) st = sp / (1.0 + gr);
    f *= exp(-(kpar * kpar * sp * sp + kperp * kperp * st * st) / 2.0);
  }
  if (mask & VG_F_FSGAUSS) {         // power_spectrum.py:504-553, vega/utils.py:396-420
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const double sp = p[d.idx_fs_gauss[2 * i]], st = p[d.idx_fs_gauss[2 * i + 1]];
      f *= exp(-(kpar * kpar * sp * sp + kperp * kperp * st * st) / 2.0);
    }
  }
  if (mask & VG_F_FSEXP) {           // power_spectrum.py:560-586
    const double sp = p[d.idx_fs_exp[0]], st = p[d.idx_fs_exp[1]];
    const double ep = p[d.idx_fs_exp[2]], et = p[d.idx_fs_exp[3]];
    const double gs = kpar * kpar * sp * sp + kperp * kperp * st * st;
    const double es = fabs(kpar) * ep * ep + fabs(kperp) * et * et;
    f *= exp(-gs / 2.0) * exp(-es);
  }
  if (mask & VG_F_VELO) {            // power_spectrum.py:588-636
#pragma unroll
    for (int t = 0; t < 2; ++t) {
      if (d.idx_velo[t] < 0) continue;
      const double x = kpar * p[d.idx_velo[t]];
      if (d.velo_kind == 1) f *= exp(-0.25 * x * x);
      else f *= 1.0 / sqrt(1.0 + x * x);
    }
  }
  if (mask & VG_F_KDAMP) {           // vega/utils.py:423-441
    f *= exp(-d.kdamp_scale * d.kdamp_scale * pow(k, (double)d.kdamp_power) / 2.0);
  }
  return f;
}

// Kaiser term of tracer t as (a0 + h0 F) + (a1 + h1 F) mu^2, F = F_hcd(k_par)
// (power_spectrum.py:198-222 with the effective bias/beta of :263-309 substituted).
struct KaiserCoef {
  double a0[2], a1[2], h0[2], h1[2];
  double L0;
};

// with_uv = false leaves the scale-dependent UV term out (the moment form carries it as a static
// function of k: vg_moments.cuh)
__host__ __device__ __forceinline__ KaiserCoef kaiser_setup(const vg_eval_desc& d, const double* p, double k,
                                                            bool with_uv = true) {
  KaiserCoef c;
#pragma unroll
  for (int t = 0; t < 2; ++t) {
    double b, beta;
    tracer_bias_beta(d, p, t, b, beta);
    if (with_uv && d.uv_on[t] && (d.uvb || d.heii)) uv_bias_beta(d, p, k, b, beta);
    c.a0[t] = b;
    c.a1[t] = b * beta;
    c.h0[t] = 0.0;
    c.h1[t] = 0.0;
    if (d.hcd_model && d.hcd_on[t]) {
      const double bh = p[d.idx_bias_hcd], betah = p[d.idx_beta_hcd];
      c.h0[t] = bh;
      c.h1[t] = bh * betah;
    } else if (d.fast_metals) {   // (1 + beta mu^2) without the bias
      c.a0[t] = 1.0;
      c.a1[t] = beta;
    }
  }
  c.L0 = par(p, d.idx_L0, 1.0);
  return c;
}

__device__ __forceinline__ double kaiser_eval(const vg_eval_desc& d, const KaiserCoef& c,
                                              double F, double mu2) {
  double t0b = fma(c.h0[0], F, c.a0[0]);
  double t0 = fma(fma(c.h1[0], F, c.a1[0]), mu2, t0b);
  double t1b = fma(c.h0[1], F, c.a0[1]);
  double t1 = fma(fma(c.h1[1], F, c.a1[1]), mu2, t1b);
  if (d.fast_metals && d.hcd_model) {
    if (d.hcd_on[0]) t0 /= t0b;
    if (d.hcd_on[1]) t1 /= t1b;
  }
  return t0 * t1;
}

constexpr int PK_KLANES = 32, PK_GROUPS = 8;   // CTA = 32 k lanes x (MG mu groups x EG vector groups)
constexpr int PK_UNROLL = 4;                   // mu rows whose grid loads are issued together

// Legendre weights are applied as L_l(mu) inside the loop and dmu*(2l+1) once at the end
__device__ __forceinline__ void legendre_0246(double mu2, double* L) {
  L[0] = 1.0;
  L[1] = fma(1.5, mu2, -0.5);
  L[2] = fma(fma(4.375, mu2, -3.75), mu2, 0.375);
  L[3] = fma(fma(fma(14.4375, mu2, -19.6875), mu2, 6.5625), mu2, -0.3125);
}

enum { PK_MODE_STATIC = 0, PK_MODE_VELO = 1, PK_MODE_GENERIC = 2 };

// Fused K1+K2a.  grid = (ceil(B / (NB*EG)), ceil(nk/32), n_ids), block = (32, MG, EG),
// MG*EG = 8.  A thread owns one k, the mu rows i = mg, mg+MG, ... and NB parameter vectors;
// the EG vector groups of a CTA walk the same (mu, k) addresses, so the static grids are
// fetched from L2 once per NB*EG vectors.  MODE selects how the per-vector factors are
// evaluated: none, velocity dispersion only (parameters hoisted into registers), or the
// generic path that reads parameters per point.
// out[b][out_slot[eval]][l][nk_pad], n_slots slots per vector (pad columns are never written
// and stay zero)
template <int NB, int MODE>
__global__ void __launch_bounds__(PK_KLANES * PK_GROUPS, 2)
pk_ell_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, int n_eval,
              const int* __restrict__ eval_ids, const int* __restrict__ out_slot, int n_slots,
              const double* __restrict__ pfull, int npf, int B,
              double* __restrict__ out, int* __restrict__ status) {
  __shared__ vg_eval_desc d;
  __shared__ double red[PK_GROUPS][NB * PK_MAX_ELL][PK_KLANES + 1];
  const int tx = threadIdx.x, mg = threadIdx.y, eg = threadIdx.z;
  const int MG = blockDim.y, EG = blockDim.z;
  const int tid = (eg * MG + mg) * PK_KLANES + tx;
  const int eval = eval_ids[blockIdx.z];
  const int slot = out_slot[eval];
  {
    const int* src = reinterpret_cast<const int*>(evals + eval);
    int* dst = reinterpret_cast<int*>(&d);
    for (int i = tid; i < (int)(sizeof(vg_eval_desc) / 4); i += PK_KLANES * PK_GROUPS) dst[i] = src[i];
  }
  __syncthreads();
  const int ik_raw = blockIdx.y * PK_KLANES + tx;
  const int ik = min(ik_raw, g.nk - 1);
  const double k = g.k[ik];
  const double pk_fid_k = g.pk_fid[ik];
  const int b0 = (blockIdx.x * EG + eg) * NB;

  // per-vector Kaiser coefficients: term_t = (a0 + h0 F) + (a1 + h1 F) mu^2
  double a0[NB][2], a1[NB][2], h0[NB][2], h1[NB][2], L0[NB], sv[NB][2];
  const double* pp[NB];
#pragma unroll
  for (int e = 0; e < NB; ++e) {
    pp[e] = pfull + (long long)min(b0 + e, B - 1) * npf;
    const KaiserCoef kc = kaiser_setup(d, pp[e], k);
#pragma unroll
    for (int t = 0; t < 2; ++t) { a0[e][t] = kc.a0[t]; a1[e][t] = kc.a1[t]; h0[e][t] = kc.h0[t]; h1[e][t] = kc.h1[t]; }
    L0[e] = kc.L0;
    if (MODE == PK_MODE_VELO) {
#pragma unroll
      for (int t = 0; t < 2; ++t) sv[e][t] = d.idx_velo[t] >= 0 ? pp[e][d.idx_velo[t]] : 0.0;
    }
  }
  const double* S = d.static_grid >= 0 ? g.grids[d.static_grid] + ik : nullptr;
  const double* Fg = (d.hcd_model && d.hcd_grid >= 0) ? g.grids[d.hcd_grid] + ik : nullptr;
  const bool hcd_any = d.hcd_model != 0;
  const bool hcd_dyn = d.hcd_model && d.hcd_grid < 0;
  const bool same = d.same_tracer != 0;
  const bool divide0 = d.fast_metals && d.hcd_model && d.hcd_on[0];
  const bool divide1 = d.fast_metals && d.hcd_model && d.hcd_on[1];
  const bool std_ells = g.std_ells != 0;
  const int velo_kind = d.velo_kind;
  const uint32_t dyn_mask = d.dyn_mask;
  const double inv_nmu = 1.0 / (double)g.nmu;

  double acc[NB][PK_MAX_ELL];
#pragma unroll
  for (int e = 0; e < NB; ++e)
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) acc[e][l] = 0.0;
  int bad[NB];
#pragma unroll
  for (int e = 0; e < NB; ++e) bad[e] = 0;

  const int step = MG * PK_UNROLL;
  for (int i0 = mg; i0 < g.nmu; i0 += step) {
    double s[PK_UNROLL], fs[PK_UNROLL];
#pragma unroll
    for (int u = 0; u < PK_UNROLL; ++u) {
      const int i = i0 + u * MG;
      const bool in = i < g.nmu;
      s[u] = (S && in) ? __ldg(S + (long long)i * g.nk) : 1.0;
      fs[u] = (Fg && in) ? __ldg(Fg + (long long)i * g.nk) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < PK_UNROLL; ++u) {
      const int i = i0 + u * MG;
      if (i >= g.nmu) break;
      const double mu = ((double)i + 0.5) / (double)g.nmu;   // == g.mu[i] (power_spectrum.py:77)
      const double mu2 = mu * mu;
      double w[PK_MAX_ELL];
      if (std_ells) {
        legendre_0246(mu2, w);
      } else {
#pragma unroll
        for (int l = 0; l < PK_MAX_ELL; ++l) w[l] = l < g.n_ell ? g.wleg[l * g.nmu + i] : 0.0;
      }
      const double kpar = k * mu;
#pragma unroll
      for (int e = 0; e < NB; ++e) {
        double F = fs[u];
        if (hcd_dyn) F = hcd_profile(d, g, L0[e], kpar);
        double t0, t1;
        if (hcd_any) {
          const double t0b = fma(h0[e][0], F, a0[e][0]);
          t0 = fma(fma(h1[e][0], F, a1[e][0]), mu2, t0b);
          if (divide0) t0 /= t0b;
          if (same) {
            t1 = t0;
          } else {
            const double t1b = fma(h0[e][1], F, a0[e][1]);
            t1 = fma(fma(h1[e][1], F, a1[e][1]), mu2, t1b);
            if (divide1) t1 /= t1b;
          }
        } else {
          t0 = fma(a1[e][0], mu2, a0[e][0]);
          t1 = same ? t0 : fma(a1[e][1], mu2, a0[e][1]);
        }
        double P = t0 * t1 * s[u];
        if (MODE == PK_MODE_VELO) {
#pragma unroll
          for (int t = 0; t < 2; ++t) {
            if (d.idx_velo[t] < 0) continue;
            const double x = kpar * sv[e][t];
            if (velo_kind == 1) P *= exp(-0.25 * x * x);
            else P *= rsqrt(fma(x, x, 1.0));
          }
        } else if (MODE == PK_MODE_GENERIC) {
          P *= pk_factors(dyn_mask, d, pp[e], k, mu, pk_fid_k, &bad[e]);
        }
#pragma unroll
        for (int l = 0; l < PK_MAX_ELL; ++l) acc[e][l] = fma(w[l], P, acc[e][l]);
      }
    }
  }
  (void)inv_nmu;
  // deterministic reduction over the mu groups, one vector group at a time
  for (int turn = 0; turn < EG; ++turn) {
    __syncthreads();
    if (eg == turn) {
#pragma unroll
      for (int e = 0; e < NB; ++e)
#pragma unroll
        for (int l = 0; l < PK_MAX_ELL; ++l) red[mg][e * PK_MAX_ELL + l][tx] = acc[e][l];
    }
    __syncthreads();
    if (eg == turn) {
      const double pk_lin = g.pk_lin[d.pk_lin_sel][ik];
      for (int o = mg; o < NB * PK_MAX_ELL; o += MG) {
        const int e = o / PK_MAX_ELL, l = o % PK_MAX_ELL;
        if (b0 + e >= B || l >= g.n_ell || ik_raw >= g.nk) continue;
        double sum = 0.0;
        for (int m = 0; m < MG; ++m) sum += red[m][o][tx];
        if (std_ells) sum *= (double)(4 * l + 1) / (double)g.nmu;   // dmu * (2 ell + 1), ell = 2l
        out[(((long long)(b0 + e) * n_slots + slot) * g.n_ell + l) * g.nk_pad + ik] = sum * pk_lin;
      }
    }
  }
#pragma unroll
  for (int e = 0; e < NB; ++e)
    if (bad[e] && b0 + e < B) atomicOr(status + b0 + e, bad[e]);
}

// ---- fast path of the fused kernel ------------------------------------------------------
// Same thread mapping as pk_ell_kernel with a fixed CTA shape (32 k lanes x 2 mu groups x 4
// vector groups, NB = 2 vectors per thread), but every option the generic kernel tests per
// grid point is a template parameter and the inner loop is branch free:
//  * mu and the four Legendre weights of every mu row sit in shared memory (warp-broadcast
//    LDS with immediate offsets instead of per-row global loads and 64-bit address math);
//  * the loop over mu rows has an exact trip count (nmu % 10 == 0) and the grid loads of 5
//    rows are issued before their arithmetic;
//  * velocity dispersion is ONE special-function evaluation per point and vector:
//    Lorentzians of both tracers as rsqrt((1 + x0^2)(1 + x1^2)), Gaussians as exp(g mu^2).
// KIND:
//   0 no HCD, same tracer      t^2            1 no HCD, two tracers      tA * tB
//   2 HCD grid, same tracer    tH^2           3 HCD grid on one tracer   tH * tP
//   4 as 3 with the Kaiser term of the HCD tracer divided by its effective bias
//     (fast_metals, power_spectrum.py:219-220 with the effective bias of :305-307)
// VMODE: 0 no per-vector factor, 1 Lorentzian on one tracer, 2 Lorentzian on both, 3 Gaussian
constexpr int PK_FAST_U = 5, PK_FAST_MG = 2, PK_FAST_EG = 4;
enum { PK_V_NONE = 0, PK_V_LOR1 = 1, PK_V_LOR2 = 2, PK_V_GAUSS = 3 };

// 1/sqrt(a) for finite a >= 1: MUFU.RSQ64H seed on the high word (relative error e < 2^-19)
// and one third-order correction y (1 + e/2 + 3 e^2/8), leaving O(e^3) < 2^-58; returned as
// (y, q) with 1/sqrt(a) = y + y q so that the caller folds y into its own product.
__device__ __forceinline__ void rsqrt_ge1_parts(double a, double& y, double& q) {
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(a));
  const double e = fma(-a * y, y, 1.0);
  q = e * fma(0.375, e, 0.5);
}

inline size_t pk_fast_smem_bytes(int nmu) {
  return (size_t)nmu * 5 * 8 + (size_t)PK_GROUPS * 2 * PK_MAX_ELL * (PK_KLANES + 1) * 8;
}

// TWIN = 1: the two accumulator sets of a thread belong to ONE vector and to TWO evaluations
// (eval_ids[z] and twin_ids[z], vg_eval_desc::pk_twin) that differ only by the static grid and
// the linear P(k): Kaiser / HCD / velocity dispersion are evaluated once per grid point for
// both (23 instead of 2 x 18 FP64 instructions per point and vector with two Lorentzians).
template <int NB, int VMODE, int KIND, int TWIN>
__global__ void __launch_bounds__(PK_KLANES * PK_GROUPS, 2)
pk_ell_fast_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, int n_eval,
                   const int* __restrict__ eval_ids, const int* __restrict__ twin_ids,
                   const int* __restrict__ out_slot, int n_slots,
                   const double* __restrict__ pfull, int npf, int B,
                   double* __restrict__ out) {
  constexpr int NV = TWIN ? 1 : NB;      // vectors per thread
  constexpr int MG = PK_FAST_MG, EG = PK_FAST_EG, U = PK_FAST_U;
  extern __shared__ __align__(16) double pk_sm[];
  __shared__ vg_eval_desc d;
  const int nmu = g.nmu;
  double2* wt = reinterpret_cast<double2*>(pk_sm);          // [nmu][2]: (w0,w1), (w2,w3)
  double* mus = pk_sm + 4 * nmu;                            // [nmu]
  typedef double RedRow[NB * PK_MAX_ELL][PK_KLANES + 1];
  RedRow* red = reinterpret_cast<RedRow*>(pk_sm + 5 * nmu); // [PK_GROUPS]
  const int tx = threadIdx.x, mg = threadIdx.y, eg = threadIdx.z;
  const int tid = (eg * MG + mg) * PK_KLANES + tx;
  const int eval = eval_ids[blockIdx.z];
  const int slot = out_slot[eval];
  const int eval2 = TWIN ? twin_ids[blockIdx.z] : eval;
  {
    const int* src = reinterpret_cast<const int*>(evals + eval);
    int* dst = reinterpret_cast<int*>(&d);
    for (int i = tid; i < (int)(sizeof(vg_eval_desc) / 4); i += PK_KLANES * PK_GROUPS) dst[i] = src[i];
    for (int i = tid; i < nmu; i += PK_KLANES * PK_GROUPS) {
      wt[2 * i] = make_double2(g.wleg[i], g.wleg[nmu + i]);
      wt[2 * i + 1] = make_double2(g.wleg[2 * nmu + i], g.wleg[3 * nmu + i]);
      mus[i] = g.mu[i];
    }
  }
  __syncthreads();
  const int ik_raw = blockIdx.y * PK_KLANES + tx;
  const int ik = min(ik_raw, g.nk - 1);
  const double k = g.k[ik];
  const int b0 = (blockIdx.x * EG + eg) * NV;
  constexpr bool HCD = KIND >= 2;
  constexpr bool TWO = KIND == 1 || KIND >= 3;
  const int tH = (KIND >= 3 && !d.hcd_on[0]) ? 1 : 0;   // tracer carrying the HCD term
  const int tP = 1 - tH;

  double aH0[NV], aH1[NV], hH0[NV], hH1[NV], aP0[NV], aP1[NV], c0[NV], c1[NV];
#pragma unroll
  for (int e = 0; e < NV; ++e) {
    const double* p = pfull + (long long)min(b0 + e, B - 1) * npf;
    const KaiserCoef kc = kaiser_setup(d, p, k);
    aH0[e] = tH ? kc.a0[1] : kc.a0[0]; aH1[e] = tH ? kc.a1[1] : kc.a1[0];
    hH0[e] = tH ? kc.h0[1] : kc.h0[0]; hH1[e] = tH ? kc.h1[1] : kc.h1[0];
    aP0[e] = tP ? kc.a0[1] : kc.a0[0]; aP1[e] = tP ? kc.a1[1] : kc.a1[0];
    // (k sigma_t)^2 per tracer; a Gaussian pair collapses to one exponent -(x0^2 + x1^2)/4
    const double s0 = (VMODE != PK_V_NONE && d.idx_velo[0] >= 0) ? k * p[d.idx_velo[0]] : 0.0;
    const double s1 = (VMODE != PK_V_NONE && d.idx_velo[1] >= 0) ? k * p[d.idx_velo[1]] : 0.0;
    if (VMODE == PK_V_GAUSS) { c0[e] = -0.25 * (s0 * s0 + s1 * s1); c1[e] = 0.0; }
    else if (VMODE == PK_V_LOR1) { c0[e] = s0 * s0 + s1 * s1; c1[e] = 0.0; }   // one of them is 0
    else { c0[e] = s0 * s0; c1[e] = s1 * s1; }
  }
  const bool has_s = d.static_grid >= 0;
  const double* __restrict__ S = has_s ? g.grids[d.static_grid] : g.k;
  const int sg2 = TWIN ? evals[eval2].static_grid : -1;
  const bool has_s2 = sg2 >= 0;
  const double* __restrict__ S2 = has_s2 ? g.grids[sg2] : g.k;
  const double* __restrict__ Fg = HCD ? g.grids[d.hcd_grid] : g.k;
  const unsigned rs = (unsigned)(MG * g.nk);      // elements between the rows of one thread
  unsigned off = (unsigned)(mg * g.nk + ik);
  const double2* wrow = wt + 2 * mg;
  const double* mrow = mus + mg;

  double acc[NB][PK_MAX_ELL];
#pragma unroll
  for (int e = 0; e < NB; ++e)
#pragma unroll
    for (int l = 0; l < PK_MAX_ELL; ++l) acc[e][l] = 0.0;

  const int n_batches = nmu / (MG * U);
  for (int jb = 0; jb < n_batches; ++jb) {
    double s[U], f[U], s2[TWIN ? U : 1];
#pragma unroll
    for (int u = 0; u < U; ++u) {   // the L2-latency loads of the whole batch first
      s[u] = has_s ? __ldg(S + (off + u * rs)) : 1.0;
      f[u] = HCD ? __ldg(Fg + (off + u * rs)) : 0.0;
      if (TWIN) s2[u] = has_s2 ? __ldg(S2 + (off + u * rs)) : 1.0;
    }
    off += U * rs;
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const double mu = mrow[u * MG];
      const double2 wa = wrow[u * MG * 2], wb = wrow[u * MG * 2 + 1];
      const double mu2 = mu * mu;
#pragma unroll
      for (int e = 0; e < NV; ++e) {
        double t;
        if (HCD) {
          const double tb = fma(hH0[e], f[u], aH0[e]);
          t = fma(fma(hH1[e], f[u], aH1[e]), mu2, tb);
          if (KIND == 4) t /= tb;
        } else {
          t = fma(aH1[e], mu2, aH0[e]);
        }
        double P = TWO ? t * fma(aP1[e], mu2, aP0[e]) : t * t;
        if (!TWIN) P *= s[u];
        if (VMODE == PK_V_LOR1 || VMODE == PK_V_LOR2) {
          double a = fma(c0[e], mu2, 1.0);
          if (VMODE == PK_V_LOR2) a *= fma(c1[e], mu2, 1.0);
          double y, q;
          rsqrt_ge1_parts(a, y, q);
          const double Py = P * y;
          P = fma(Py, q, Py);
        } else if (VMODE == PK_V_GAUSS) {
          P *= exp(c0[e] * mu2);
        }
        if (TWIN) {
          const double P0 = P * s[u], P1 = P * s2[u];
          acc[0][0] = fma(wa.x, P0, acc[0][0]);
          acc[0][1] = fma(wa.y, P0, acc[0][1]);
          acc[0][2] = fma(wb.x, P0, acc[0][2]);
          acc[0][3] = fma(wb.y, P0, acc[0][3]);
          acc[1][0] = fma(wa.x, P1, acc[1][0]);
          acc[1][1] = fma(wa.y, P1, acc[1][1]);
          acc[1][2] = fma(wb.x, P1, acc[1][2]);
          acc[1][3] = fma(wb.y, P1, acc[1][3]);
        } else {
          acc[e][0] = fma(wa.x, P, acc[e][0]);
          acc[e][1] = fma(wa.y, P, acc[e][1]);
          acc[e][2] = fma(wb.x, P, acc[e][2]);
          acc[e][3] = fma(wb.y, P, acc[e][3]);
        }
      }
    }
    wrow += U * MG * 2;
    mrow += U * MG;
  }
  for (int turn = 0; turn < EG; ++turn) {
    __syncthreads();
    if (eg == turn) {
#pragma unroll
      for (int e = 0; e < NB; ++e)
#pragma unroll
        for (int l = 0; l < PK_MAX_ELL; ++l) red[mg][e * PK_MAX_ELL + l][tx] = acc[e][l];
    }
    __syncthreads();
    if (eg == turn) {
      const double pk_lin = g.pk_lin[d.pk_lin_sel][ik];
      const double pk_lin2 = TWIN ? g.pk_lin[evals[eval2].pk_lin_sel][ik] : pk_lin;
      const int slot2 = TWIN ? out_slot[eval2] : slot;
      for (int o = mg; o < NB * PK_MAX_ELL; o += MG) {
        const int e = o / PK_MAX_ELL, l = o % PK_MAX_ELL;
        const int b = TWIN ? b0 : b0 + e;          // TWIN: e selects the evaluation, not the vector
        if (b >= B || l >= g.n_ell || ik_raw >= g.nk) continue;
        double sum = 0.0;
        for (int m = 0; m < MG; ++m) sum += red[m][o][tx];
        const int sl = (TWIN && e == 1) ? slot2 : slot;
        out[(((long long)b * n_slots + sl) * g.n_ell + l) * g.nk_pad + ik] = sum * ((TWIN && e == 1) ? pk_lin2 : pk_lin);
      }
    }
  }
}

// Materialising K1 (HBM-write bound): P(k,mu) of ONE evaluation for every vector of the
// batch, out[b][nmu][nk].  grid = (ceil(nk/128), nmu, B), block = 128.
__global__ void __launch_bounds__(128)
pk_grid_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, int eval,
               const double* __restrict__ pfull, int npf, double* __restrict__ out,
               int* __restrict__ status) {
  const vg_eval_desc& d = evals[eval];
  const int ik = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y, b = blockIdx.z;
  if (ik >= g.nk) return;
  const double* p = pfull + (long long)b * npf;
  const double k = g.k[ik], mu = g.mu[i];
  KaiserCoef kc = kaiser_setup(d, p, k);
  double F = 0.0;
  if (d.hcd_model)
    F = d.hcd_grid >= 0 ? g.grids[d.hcd_grid][(long long)i * g.nk + ik] : hcd_profile(d, g, kc.L0, k * mu);
  double P = g.pk_lin[d.pk_lin_sel][ik] * kaiser_eval(d, kc, F, mu * mu);
  if (d.static_grid >= 0) P *= g.grids[d.static_grid][(long long)i * g.nk + ik];
  int bad = 0;
  if (d.dyn_mask) P *= pk_factors(d.dyn_mask, d, p, k, mu, g.pk_fid[ik], &bad);
  if (bad) atomicOr(status + b, bad);
  out[((long long)b * g.nmu + i) * g.nk + ik] = P;
}

// Plan-time: static factor grid of one evaluation from the fiducial parameters
// (mode 0: product of d.static_mask factors; mode 1: F_hcd(k_par) at the fiducial L0).
__global__ void __launch_bounds__(128)
static_grid_kernel(PkGridDev g, const vg_eval_desc* __restrict__ evals, int eval, int mode,
                   const double* __restrict__ base, double* __restrict__ out, int* __restrict__ bad_out) {
  const vg_eval_desc& d = evals[eval];
  const int ik = blockIdx.x * 128 + threadIdx.x, i = blockIdx.y;
  if (ik >= g.nk) return;
  const double k = g.k[ik], mu = g.mu[i];
  int bad = 0;
  double v;
  if (mode == 0) v = pk_factors(d.static_mask, d, base, k, mu, g.pk_fid[ik], &bad);
  else v = hcd_profile(d, g, par(base, d.idx_L0, 1.0), k * mu);
  if (bad) atomicOr(bad_out, bad);
  out[(long long)i * g.nk + ik] = v;
}

// pfull[b][:] = base[:] with the sampled columns overwritten by theta[b][:]
// (the merge of vega/vega_interface.py:389-404)
__global__ void expand_theta_kernel(const double* __restrict__ base, int npf,
                                    const double* __restrict__ theta, int n_theta,
                                    const int* __restrict__ cols, long long B,
                                    double* __restrict__ pfull) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * npf) return;
  pfull[idx] = base[idx % npf];
}
// offset / force_one (nullable, per theta column): the blinding transform of
// vega/vega_interface.py:405-419 -- theta += pi - exp(v^2) for the blinded parameters
// (vega/utils.py:375-393) and the BLIND_FIXED_PARS pinned to 1 whatever the caller passed
__global__ void scatter_theta_kernel(const double* __restrict__ theta, int n_theta,
                                     const int* __restrict__ cols, long long B, int npf,
                                     const double* __restrict__ offset, const int* __restrict__ force_one,
                                     double* __restrict__ pfull) {
  long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= B * n_theta) return;
  long long b = idx / n_theta;
  int j = (int)(idx % n_theta);
  double v = theta[idx];
  if (offset) v += offset[j];
  if (force_one && force_one[j]) v = 1.0;
  pfull[b * npf + cols[j]] = v;
}

}  // namespace vg
