// K3: Alcock-Paczynski rescaling of the (r,mu) bins, cubic-spline read-out of xi_l(r),
// Legendre sum, redshift evolution and the additive systematics, for a batch of vectors.
//
// Restates vega/correlation_func.py:117-236 (compute / compute_core / _rescale_coords),
// :276-370 (bias evolution), :446-551 (QSO radiation, relativistic, asymmetry),
// :628-686 (UV shot noise) and the read-out of vega/pktoxi.py:144-163
// (interp1d(kind='cubic') in log r == the not-a-knot interpolating cubic spline; its
// second derivatives arrive from the Hankel GEMM as a linear function of P_l(k)).
//
// One CTA handles one (vector, evaluation) pair and a chunk of bins; the knot table of
// that pair ([n_ell][2][n_out_pad] doubles, ~52 KB) is staged in shared memory with one
// bulk async copy (TMA, cp.async.bulk + mbarrier) per multipole, then each thread gathers
// its bins from shared memory.
#pragma once
#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_pk.cuh"

namespace vg {

struct HankelDev {
  int n_ell, n_out, n_out_pad, bounds_error;
  int uniform;             // knots are log-uniform to 1e-7 of a spacing (FFTLog grids are): the fast
                           // kernel finds the interval from one multiply and at most one step
  int ells[PK_MAX_ELL];
  const double* xk;        // [n_ell][n_out_pad] knots (log r)
  const double* inv_h;     // [n_ell][n_out_pad] 1 / (xk[j+1] - xk[j]), host-computed
  // per multipole: ends of the knot vector and (n_out-1)/(x1-x0), the first guess of the
  // interval search (host-computed, so they sit in the constant bank, not in registers)
  double x0[PK_MAX_ELL], x1[PK_MAX_ELL], inv_dx[PK_MAX_ELL];
};

struct StaticXiDev {       // [which = rel l=1, rel l=3, asy l=0, asy l=2][pk_lin_sel]
  int n[12];
  const double* tab[12];   // [3][n]: knots, values, second derivatives
};

struct XiArgs {
  HankelDev h;
  const vg_eval_desc* evals; int n_eval;
  const double* pfull; int npf; int B;
  // knot tables [B][n][n_ell][2*n_out_pad] from the Hankel GEMM (n_full per vector) and from
  // the moment form (n_mom per vector); tab_slot[eval] >= 0: index in the first, < 0:
  // -(index + 1) in the second
  const double* tables; int n_full;
  const double* tables_mom; int n_mom;
  long long mom_stride_b, mom_stride_j;   // element strides of tables_mom over vector / table
  const int* tab_slot;
  const int* gather_ids;   // evaluations gathered by this launch (grid.y entries)
  const int* metal_map;    // metal slot (out_slot - 2) -> column block of xmet
  int unit_bias;           // 1: leave b1*b2 / extra bias out of the weight (finalize pass of the
                           //    frozen fast_metals terms); 2: weight 1 altogether (CorrelationFunction.compute
                           //    on a caller's P(k,mu): vega/correlation_func.py:117-161 knows no weights)
  const double* const* coords;   // coordinate tables, each [6][n_model]
  const double* opz_eff2;  // per coordinate table: (1+z_eff)^2
  int n_model;
  StaticXiDev sx;
  const double* uv_tau; const double* uv_a; int uv_n;
  // outputs
  double* core;            // [2][B][ld_model]   slot 0 peak, 1 smooth
  double* xmet;            // [B][n_metal*ld_model]
  int ld_model, n_metal;
  double* flat_out;        // non-null (plan time): out = flat_out[(blockIdx.y * B + b) * ld_model]
  int* status;
  int bins_per_cta;
  // fast kernel: per-evaluation rows evolfac[e][ld_model] = bias evolution x growth of every bin,
  // built at finalize for the evaluations whose exponents are not sampled (vg_eval_desc::static_evol)
  const double* evolfac;
};

__device__ __forceinline__ double legendre_eval(int l, double x) {
  const double x2 = x * x;
  switch (l) {
    case 0: return 1.0;
    case 1: return x;
    case 2: return 1.5 * x2 - 0.5;
    case 3: return (2.5 * x2 - 1.5) * x;
    case 4: return (35.0 * x2 * x2 - 30.0 * x2 + 3.0) * 0.125;
    case 6: return (((231.0 * x2 - 315.0) * x2 + 105.0) * x2 - 5.0) * 0.0625;
    case 5: return ((63.0 * x2 - 70.0) * x2 + 15.0) * x * 0.125;
    default: {   // recurrence for anything higher
      double p0 = 1.0, p1 = x;
      for (int n = 2; n <= l; ++n) { double pn = ((2 * n - 1) * x * p1 - (n - 1) * p0) / n; p0 = p1; p1 = pn; }
      return p1;
    }
  }
}

// interval of x in the (almost uniform) knot vector; returns -1/-2 when x is outside
__device__ __forceinline__ int knot_interval(const double* xk, int n, double x, bool* outside) {
  const double x0 = xk[0], x1 = xk[n - 1];
  *outside = !(x >= x0 && x <= x1);
  int j = (int)floor((x - x0) * (double)(n - 1) / (x1 - x0));
  j = max(0, min(n - 2, j));
  while (j > 0 && x < xk[j]) --j;
  while (j < n - 2 && x >= xk[j + 1]) ++j;
  return j;
}
// same with the first guess from a precomputed (n-1)/(x1-x0); the correction loops make the
// result independent of the guess
__device__ __forceinline__ int knot_interval_inv(const double* xk, int n, double x, double x0, double x1,
                                                 double inv_dx, bool* outside) {
  *outside = !(x >= x0 && x <= x1);
  int j = (int)floor((x - x0) * inv_dx);
  j = max(0, min(n - 2, j));
  while (j > 0 && x < xk[j]) --j;
  while (j < n - 2 && x >= xk[j + 1]) ++j;
  return j;
}

__device__ __forceinline__ double spline_eval(const double* xk, const double* y, const double* m,
                                              int j, double x) {
  const double h = xk[j + 1] - xk[j];
  const double t = (x - xk[j]) / h, a = 1.0 - t;
  return a * y[j] + t * y[j + 1] + ((a * a * a - a) * m[j] + (t * t * t - t) * m[j + 1]) * (h * h * (1.0 / 6.0));
}

// (ap, at) from the fit parametrisation: vega/scale_parameters.py:79-125
__device__ __forceinline__ void scale_ap_at(int form, const int* idx, const double* p,
                                            double& ap, double& at) {
  ap = 1.0; at = 1.0;
  if (form == 1) { ap = p[idx[0]]; at = p[idx[1]]; }
  else if (form == 2) { double a = p[idx[0]], e = p[idx[1]]; ap = a * (1 + e) * (1 + e); at = a / (1 + e); }
  else if (form == 3) { double phi = p[idx[0]], al = p[idx[1]]; ap = al / sqrt(phi); at = al * sqrt(phi); }
}

// vega/correlation_func.py:200-236
__device__ __forceinline__ void rescale_coords(double r, double mu, double ap, double at, double drp,
                                               double& r2, double& mu2) {
  if (r == 0.0) { r2 = 0.0; mu2 = 0.0; return; }
  const double rp = ap * (r * mu + drp);
  const double rt = at * (r * sqrt(1.0 - mu * mu));
  r2 = sqrt(rp * rp + rt * rt);
  mu2 = rp / r2;
}

// same with the reciprocal knot spacing supplied (one multiplication instead of a division)
__device__ __forceinline__ double spline_eval_inv(const double* xk, const double* y, const double* m,
                                                  int j, double x, double inv_h) {
  const double h = xk[j + 1] - xk[j];
  const double t = (x - xk[j]) * inv_h, a = 1.0 - t;
  return a * y[j] + t * y[j + 1] + ((a * a * a - a) * m[j] + (t * t * t - t) * m[j + 1]) * (h * h * (1.0 / 6.0));
}

__device__ __forceinline__ double static_xi_eval(const StaticXiDev& sx, int which, double logr) {
  const int n = sx.n[which];
  const double* t = sx.tab[which];
  bool outside;
  int j = knot_interval(t, n, logr, &outside);
  return spline_eval(t, t + n, t + 2 * n, j, logr);
}

__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return (unsigned)__cvta_generic_to_shared(p);
}

__global__ void __launch_bounds__(256) xi_gather_kernel(XiArgs a) {
  extern __shared__ __align__(128) double sm_tab[];   // [n_ell][2*n_out_pad]
  __shared__ __align__(8) unsigned long long bar;
  __shared__ vg_eval_desc d;
  const int tid = threadIdx.x;
  const int e = a.gather_ids[blockIdx.y], b = blockIdx.z;
  {
    const int* src = reinterpret_cast<const int*>(a.evals + e);
    int* dst = reinterpret_cast<int*>(&d);
    for (int i = tid; i < (int)(sizeof(vg_eval_desc) / 4); i += 256) dst[i] = src[i];
  }
  const int row = 2 * a.h.n_out_pad;
  const unsigned row_bytes = (unsigned)row * 8u;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
                 ::"r"(smem_u32(&bar)), "r"(row_bytes * (unsigned)a.h.n_ell) : "memory");
    const int slot = a.tab_slot[e];
    const double* src = slot >= 0 ? a.tables + ((long long)b * a.n_full + slot) * a.h.n_ell * row
                                  : a.tables_mom + b * a.mom_stride_b + (-slot - 1) * a.mom_stride_j;
    for (int l = 0; l < a.h.n_ell; ++l) {
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
          ::"r"(smem_u32(sm_tab + l * row)), "l"(src + (long long)l * row), "r"(row_bytes),
          "r"(smem_u32(&bar)) : "memory");
    }
  }
  // per-vector scalars while the copy is in flight
  const double* p = a.pfull + (long long)b * a.npf;
  double ap, at;
  scale_ap_at(d.scale_form, d.idx_scale, p, ap, at);
  const double drp = par(p, d.idx_drp, 0.0);
  const double* ct = a.coords[d.coord_set];
  const double opz_eff2 = a.opz_eff2[d.coord_set];
  const int n = a.n_model;
  // (1+z)^alpha evolution per tracer; rows 2 and 5 of the coordinate table hold
  // log((1+z_t)/(1+z_eff)) of tracer 1 and 2 (equal unless new-bias-evolution separates the
  // redshifts of the two tracers of a cross-correlation, vega/correlation_func.py:238-299)
  const double zexp0 = d.zevol_kind[0] == 0 ? p[d.idx_alpha[0]] : 0.0;
  const double zexp1 = d.zevol_kind[1] == 0 ? p[d.idx_alpha[1]] : 0.0;
  const bool croom = d.zevol_kind[0] == 1 || d.zevol_kind[1] == 1;
  const int n_croom = (d.zevol_kind[0] == 1) + (d.zevol_kind[1] == 1);
  double weight = a.unit_bias == 2 ? 1.0 : d.weight_const;
  if (d.weight_bias && !a.unit_bias) {
    double b1, be1, b2, be2;
    tracer_bias_beta(d, p, 0, b1, be1);
    tracer_bias_beta(d, p, 1, b2, be2);
    weight *= b1 * b2;
  }
  if (d.idx_weight_extra >= 0 && !a.unit_bias) weight *= p[d.idx_weight_extra];
  if (d.idx_weight_extra2 >= 0 && !a.unit_bias) weight *= p[d.idx_weight_extra2];
  double* out;
  if (a.flat_out) out = a.flat_out + ((long long)blockIdx.y * a.B + b) * a.ld_model;
  else if (d.out_slot < 2) out = a.core + ((long long)d.out_slot * a.B + b) * a.ld_model;
  else out = a.xmet + (long long)b * a.n_metal * a.ld_model + (long long)a.metal_map[d.out_slot - 2] * a.ld_model;

  {  // wait for the table (phase 0)
    unsigned done = 0;
    while (!done) {
      asm volatile(
          "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
          : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
  }

  int bad = 0;
  const int i0 = blockIdx.x * a.bins_per_cta;
  const int i1 = min(a.ld_model, i0 + a.bins_per_cta);
  for (int i = i0 + tid; i < i1; i += 256) {
    if (i >= n) { out[i] = 0.0; continue; }   // keep the GEMM pad columns finite
    const double r = ct[i], mu = ct[n + i];
    double r2, mu2;
    rescale_coords(r, mu, ap, at, drp, r2, mu2);
    double xi = 0.0;
    if (r2 != 0.0) {
      const double x = log(r2);
      for (int l = 0; l < a.h.n_ell; ++l) {
        if (d.single_ell >= 0 && a.h.ells[l] != d.single_ell) continue;
        const double* xk = a.h.xk + l * a.h.n_out_pad;
        bool outside;
        const int j = knot_interval_inv(xk, a.h.n_out, x, a.h.x0[l], a.h.x1[l], a.h.inv_dx[l], &outside);
        if (outside && a.h.bounds_error) bad |= VG_STATUS_BOUNDS;
        const double v = spline_eval_inv(xk, sm_tab + l * row, sm_tab + l * row + a.h.n_out_pad, j, x,
                                         __ldg(a.h.inv_h + l * a.h.n_out_pad + j));
        xi += d.single_ell >= 0 ? v : v * legendre_eval(a.h.ells[l], mu2);
      }
    }
    // redshift evolution of the biases (:276-370) and growth (:372-403)
    double evol = exp(fma(zexp0, ct[2 * n + i], zexp1 * ct[5 * n + i]));
    if (croom) {
      const double p0 = p[d.idx_croom[0]], p1 = p[d.idx_croom[1]];
      const double c = (p0 + p1 * ct[3 * n + i]) / (p0 + p1 * opz_eff2);
      evol *= n_croom == 2 ? c * c : c;
    }
    xi *= evol;
    xi *= ct[4 * n + i];

    if (d.radiation && !d.is_peak) {        // :446-489
      double rp, rt;
      if (d.rescale_sys) { rp = r2 * mu2 + drp; rt = r2 * sqrt(1.0 - mu2 * mu2); }
      else { rp = r * mu + drp; rt = r * sqrt(1.0 - mu * mu); }
      const double rs = sqrt(rp * rp + rt * rt), ms = rp / rs;
      double v = p[d.idx_qso_rad[0]] / (rs * rs) * (1.0 - p[d.idx_qso_rad[1]] * (1.0 - ms * ms));
      v *= exp(-rs * ((1.0 + ms) / p[d.idx_qso_rad[2]] + 1.0 / p[d.idx_qso_rad[3]]));
      xi += v;
    }
    if (d.rel || d.asy) {                   // :491-551
      double ap3, at3, r3, mu3;
      scale_ap_at(d.scale_form_sys, d.idx_scale_sys, p, ap3, at3);
      rescale_coords(r, mu, ap3, at3, drp, r3, mu3);
      const double lx = log(r3);
      if (d.rel) {
        xi += p[d.idx_Arel[0]] * static_xi_eval(a.sx, 0 * 3 + d.pk_lin_sel, lx) * legendre_eval(1, mu3)
            + p[d.idx_Arel[1]] * static_xi_eval(a.sx, 1 * 3 + d.pk_lin_sel, lx) * legendre_eval(3, mu3);
      }
      if (d.asy) {
        const double t0 = static_xi_eval(a.sx, 2 * 3 + d.pk_lin_sel, lx), t2 = static_xi_eval(a.sx, 3 * 3 + d.pk_lin_sel, lx);
        xi += (p[d.idx_Aasy[0]] * t0 - p[d.idx_Aasy[1]] * t2) * r3 * legendre_eval(1, mu3)
            + p[d.idx_Aasy[2]] * t2 * r3 * legendre_eval(3, mu3);
      }
    }
    if (d.uv_shot) {                        // :628-686
      const double lam = p[d.idx_uvshot[0]], bg = p[d.idx_uvshot[1]];
      const double rr = d.rescale_sys ? sqrt(r2 * r2 + mu2 * mu2) : r;
      const double tau = rr / lam;
      double av;
      if (tau < a.uv_tau[0]) av = a.uv_a[0];
      else if (tau > a.uv_tau[a.uv_n - 1]) av = 0.0;
      else {
        int lo = 0, hi = a.uv_n - 1;
        while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (a.uv_tau[mid] <= tau) lo = mid; else hi = mid; }
        av = (a.uv_a[hi] - a.uv_a[lo]) / (a.uv_tau[hi] - a.uv_tau[lo]) * (tau - a.uv_tau[lo]) + a.uv_a[lo];
      }
      xi += bg * bg * p[d.idx_uvshot[2]] * lam / rr * av;
    }
    if (!isfinite(xi)) bad |= VG_STATUS_NONFINITE;
    out[i] = weight * xi;
  }
  if (bad) atomicOr(a.status + b, bad);
}

// Fast variant of xi_gather_kernel for the standard case: multipoles (0, 2, 4, 6), log-uniform
// knots, no relativistic / asymmetry / UV-shot term, power-law bias evolution, all multipoles
// summed.  Same arithmetic, restructured so that a bin costs ~1/4 of the instructions:
//  * the multipole loop is unrolled with its per-multipole constants in the constant bank;
//  * the knot interval comes from one multiply and at most one +-1 step (the generic kernel
//    walks two correction loops), the reciprocal spacing from the mean spacing and one Newton
//    step (relative error eps^2, eps < 1e-7 the checked non-uniformity) instead of a table read;
//  * r mu and r sqrt(1 - mu^2) of every bin are rows 6 and 7 of the coordinate table, r' and mu'
//    come from one rsqrt, log r' = log(r'^2) / 2;
//  * when the evolution exponents are not sampled the per-bin factor (1+z)^alpha x growth is read
//    from a row built at finalize instead of an exp() per bin.
__global__ void __launch_bounds__(256) xi_gather_fast_kernel(XiArgs a) {
  extern __shared__ __align__(128) double sm_tab[];   // [4][2*n_out_pad]
  __shared__ __align__(8) unsigned long long bar;
  __shared__ vg_eval_desc d;
  const int tid = threadIdx.x;
  const int e = a.gather_ids[blockIdx.y], b = blockIdx.z;
  {
    const int* src = reinterpret_cast<const int*>(a.evals + e);
    int* dst = reinterpret_cast<int*>(&d);
    for (int i = tid; i < (int)(sizeof(vg_eval_desc) / 4); i += 256) dst[i] = src[i];
  }
  const int row = 2 * a.h.n_out_pad;
  const unsigned row_bytes = (unsigned)row * 8u;
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;\n" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
  }
  __syncthreads();
  if (tid == 0) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n"
                 ::"r"(smem_u32(&bar)), "r"(row_bytes * 4u) : "memory");
    const int slot = a.tab_slot[e];
    const double* src = slot >= 0 ? a.tables + ((long long)b * a.n_full + slot) * 4 * row
                                  : a.tables_mom + b * a.mom_stride_b + (-slot - 1) * a.mom_stride_j;
#pragma unroll
    for (int l = 0; l < 4; ++l) {
      asm volatile(
          "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n"
          ::"r"(smem_u32(sm_tab + l * row)), "l"(src + (long long)l * row), "r"(row_bytes),
          "r"(smem_u32(&bar)) : "memory");
    }
  }
  const double* p = a.pfull + (long long)b * a.npf;
  double ap, at;
  scale_ap_at(d.scale_form, d.idx_scale, p, ap, at);
  const double drp = par(p, d.idx_drp, 0.0);
  const double* ct = a.coords[d.coord_set];
  const int n = a.n_model;
  const bool evol_static = d.static_evol != 0 && a.evolfac != nullptr;
  const double* evolrow = a.evolfac + (long long)e * a.ld_model;
  const double zexp0 = p[d.idx_alpha[0]], zexp1 = p[d.idx_alpha[1]];
  double weight = a.unit_bias == 2 ? 1.0 : d.weight_const;
  if (d.weight_bias && !a.unit_bias) {
    double b1, be1, b2, be2;
    tracer_bias_beta(d, p, 0, b1, be1);
    tracer_bias_beta(d, p, 1, b2, be2);
    weight *= b1 * b2;
  }
  if (d.idx_weight_extra >= 0 && !a.unit_bias) weight *= p[d.idx_weight_extra];
  if (d.idx_weight_extra2 >= 0 && !a.unit_bias) weight *= p[d.idx_weight_extra2];
  const bool radiation = d.radiation && !d.is_peak;
  double rad0 = 0.0, rad1 = 0.0, rad2 = 0.0, rad3 = 0.0;
  if (radiation) {
    rad0 = p[d.idx_qso_rad[0]]; rad1 = p[d.idx_qso_rad[1]];
    rad2 = 1.0 / p[d.idx_qso_rad[2]]; rad3 = 1.0 / p[d.idx_qso_rad[3]];
  }
  double* out;
  if (a.flat_out) out = a.flat_out + ((long long)blockIdx.y * a.B + b) * a.ld_model;
  else if (d.out_slot < 2) out = a.core + ((long long)d.out_slot * a.B + b) * a.ld_model;
  else out = a.xmet + (long long)b * a.n_metal * a.ld_model + (long long)a.metal_map[d.out_slot - 2] * a.ld_model;
  const int n_out = a.h.n_out, n_pad = a.h.n_out_pad;

  {  // wait for the table (phase 0)
    unsigned done = 0;
    while (!done) {
      asm volatile(
          "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}\n"
          : "=r"(done) : "r"(smem_u32(&bar)), "r"(0u) : "memory");
    }
  }

  int bad = 0;
  const int i0 = blockIdx.x * a.bins_per_cta;
  const int i1 = min(a.ld_model, i0 + a.bins_per_cta);
  for (int i = i0 + tid; i < i1; i += 256) {
    if (i >= n) { out[i] = 0.0; continue; }
    const double rp0 = ct[6 * n + i], rt0 = ct[7 * n + i];
    const double fac = evol_static ? evolrow[i]
                                   : exp(fma(zexp0, ct[2 * n + i], zexp1 * ct[5 * n + i])) * ct[4 * n + i];
    const double rp = ap * (rp0 + drp), rt = at * rt0;
    const double s = fma(rp, rp, rt * rt);
    double xi = 0.0, r2 = 0.0, mu2 = 0.0;
    if (!(rp0 == 0.0 && rt0 == 0.0) && s > 0.0) {        // r == 0 bins stay 0 (vega/pktoxi.py:147-150)
      const double inv = rsqrt(s);
      r2 = s * inv; mu2 = rp * inv;
      const double x = 0.5 * log(s);
      const double m2 = mu2 * mu2;
      double L[4];
      L[0] = 1.0;
      L[1] = fma(1.5, m2, -0.5);
      L[2] = fma(fma(4.375, m2, -3.75), m2, 0.375);
      L[3] = fma(fma(fma(14.4375, m2, -19.6875), m2, 6.5625), m2, -0.3125);
#pragma unroll
      for (int l = 0; l < 4; ++l) {
        const double x0 = a.h.x0[l], c = a.h.inv_dx[l];
        if (!(x >= x0 && x <= a.h.x1[l]) && a.h.bounds_error) bad |= VG_STATUS_BOUNDS;
        int j = __double2int_rd((x - x0) * c);
        j = max(0, min(n_out - 2, j));
        const double* xk = a.h.xk + l * n_pad;
        double lo = __ldg(xk + j), hi = __ldg(xk + j + 1);
        if (x < lo && j > 0) { --j; hi = lo; lo = __ldg(xk + j); }
        else if (x >= hi && j < n_out - 2) { ++j; lo = hi; hi = __ldg(xk + j + 1); }
        const double h = hi - lo;
        const double t0 = (x - lo) * c;
        const double t = t0 * fma(-h, c, 2.0), am = 1.0 - t;
        const double* y = sm_tab + l * row;
        const double* mm = y + n_pad;
        const double v = am * y[j] + t * y[j + 1] +
                         ((am * am * am - am) * mm[j] + (t * t * t - t) * mm[j + 1]) * (h * h * (1.0 / 6.0));
        xi = fma(v, L[l], xi);
      }
    }
    xi *= fac;
    if (radiation) {        // vega/correlation_func.py:446-489
      double rps, rts;
      if (d.rescale_sys) { rps = r2 * mu2 + drp; rts = r2 * sqrt(1.0 - mu2 * mu2); }
      else { rps = rp0 + drp; rts = rt0; }
      const double ss = fma(rps, rps, rts * rts);
      const double rs = sqrt(ss), ms = rps / rs;
      double v = rad0 / ss * (1.0 - rad1 * (1.0 - ms * ms));
      v *= exp(-rs * ((1.0 + ms) * rad2 + rad3));
      xi += v;
    }
    if (!isfinite(xi)) bad |= VG_STATUS_NONFINITE;
    out[i] = weight * xi;
  }
  if (bad) atomicOr(a.status + b, bad);
}

// evolfac[e][i] = bias evolution x growth of bin i for evaluation e at the plan parameters
// (vega/correlation_func.py:276-403): same expression as the gather kernels evaluate per vector
__global__ void evol_rows_kernel(const vg_eval_desc* __restrict__ evals, int n_eval, const double* __restrict__ base,
                                 const double* const* __restrict__ coords, const double* __restrict__ opz_eff2,
                                 int n, int ld, double* __restrict__ evolfac) {
  const int e = blockIdx.y;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ld) return;
  const vg_eval_desc& d = evals[e];
  double v = 0.0;
  if (i < n && d.static_evol) {
    const double* ct = coords[d.coord_set];
    const double zexp0 = d.zevol_kind[0] == 0 ? base[d.idx_alpha[0]] : 0.0;
    const double zexp1 = d.zevol_kind[1] == 0 ? base[d.idx_alpha[1]] : 0.0;
    v = exp(fma(zexp0, ct[2 * n + i], zexp1 * ct[5 * n + i]));
    const int n_croom = (d.zevol_kind[0] == 1) + (d.zevol_kind[1] == 1);
    if (n_croom) {
      const double p0 = base[d.idx_croom[0]], p1 = base[d.idx_croom[1]];
      const double c = (p0 + p1 * ct[3 * n + i]) / (p0 + p1 * opz_eff2[d.coord_set]);
      v *= n_croom == 2 ? c * c : c;
    }
    v *= ct[4 * n + i];
  }
  evolfac[(long long)e * ld + i] = v;
}

// ---- per-stage entry points with CALLER-SUPPLIED arrays (PktoXi / CorrelationFunction call shapes) ----
// Multipoles of a P(k,mu) grid handed in by the caller (PktoXi.compute_pk_ells, vega/pktoxi.py:79-97):
// out[b][l][k] = sum_i wleg[l][i] P[b][i][k], wleg = dmu (2l+1) L_l(mu_i).  grid (ceil(nk/128), n_ell, B).
__global__ void __launch_bounds__(128)
project_grid_kernel(const double* __restrict__ pk, int nmu, int nk, int nk_pad, int n_ell,
                    const double* __restrict__ wleg, double* __restrict__ out) {
  const int ik = blockIdx.x * 128 + threadIdx.x, l = blockIdx.y;
  const long long b = blockIdx.z;
  if (ik >= nk_pad) return;
  double acc = 0.0;
  if (ik < nk) {
    const double* p = pk + b * (long long)nmu * nk + ik;
    const double* w = wleg + (long long)l * nmu;
    for (int i = 0; i < nmu; ++i) acc = fma(w[i], p[(long long)i * nk], acc);
  }
  out[(b * n_ell + l) * nk_pad + ik] = acc;
}

// xi(r, mu) on arbitrary coordinates from knot tables (PktoXi.compute, vega/pktoxi.py:99-163): spline
// read-out in log r per multipole, times L_l(mu), summed -- or one multipole alone (single_ell >= 0).
// r == 0 gives 0; r outside the transform's range flags VG_STATUS_BOUNDS.  grid (ceil(n/256), B).
__global__ void __launch_bounds__(256)
xi_plain_kernel(HankelDev h, const double* __restrict__ tables, const double* __restrict__ r,
                const double* __restrict__ mu, int n, int single_ell, double* __restrict__ out,
                int* __restrict__ status) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  const long long b = blockIdx.y;
  if (i >= n) return;
  const int row = 2 * h.n_out_pad;
  const double* tab = tables + b * (long long)h.n_ell * row;
  double xi = 0.0;
  int bad = 0;
  if (r[i] != 0.0) {
    const double x = log(r[i]);
    for (int l = 0; l < h.n_ell; ++l) {
      if (single_ell >= 0 && h.ells[l] != single_ell) continue;
      const double* xk = h.xk + l * h.n_out_pad;
      bool outside;
      const int j = knot_interval_inv(xk, h.n_out, x, h.x0[l], h.x1[l], h.inv_dx[l], &outside);
      if (outside && h.bounds_error) bad |= VG_STATUS_BOUNDS;
      const double v = spline_eval(xk, tab + l * row, tab + l * row + h.n_out_pad, j, x);
      xi += single_ell >= 0 ? v : v * legendre_eval(h.ells[l], mu[i]);
    }
  }
  out[b * n + i] = xi;
  if (bad) atomicOr(status + b, bad);
}

}  // namespace vg
