// new_metals: pair histograms behind the metal matrices (vega/metals.py:502-752).
//
// The reference forms every pixel pair of the two stacked-weight tables on the host
// (N1 x N2 ~ 1e7 pairs per metal correlation) and calls np.histogram2d / np.histogram on them
// (:543-556, :566-584, :625-637, :716-728).  Here one kernel walks the pair table once and
// accumulates all of those histograms in shared memory:
//   H[a][t]      sum of w over pairs with assumed rp in bin a and true rp in bin t
//   per_true[t]  sum of w by true-rp bin (pairs whose assumed rp falls outside still count)
//   per_assumed[0..2][a]   sum of w, w * rp_assumed, w * z_true_pair by assumed-rp bin
//   ratio[q]     sum of w / d_true^2 * (|rp_true| < 20) by bin of d_assumed / d_true
// Bins follow numpy: [e_i, e_{i+1}) with the last bin closed on the right; the index is a
// uniform-grid guess corrected against the actual edge values, so assignment is identical
// to np.searchsorted on the same edges.  Accumulation order differs (atomics): the sums
// agree with numpy to rounding (tests: 1e-12 relative).
#pragma once
#include "vg_common.cuh"

namespace vg {

struct PairSide {
  const double* r_true; const double* r_assumed; const double* z_true; const double* z_assumed;
  const double* weight; int n;
};
struct PairArgs {
  PairSide s1, s2;
  int signed_rp;             // a discrete tracer is involved: keep the sign of rp (:472-473)
  double zmin, zmax;
  const double* rp_edges; int nb;
  const double* ratio_edges; int nr;     // nr == 0: skip the distance-ratio histogram
  double* H; double* per_true; double* per_assumed; double* ratio;
  double* minmax;            // [2 * gridDim.x * gridDim.y] per-CTA min / max of the ratio (pass 0)
};

__device__ __forceinline__ int edge_bin(const double* e, int n, double x) {
  const double lo = e[0], hi = e[n];
  if (!(x >= lo && x <= hi)) return -1;
  int g = (int)((x - lo) / (hi - lo) * n);
  g = max(0, min(n - 1, g));
  while (g > 0 && x < e[g]) --g;
  while (g < n - 1 && x >= e[g + 1]) ++g;
  return g;
}

constexpr int PAIR_TI = 32, PAIR_TJ = 256;

// pass 0: range of d_assumed / d_true over ALL pairs (np.histogram with an integer bin count
// takes its range from the data, weights or not)
__global__ void __launch_bounds__(PAIR_TJ) pair_ratio_range_kernel(PairArgs a) {
  __shared__ double smin[PAIR_TJ / 32], smax[PAIR_TJ / 32];
  const int j = blockIdx.x * PAIR_TJ + threadIdx.x;
  double lo = 1e300, hi = -1e300;
  if (j < a.s2.n) {
    const double rt2 = a.s2.r_true[j], ra2 = a.s2.r_assumed[j];
    const int i1 = min(a.s1.n, (blockIdx.y + 1) * PAIR_TI);
    for (int i = blockIdx.y * PAIR_TI; i < i1; ++i) {
      const double q = ((a.s1.r_assumed[i] + ra2) / 2) / ((a.s1.r_true[i] + rt2) / 2);
      lo = fmin(lo, q); hi = fmax(hi, q);
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = lo; smax[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < PAIR_TJ / 32; ++w) { lo = fmin(lo, smin[w]); hi = fmax(hi, smax[w]); }
    const long long cta = (long long)blockIdx.y * gridDim.x + blockIdx.x;
    a.minmax[2 * cta] = lo; a.minmax[2 * cta + 1] = hi;
  }
}

// pass 1: all histograms.  Dynamic shared memory: nb*nb + 4*nb + nr doubles.
__global__ void __launch_bounds__(PAIR_TJ) pair_histogram_kernel(PairArgs a) {
  extern __shared__ double sh[];
  const int nb = a.nb, nr = a.nr;
  double* sH = sh;
  double* sTrue = sH + nb * nb;
  double* sAss = sTrue + nb;          // [3][nb]
  double* sRatio = sAss + 3 * nb;
  const int total = nb * nb + 4 * nb + nr;
  for (int t = threadIdx.x; t < total; t += PAIR_TJ) sh[t] = 0.0;
  __syncthreads();
  const int j = blockIdx.x * PAIR_TJ + threadIdx.x;
  if (j < a.s2.n) {
    const double rt2 = a.s2.r_true[j], ra2 = a.s2.r_assumed[j];
    const double zt2 = a.s2.z_true[j], za2 = a.s2.z_assumed[j], w2 = a.s2.weight[j];
    const int i1 = min(a.s1.n, (blockIdx.y + 1) * PAIR_TI);
    for (int i = blockIdx.y * PAIR_TI; i < i1; ++i) {
      const double rt1 = a.s1.r_true[i], ra1 = a.s1.r_assumed[i];
      const double zpair = (a.s1.z_assumed[i] + za2) / 2.;
      double w = a.s1.weight[i] * w2;
      w *= (zpair >= a.zmin && zpair <= a.zmax) ? 1.0 : 0.0;
      double rp_t = rt1 - rt2, rp_a = ra1 - ra2;
      if (!a.signed_rp) { rp_t = fabs(rp_t); rp_a = fabs(rp_a); }
      const int bt = edge_bin(a.rp_edges, nb, rp_t), ba = edge_bin(a.rp_edges, nb, rp_a);
      if (w != 0.0) {
        if (bt >= 0) atomicAdd(sTrue + bt, w);
        if (ba >= 0) {
          atomicAdd(sAss + ba, w);
          atomicAdd(sAss + nb + ba, w * rp_a);
          atomicAdd(sAss + 2 * nb + ba, w * ((a.s1.z_true[i] + zt2) / 2.));
          if (bt >= 0) atomicAdd(sH + ba * nb + bt, w);
        }
      }
      if (nr > 0) {
        const double dt = (rt1 + rt2) / 2, da = (ra1 + ra2) / 2;
        const double wr = w / (dt * dt) * (fabs(rp_t) < 20. ? 1.0 : 0.0);
        if (wr != 0.0) {
          const int bq = edge_bin(a.ratio_edges, nr, da / dt);
          if (bq >= 0) atomicAdd(sRatio + bq, wr);
        }
      }
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < total; t += PAIR_TJ) {
    const double v = sh[t];
    if (v == 0.0) continue;
    double* dst = t < nb * nb ? a.H + t
                : t < nb * nb + nb ? a.per_true + (t - nb * nb)
                : t < nb * nb + 4 * nb ? a.per_assumed + (t - nb * nb - nb)
                : a.ratio + (t - nb * nb - 4 * nb);
    atomicAdd(dst, v);
  }
}

}  // namespace vg
