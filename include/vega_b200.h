/* vega_b200 -- C ABI of the B200 model-evaluation + likelihood path.
 *
 * The reference (andreicuceu/vega) is pure Python and has no FFI; this header is the
 * boundary a maintainer would bind (ctypes stub in INTEGRATION.md) to replace
 *   VegaInterface.compute_model / chi2 / log_lik   vega/vega_interface.py:208-387
 *   Model.compute                                  vega/model.py:157-187
 *   PowerSpectrum.compute                          vega/power_spectrum.py:87-196
 *   PktoXi.compute / compute_pk_ells               vega/pktoxi.py:79-163
 *   CorrelationFunction.compute                    vega/correlation_func.py:117-198
 *   Metals.compute                                 vega/metals.py:258-336
 *   BroadbandPolynomials.compute                   vega/broadband_poly.py:74-198
 * for a BATCH of parameter vectors.  Plain pointers and sizes only, no torch types.
 *
 * Conventions: every function returns 0 on success (VG_ERR_* otherwise, text from
 * vg_last_error()).  All arithmetic is FP64.  "dev" pointers are device pointers, "host"
 * pointers are host pointers; the caller owns every buffer it passes.  Work is queued on
 * the given CUDA stream (a cudaStream_t passed as void*) and is asynchronous with respect
 * to the host unless stated.  status[b] != 0 means the reference would have raised a
 * VegaModelError for that vector (vega/vega_interface.py:268-279): its chi2 is 1e100.
 */
#ifndef VEGA_B200_H
#define VEGA_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Return codes of every entry point (text of the last failure: vg_last_error()) */
#define VG_OK 0
#define VG_ERR_CUDA 1    /* a CUDA call failed, or there is no CUDA device (no CPU fallback) */
#define VG_ERR_ARG 2     /* bad argument */
#define VG_ERR_STATE 3   /* call out of order (e.g. a new data mask after vg_plan_finalize) */

/* Per-vector status bits written by the kernels into status[b]: the model-domain failures the
 * reference raises as exceptions (vega/utils.py:444-453) and VegaInterface.chi2 turns into 1e100
 * (vega/vega_interface.py:268-279). */
#define VG_STATUS_BOUNDS 1      /* r outside the FFTLog range  -> VegaBoundsError */
#define VG_STATUS_ARINYO 2      /* NaN/Inf in the Arinyo D_NL  -> VegaArinyoError */
#define VG_STATUS_NONFINITE 4   /* a non-finite model bin */
#define VG_STATUS_DEVICE 8      /* internal: a kernel pipeline timed out -- the values are not valid */

#define VG_MAX_BB_TERMS 8

/* P(k,mu) factor bits (vega/power_spectrum.py:118-194) */
#define VG_F_ARINYO   (1u << 0)   /* compute_dnl_arinyo      :435-479 */
#define VG_F_MCDONALD (1u << 1)   /* compute_dnl_mcdonald    :419-433 */
#define VG_F_GK       (1u << 2)   /* compute_Gk              :481-502 */
#define VG_F_PEAKNL   (1u << 3)   /* compute_peak_nl         :382-417 */
#define VG_F_FSGAUSS  (1u << 4)   /* fullshape gauss         :504-553 */
#define VG_F_FSEXP    (1u << 5)   /* fullshape exp           :560-586 */
#define VG_F_VELO     (1u << 6)   /* velocity dispersion     :588-636 */
#define VG_F_KDAMP    (1u << 7)   /* k^n damping  vega/utils.py:423-441 */
#define VG_F_MOCKGK   (1u << 8)   /* mock-bin-size G(k)      :143-160 */

/* One P(k,mu) -> xi evaluation: the (PowerSpectrum, CorrelationFunction) pair the
 * reference builds per component (peak / smooth, vega/model.py:60-66) and per metal
 * correlation (vega/metals.py:95-142).  Every idx_* is a column of the full parameter
 * vector (or -1 when the parameter is absent from the configuration). */
typedef struct vg_eval_desc {
  /* ---- P(k,mu) ---- */
  int32_t pk_lin_sel;       /* 0: pk_full - pk_smooth (peak), 1: pk_smooth, 2: pk_full,
                               3: the P(k) of vg_plan_set_direct_pk */
  int32_t is_peak;
  int32_t fast_metals;      /* Kaiser term without b1*b2 (vega/power_spectrum.py:219-220) */
  int32_t same_tracer;
  int32_t idx_bias[2], idx_beta[2], idx_bias_eta[2];   /* vega/utils.py:45-108 */
  int32_t idx_growth_rate;  /* -1: 0.970386 */
  int32_t uv_on[2];         /* UV / HeII applies to tracer t (LYA only, :111-114) */
  int32_t uvb, heii;
  int32_t idx_lambda_uv, idx_bias_gamma, idx_bias_prim, idx_lambda_heii, idx_bias_gamma_e;
  int32_t hcd_model;        /* 0 none, 1 Rogers2018, 2 fvoigt, 3 sinc  (:263-380) */
  int32_t hcd_on[2];
  int32_t idx_bias_hcd, idx_beta_hcd, idx_L0;   /* idx_L0 -1: L0 = 1 */
  int32_t hcd_grid;         /* >=0: static F(k_par) grid; -1: F evaluated per vector */
  int32_t static_grid;      /* >=0: grid holding the product of `static_mask` factors */
  uint32_t dyn_mask;        /* factors evaluated per parameter vector */
  uint32_t static_mask;     /* factors folded into `static_grid` (their parameters are fixed) */
  int32_t arinyo_mode;      /* 2: LYxLY, 1: one LY tracer (sqrt), 0: none (ones) */
  int32_t idx_arinyo[6];    /* q1 q2 kv av bv kp; q2 = -1 -> 0 */
  int32_t idx_sigmaNL[2];   /* par, per; one may be -1 -> derived with 1+f (:391-405) */
  int32_t idx_fs_gauss[4];  /* par1 per1 par2 per2: G(par1,per1)*G(par2,per2) */
  int32_t idx_fs_exp[4];    /* par_sigma per_sigma par_exp per_exp */
  int32_t velo_kind;        /* 1 gauss, 2 lorentz */
  int32_t idx_velo[2];      /* per tracer, -1 when the tracer is not discrete */
  int32_t kdamp_power;
  double gk_drp, gk_drt;    /* binning sizes; 0 skips that sinc (:493-500) */
  double kdamp_scale;
  /* ---- xi(rp,rt) ---- */
  int32_t coord_set;        /* coordinate table this evaluation is gathered on */
  int32_t scale_form;       /* 0: ap=at=1, 1: ap_at, 2: aiso_epsilon, 3: phi_alpha
                               (vega/scale_parameters.py:38-230, resolved per component) */
  int32_t idx_scale[2];
  int32_t scale_form_sys;   /* same, for the rel/asy terms (get_ap_at without corr name) */
  int32_t idx_scale_sys[2];
  int32_t idx_drp;          /* drp_{QSO} for cross-correlations, -1 otherwise */
  int32_t single_ell;       /* -1: sum of all multipoles */
  int32_t zevol_kind[2];    /* 0 standard power law, 1 Croom (:276-370) */
  int32_t idx_alpha[2];
  int32_t idx_croom[2];
  int32_t radiation;        /* QSO radiation (smooth component only, :446-489) */
  int32_t idx_qso_rad[4];   /* strength asymmetry lifetime decrease */
  int32_t rescale_sys;      /* rescale-coords-systematics */
  int32_t rel, asy;         /* relativistic / standard asymmetry (:491-551) */
  int32_t idx_Arel[2], idx_Aasy[3];
  int32_t uv_shot;          /* UVB shot noise (:628-686) */
  int32_t idx_uvshot[3];    /* lambda_uv, bias_gamma(_e), uv_shotnoise_amp */
  /* ---- output ---- */
  int32_t out_slot;         /* 0: peak core, 1: smooth core, 2+c: metal correlation c */
  int32_t weight_bias;      /* 1: multiply by b1*b2 of the pair (vega/metals.py:316-334) */
  int32_t idx_weight_extra; /* bias_{M1}_{M2} (separate-metal-auto-biases), -1 none */
  double weight_const;      /* 2 for M1 != M2 in an auto-correlation (vega/metals.py:238-239) */
  /* ---- fast_metals caches (vega/metals.py:144-207) ---- */
  int32_t table_owner;      /* >=0: this evaluation has no P(k,mu) of its own and reads the knot
                               tables of evaluation `table_owner` -- the metal x main-tracer cache
                               hit of compute_xi_metal_cross_main (:171-207), whose key holds the
                               betas but not the pair; only weight and metal matrix differ */
  /* ---- mock-bin-size smoothing (vega/power_spectrum.py:143-160) ---- */
  int32_t mock_los_mode;    /* 0 none, 1 growth: rp size * (1+f), 2 amplitude: * (1+los_smooth_amp),
                               3 only-los: transverse size 0 */
  int32_t idx_los_smooth_amp;
  double mock_bin_size;
  int32_t direct;           /* 1: the single `full` component of Model.compute_direct
                               (vega/model.py:189-208): evaluated only in direct-P(k) mode, where it
                               replaces peak, smooth and metals */
  int32_t frozen;           /* 1: metal x metal term of compute_xi_metal_metal (:144-169): M_c xi_c
                               is evaluated once, at the plan's base parameters, in
                               vg_plan_finalize; the hot path adds b1*b2 (x extra) times it */
  int32_t static_xi;        /* 1: nothing the xi stage reads (AP scale, drp, bias-evolution exponents)
                               is a sampled parameter and no additive xi term is on, so the
                               coordinates and per-bin factors are the same for every vector: in
                               moment form the gather is applied once to the static tables */
  int32_t pk_twin;          /* >=0: evaluation `pk_twin` of the same correlation has the same P(k,mu)
                               up to the linear P(k) and the static grid (peak / smooth of one
                               correlation): the two go through the per-point pass together and
                               share the Kaiser / HCD / velocity-dispersion arithmetic.  -1: none */
  int32_t static_evol;      /* 1: the bias-evolution exponents (alpha_*, croom_*) of this evaluation are
                               not sampled: (1+z)^alpha x growth per bin is a row built at finalize */
  int32_t idx_weight_extra2; /* a second parameter multiplying the weight, -1 none: bao_amp on the metal
                               terms of the PEAK component when no-metal-decomp = False
                               (vega/model.py:119-131,186) */
  int32_t uv_static;        /* 1: none of the UV / HeII parameters (lambda_uv, bias_gamma, bias_prim,
                               lambda_heii, bias_gamma_e; vega/power_spectrum.py:224-261) is sampled: the
                               scale-dependent bias term is a static function of k and the evaluation
                               can take the moment form */
} vg_eval_desc;

/* One broadband term (vega/broadband_poly.py:23-198) */
typedef struct vg_bb_term {
  int32_t kind;             /* 0: polynomial sum_ij a_ij x1^i x2^j, 1: Gaussian sky */
  int32_t position;         /* 0 pre-mul, 1 pre-add, 2 post-mul, 3 post-add */
  int32_t n_coef;           /* polynomial: number of (i,j) terms */
  int32_t design_offset;    /* polynomial: first row of the design table for this term */
  int32_t idx_coef_first;   /* unused (coefficients are listed in idx_coef) */
  int32_t idx_sky[2];       /* scale-sky, sigma-sky */
  int32_t idx_coef[32];
} vg_bb_term;

typedef struct vg_plan vg_plan;

const char* vg_last_error(void);
int vg_version(void);
/* sizeof(vg_eval_desc) / sizeof(vg_bb_term) as compiled: bindings check their mirrors */
int vg_sizeof_eval_desc(void);
int vg_sizeof_bb_term(void);

/* ---------------------------------------------------------------------------------
 * Plan = the static state VegaInterface.__init__ builds once (vega/vega_interface.py:36-206):
 * templates, grids, matrices, data, the parameter-name -> column map.  Host pointers
 * are copied; the plan owns its device memory.
 * ------------------------------------------------------------------------------- */

/* n_params: length of the full parameter vector (every entry of VegaInterface.params plus
 * derived columns); base: its fiducial values; theta_cols[n_theta]: the columns a batch row
 * theta[b, 0..n_theta) overwrites (the [sample] set).  k, pk_*: the fiducial template
 * (vega/vega_interface.py:665-703), pk_fid already rescaled to z_eff
 * (vega/power_spectrum.py:72-73).  mu[nmu]: midpoint nodes (vega/power_spectrum.py:76-78);
 * wleg[n_ell][nmu] = dmu*(2l+1)*L_l(mu) (vega/pktoxi.py:55,138). */
int vg_plan_create(int32_t n_params, const double* base, int32_t n_theta, const int32_t* theta_cols,
                   int32_t nk, const double* k, const double* pk_full, const double* pk_smooth,
                   const double* pk_fid, int32_t nmu, const double* mu, int32_t n_ell,
                   const double* wleg, int32_t idx_bao_amp, vg_plan** out);
int vg_plan_destroy(vg_plan* plan);
/* Blinding at the parameter level (vega/vega_interface.py:405-419, vega/utils.py:375-393), per
 * theta column j: theta[b][j] + offset[j] enters the model (offset = pi - exp(v^2) for a blinded
 * parameter, 0 otherwise), and force_one[j] != 0 pins the column to 1 whatever the caller passed
 * (the BLIND_FIXED_PARS).  Never called = identity. */
int vg_plan_set_theta_transform(vg_plan* plan, const double* offset, const int32_t* force_one);
/* The box the sampled parameters live in: lo[n_theta], hi[n_theta] per theta column, NaN where the
 * [sample] section gives no limit (vega/vega_interface.py:737-790 `sample_params['limits']`: the
 * limits iminuit and the samplers keep theta inside, vega/minimizer.py:60-66,
 * vega/samplers/polychord.py:106-127).  Optional, before vg_plan_finalize: only the
 * velocity-dispersion tables (VG_OPT_VELO_TABLES) read it. */
int vg_plan_set_theta_limits(vg_plan* plan, const double* lo, const double* hi);
/* Fvoigt profile table for model-hcd = fvoigt (vega/power_spectrum.py:360-380) */
int vg_plan_set_fvoigt(vg_plan* plan, int32_t n, const double* x, const double* f);
/* UVB shot-noise A(tau) table (vega/correlation_func.py:597-626) */
int vg_plan_set_uvshot_table(vg_plan* plan, int32_t n, const double* tau, const double* a);
/* Gaussian priors (vega/vega_interface.py:423-446) */
int vg_plan_set_priors(vg_plan* plan, int32_t n, const int32_t* cols, const double* mean,
                       const double* sigma);
/* A Hankel + spline operator set: for each of n_ell multipoles an operator
 * op[l][2*n_out][nk_pad] (rows 0..n_out: xi at the knots; rows n_out..2n_out: spline second
 * derivatives, both linear in P_l(k)) and the knots xk[l][n_out] = log r.  It restates
 * mcfit.P2xi + interp1d(cubic) (vega/pktoxi.py:141-152) or Pk2Mp + splrep
 * (vega/pktoxi.py:230-319).  bounds_error: 1 = out-of-range r flags VG_STATUS_BOUNDS
 * (interp1d), 0 = end-polynomial extrapolation (splev).  ells[l]: multipole order. */
int vg_plan_add_hankel(vg_plan* plan, int32_t n_ell, const int32_t* ells, int32_t n_out,
                       const double* xk, const double* op, int32_t bounds_error, int32_t* id);

int vg_plan_add_corr(vg_plan* plan, int32_t hankel_id, int32_t n_model, int32_t n_dist,
                     int32_t* corr_id);
/* Coordinate table of n_model bins: tab[6][n_model] = r, mu, log((1+z_1)/(1+z_eff)),
 * (1+z)^2, growth factor ratio squared, log((1+z_2)/(1+z_eff))
 * (vega/correlation_func.py:46-48,76-80,372-403).  z_1 = z_2 = z of the bin unless
 * new-bias-evolution gives the two tracers of a cross-correlation their own redshifts
 * (vega/correlation_func.py:238-274). */
int vg_corr_add_coords(vg_plan* plan, int32_t corr, const double* tab, double opz_eff2,
                       int32_t* coord_set);
int vg_corr_add_eval(vg_plan* plan, int32_t corr, const vg_eval_desc* desc, int32_t* eval_id);
/* Static spline tables of the relativistic / asymmetry multipoles of the linear P(k)
 * (vega/pktoxi.py:321-382): which = 0 rel l=1, 1 rel l=3, 2 asy l=0, 3 asy l=2; pk_lin_sel:
 * the isotropic P(k) of the component it is added to (0 peak, 1 smooth, 2 full);
 * tab[3][n] = knots (log r), values, second derivatives. */
int vg_corr_set_static_xi(vg_plan* plan, int32_t corr, int32_t which, int32_t pk_lin_sel, int32_t n,
                          const double* tab);
/* Metal matrix of metal slot c, [n_model][n_model] row-major (vega/metals.py:338-367).
 * Slots never set are identity (the reference's `test` flag, vega/data.py:683-684). */
int vg_corr_set_metal_matrix(vg_plan* plan, int32_t corr, int32_t slot, const double* mat);
/* rp-only metal matrix of a new_metals configuration (vega/metals.py:354-358, 679-752):
 * mat[n_rp][n_rp], applied to every rt column of the metal xi (n_rp * n_rt = model bins).  A
 * correlation uses either this form for all of its metal correlations or dense matrices. */
int vg_corr_set_metal_matrix_rp(vg_plan* plan, int32_t corr, int32_t slot, const double* mat,
                                int32_t n_rp, int32_t n_rt);
/* Distortion matrix [n_dist][n_model] (vega/model.py:143-144); never set = identity. */
int vg_corr_set_distortion(vg_plan* plan, int32_t corr, const double* dmat);
/* DESI instrumental systematics: amp * table[n_model] added to the smooth component
 * (vega/correlation_func.py:553-595, vega/model.py:133-135). */
int vg_corr_set_inst_sys(vg_plan* plan, int32_t corr, const double* table, int32_t idx_amp,
                         double amp_default);
/* Broadband terms.  design_pre[rows][n_model], design_post[rows][n_dist]: monomials
 * x1^i x2^j; sky_pre[2][n_model] / sky_post[2][n_dist]: rt and the 0<=rp<binsize window. */
int vg_corr_set_broadband(vg_plan* plan, int32_t corr, int32_t n_terms, const vg_bb_term* terms,
                          int32_t rows_pre, const double* design_pre, int32_t rows_post,
                          const double* design_post, const double* sky_pre, const double* sky_post);
/* Masked data of one correlation (vega/data.py:169-250,397-413): mask_idx[n_mask] are the
 * distorted-model bins kept by the cuts, data[n_mask] the masked data vector, invcov
 * [n_mask][n_mask] the inverse masked covariance (NULL = identity).  May be called again
 * to swap in a Monte Carlo mock (vega/vega_interface.py:296-312). */
int vg_corr_set_data(vg_plan* plan, int32_t corr, int32_t n_mask, const int32_t* mask_idx,
                     const double* data, const double* invcov);
/* A new data vector alone -- a Monte Carlo mock swapped in while mask and covariance stay
 * (vega/vega_interface.py:296-312 reads data.masked_mc_mock on every call): data[n_mask], same mask as
 * the last vg_corr_set_data.  Keeps the inverse covariance, its square root and everything built on them. */
int vg_corr_set_data_vector(vg_plan* plan, int32_t corr, int32_t n_mask, const double* data);
/* A square root of the inverse covariance just given to vg_corr_set_data: w[n_mask][n_mask] with
 * invcov = w^T w (e.g. the transposed lower Cholesky factor), or NULL to clear it.  Optional; with it
 * (and VG_OPT_WHITEN) the chi2 (d - m)^T C^-1 (d - m) of vega/vega_interface.py:294-319 is evaluated as
 * |w d - w m|^2 with w folded into the masked distortion rows, the fused metal operator, the data and the
 * mocks, so no product with C^-1 is left on the hot path.  Used when the correlation has a distortion
 * matrix, its own covariance (no global one) and no post-distortion broadband term; ignored otherwise.
 * A later vg_corr_set_data with a new inverse covariance drops it until it is given again. */
int vg_corr_set_whitening(vg_plan* plan, int32_t corr, int32_t n_mask, const double* w);
/* 1 when the chi2 path of the correlation runs on the whitened operands (after the first chi2 call) */
int vg_corr_whitened(const vg_plan* plan, int32_t corr);
/* Global covariance over all correlations (vega/vega_interface.py:888-954) */
int vg_plan_set_global_invcov(vg_plan* plan, int32_t n_total, const double* invcov);
/* Uploads everything, builds the static factor grids on the device. Synchronous. */
int vg_plan_finalize(vg_plan* plan, void* stream);
/* Largest number of vectors processed per pass (workspace is sized for it); default 1024 */
int vg_plan_set_chunk(vg_plan* plan, int64_t chunk);
/* Run-time options.  VG_OPT_MOMENTS (default 1): evaluations whose sampled P(k,mu)
 * parameters are only Kaiser / HCD amplitudes are computed in moment form -- their knot
 * tables are a 3/9/14-term combination of static tables built at finalize (exact algebra: the
 * mu projection of vega/pktoxi.py:79-97, the FFTLog and the spline are linear; 14 terms with the
 * scale-dependent UV / HeII bias of vega/power_spectrum.py:224-261 at fixed UV parameters,
 * vg_eval_desc::uv_static); 0 sends every evaluation through the per-point P(k,mu) pass. */
#define VG_OPT_MOMENTS 0
/* VG_OPT_DIRECT_PK (default 0): 1 makes vg_model_batch / vg_chi2_batch* evaluate
 * Model.compute_direct (vega/model.py:189-208) -- one `full` component from the isotropic
 * P(k) given to vg_plan_set_direct_pk, no peak/smooth split, no bao_amp, and (as in the
 * reference with no-metal-decomp) no metal terms. */
#define VG_OPT_DIRECT_PK 1
/* VG_OPT_FOLD_METALS (default 0): metal evaluations whose xi stage is in moment form
 * (vg_eval_desc::static_xi) are a 3/9-term combination of static vectors X_m, so their matrix
 * product M_c xi_c (vega/metals.py:350-358) is the same combination of Y_m = M_c X_m.  1 builds
 * nothing new (the Y_m are made at finalize either way) but takes those terms out of the gather
 * and out of the per-vector matrix product; same result to rounding (summation order).  Needs
 * VG_OPT_MOMENTS.  The bench headline keeps it off: every matrix applied per vector. */
#define VG_OPT_FOLD_METALS 2
/* VG_OPT_OZAKI (default -1 = automatic): the dense metal-matrix product (vega/metals.py:338-367 for
 * a batch) in split-integer form: both operands are cut into S signed radix-128 digits per
 * row-scaled entry, the S(S+1)/2 digit products run on the tcgen05 tensor cores (kind::i8, exact
 * s32 accumulation in TMEM) and are recombined in FP64 (csrc/vg_ozaki.cuh).  -1: per correlation,
 * the fewest digits whose worst-case truncation bound at the plan parameters is below 5e-11 max|xi|
 * (vg_corr_ozaki_slices); 3..8: that many digits; 0: the FP64 DMMA kernel (the parity reference).
 * Batches below VG_OPT_OZAKI_MIN_BATCH vectors take the DMMA kernel. */
#define VG_OPT_OZAKI 3
/* VG_OPT_OZAKI_DM (default 1): the distortion-matrix product (vega/model.py:142-143) takes the same
 * split-integer kernel, with the digits VG_OPT_OZAKI selects (automatic: its own bound against
 * max|D xi| at the plan parameters, vg_corr_ozaki_dm_slices); 0 keeps it on the FP64 DMMA kernel. */
#define VG_OPT_OZAKI_DM 4
/* VG_OPT_VELO_TABLES (default 1): evaluations whose only sampled non-polynomial P(k,mu) factor is ONE
 * velocity dispersion (vega/power_spectrum.py:588-636) with finite limits (vg_plan_set_theta_limits)
 * take the moment form too: their static knot tables K(sigma_v) are a Chebyshev series in
 * log(sigma_v^2 + eps^2) over [lo, hi], built at finalize from 64 per-point passes and CHECKED there
 * against the per-point pass on 8 other values of sigma_v (order count = the smallest that meets
 * 2e-12 max|xi| per output slot; none meeting it = no tables).  The hot path is then one FP64 GEMM
 * per evaluation instead of the 814 000-point pass and its Hankel rows.  Used for calls whose vectors
 * all lie inside [lo, hi] (checked per call: on the host for host theta, one small kernel and a
 * 4-byte read-back for device theta); anything else, and 0, keeps the per-point pass (the parity
 * reference of the tables). */
#define VG_OPT_VELO_TABLES 5
/* VG_OPT_FUSE_DM_METALS (default 1): on the chi2 path the metal matrices are applied as the plan-time
 * product (masked rows of the distortion matrix) x M_c -- D (M_c xi_c) = (D M_c) xi_c,
 * vega/model.py:142-143 with vega/metals.py:350-358 -- so the per-vector product has n_mask output
 * bins instead of n_model and the distortion product keeps the core terms only.  Needs dense metal
 * matrices, a distortion matrix and no multiplicative pre-distortion broadband term; the model entry
 * points (all distorted bins) keep the two products.  0: the two-step form (its parity reference). */
#define VG_OPT_FUSE_DM_METALS 6
/* VG_OPT_WHITEN (default 1): use the square root given to vg_corr_set_whitening (see there); 0 keeps the
 * product with C^-1 (the parity reference). */
#define VG_OPT_WHITEN 7
/* VG_OPT_OZAKI_MIN_BATCH (default 64, the measured crossover): chunks with fewer parameter vectors take the FP64 DMMA kernel for
 * the products VG_OPT_OZAKI / VG_OPT_OZAKI_DM name (a thin batch streams the FP64 operator once; the
 * split-integer kernel pays for all digit products of a 64-vector tile whatever the batch). */
#define VG_OPT_OZAKI_MIN_BATCH 8
/* VG_OPT_LANES (default 1): the correlations of a chunk are spread over up to four lanes (correlation
 * i on lane i % 4; lanes 1..3 are streams of the plan with their own workspace, forked from and joined
 * to the caller's stream, inside the CUDA graph of a small pass too); the chi2 of the lanes are added at
 * the end (same values: the per-correlation arithmetic does not change).  A single vector or a gradient
 * stencil -- latency-bound kernels -- gains 15-40 %, a 1024-vector chunk 2 %.  Chunks above VG_LANE_MAX_B
 * (environment, default 1024) vectors and passes under vg_profile_enable stay on one stream.  0 keeps
 * every correlation on the caller's stream. */
#define VG_OPT_LANES 9
int vg_plan_set_option(vg_plan* plan, int32_t option, int32_t value);
/* Isotropic P(k) [nk] for the direct mode (e.g. from a Boltzmann code, the `direct_pk`
 * argument of VegaInterface.compute_model / chi2 / log_lik); host pointer, copied. */
int vg_plan_set_direct_pk(vg_plan* plan, const double* pk_host);
/* Number of evaluations of a correlation served by the moment form (after finalize; 0 when
 * the option is off) -- the remaining ones go through pk_ell + the Hankel GEMM */
int vg_corr_n_moment(const vg_plan* plan, int32_t corr);
/* Evaluations of the correlation served by velocity-dispersion tables (after finalize; 0 when off or
 * when no order count met the tolerance); orders (nullable): Chebyshev orders kept, error
 * (nullable): the largest |xi(tables) - xi(per point)| / max|xi| measured at finalize. */
int vg_corr_velo_tables(const vg_plan* plan, int32_t corr, int32_t* orders, double* error);
/* 1 when the last vg_model_batch / vg_chi2_batch* call ran on the velocity-dispersion tables (every
 * vector inside the tabulated range), 0 when it took the per-point pass */
int vg_plan_velo_tables_active(const vg_plan* plan);
/* Evaluations of the correlation whose xi stage (vega/correlation_func.py:117-236) is in moment
 * form too (vg_eval_desc::static_xi): read-out of static per-bin tables, no gather. */
int vg_corr_n_static_xi(const vg_plan* plan, int32_t corr);
/* Metal evaluations VG_OPT_FOLD_METALS = 1 serves from static vectors (after finalize). */
int vg_corr_n_folded(const vg_plan* plan, int32_t corr);
/* Digits the split-integer metal GEMM of a correlation runs with under the current VG_OPT_OZAKI
 * (0: the correlation has no dense metal matrices or the option is off); bound_per_digits[9]
 * (nullable) receives, for S = 3..8, the truncation bound relative to max|xi_smooth| at the plan
 * parameters. */
int vg_corr_ozaki_slices(const vg_plan* plan, int32_t corr, double* bound_per_digits);
/* The same for the distortion-matrix product (bound relative to max|D xi|; 0: DMMA) */
int vg_corr_ozaki_dm_slices(const vg_plan* plan, int32_t corr, double* bound_per_digits);
/* The fused distortion x metal operator of a correlation (VG_OPT_FUSE_DM_METALS): returns its rows
 * (n_mask; 0 when it was not built or the option is off), digits (nullable): digits of the split-integer
 * GEMM on it (0: DMMA), bound_per_digits[9] (nullable): truncation bound relative to max|D xi|. */
int vg_corr_fused_metal_operator(const vg_plan* plan, int32_t corr, int32_t* digits, double* bound_per_digits);
/* Status bits raised while building the static grids (e.g. VG_STATUS_ARINYO): they apply
 * to every vector */
int vg_plan_static_status(const vg_plan* plan);
/* chi2 passes of at most 96 vectors are replayed as CUDA graphs (after two ordinary calls per batch
 * size; VG_GRAPHS=0 in the environment turns it off): number of replays so far */
int64_t vg_plan_graph_replays(const vg_plan* plan);
/* Bytes of device memory held by the plan: static operands, the whitened / fused operators and the digit
 * planes built so far, Monte Carlo mocks, the current workspace and its per-lane copies (VG_OPT_LANES) */
int64_t vg_plan_device_bytes(const vg_plan* plan);

/* ---------------------------------------------------------------------------------
 * Hot path.  theta_dev[B][n_theta] row-major device memory.
 * ------------------------------------------------------------------------------- */
/* compute_model for a batch: xi_out_dev[c] -> [B][n_dist_c] (vega/vega_interface.py:208-248) */
int vg_model_batch(vg_plan* plan, const double* theta_dev, int64_t B, double* const* xi_out_dev,
                   int32_t* status_dev, void* stream);
/* chi2 for a batch incl. priors (vega/vega_interface.py:250-325); status != 0 -> 1e100 */
int vg_chi2_batch(vg_plan* plan, const double* theta_dev, int64_t B, double* chi2_dev,
                  int32_t* status_dev, void* stream);
/* Monte Carlo mocks (vega/analysis.py:195-308, bin/run_vega_mc_fits_mpi.py:11-78): n_mocks masked
 * data vectors [n_mocks][n_mask] of one correlation (after vg_corr_set_data, whose invcov
 * is then the scaled one, vega/data.py:689-760).  vg_chi2_batch_mc evaluates vector b against
 * mock mock_idx_dev[b] of every correlation, so many mock fits advance in one batch. */
int vg_corr_set_mocks(vg_plan* plan, int32_t corr, int32_t n_mocks, const double* mocks);
int vg_chi2_batch_mc(vg_plan* plan, const double* theta_dev, const int32_t* mock_idx_dev, int64_t B,
                     double* chi2_dev, int32_t* status_dev, void* stream);
/* Same with HOST buffers: pinned staging + H2D of theta, D2H of chi2/status; synchronous. */
int vg_chi2_batch_host(vg_plan* plan, const double* theta_host, int64_t B, double* chi2_host,
                       int32_t* status_host, void* stream);
/* Per-stage taps for parity tests. stage: 0 P(k,mu) of one evaluation [B][nmu][nk];
 * 1 multipoles [B][n_eval][n_ell][nk_pad]; 2 knot tables [B][n_eval][n_ell][2][n_out];
 * 3 per-evaluation xi [B][n_eval][n_model]. */
int vg_stage_batch(vg_plan* plan, int32_t corr, int32_t eval, int32_t stage,
                   const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev,
                   void* stream);
int64_t vg_stage_size(const vg_plan* plan, int32_t corr, int32_t stage);
/* ---------------------------------------------------------------------------------
 * Per-stage call shapes on CALLER-SUPPLIED arrays (the reference's PowerSpectrum.compute,
 * PktoXi.compute / compute_pk_ells, CorrelationFunction.compute and Model.compute take their
 * P(k) / P(k,mu) as arguments).  Synchronous; B <= the chunk size.
 * ------------------------------------------------------------------------------- */
/* New templates for Model.compute(pars, pk_full, pk_smooth) (vega/model.py:157-187): replaces the
 * linear P(k) of the peak (pk_full - pk_smooth), smooth and full components and re-derives
 * everything vg_plan_finalize derives from them.  Host pointers [nk]. */
int vg_plan_set_templates(vg_plan* plan, const double* pk_full, const double* pk_smooth, void* stream);
/* PowerSpectrum.compute(pk_lin, params, fast_metals) (vega/power_spectrum.py:87-196) of the
 * PowerSpectrum object evaluation `eval` of the correlation stands for: out_dev[B][nmu][nk].
 * pk_lin_host [nk] (NULL: the evaluation's own linear P(k)); fast_metals 0 / 1, -1 = as configured. */
int vg_pk_batch(vg_plan* plan, int32_t corr, int32_t eval, const double* pk_lin_host, int32_t fast_metals,
                const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev, void* stream);
/* PktoXi.compute(r_grid, mu_grid, pk, single_ell) and compute_pk_ells(pk) (vega/pktoxi.py:79-163) for
 * B caller-supplied P(k,mu) grids pk_grid_dev[B][nmu][nk] through transform `hankel_id`:
 * xi_out_dev[B][n] on the coordinates r_host[n], mu_host[n] (n = 0: skipped), ells_out_dev
 * [B][n_ell][nk_pad] (nullable).  Works on a plan that is not finalized (a stand-alone PktoXi). */
int vg_pktoxi_batch(vg_plan* plan, int32_t hankel_id, const double* pk_grid_dev, int64_t B, const double* r_host,
                    const double* mu_host, int32_t n, int32_t single_ell, double* xi_out_dev, double* ells_out_dev,
                    int32_t* status_dev, void* stream);
/* CorrelationFunction.compute(pk, pk_lin, PktoXi_obj, params) (vega/correlation_func.py:117-198) of
 * the CorrelationFunction object evaluation `eval` stands for, on caller-supplied P(k,mu) grids:
 * AP rescaling, spline read-out, bias evolution, growth and the additive terms; out_dev[B][n_model].
 * pk_lin_sel: which template (0 peak, 1 smooth, 2 full) the relativistic / asymmetry tables are
 * taken for (-1: the evaluation's own). */
int vg_xi_from_grid_batch(vg_plan* plan, int32_t corr, int32_t eval, const double* pk_grid_dev, int32_t pk_lin_sel,
                          const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev, void* stream);

/* Per-stage device timing with CUDA events on the launch stream (bench.py roofline figures).
 * Tags: 0 P(k,mu)->multipoles, 1 Hankel GEMM, 2 xi gather kernels, 3 dense metal GEMM (DMMA or the
 * split-integer tcgen05 kernel, with its split-K reduction), 4 distortion GEMM, 5 C^-1 GEMM,
 * 6 elementwise stages, 7 moment-form knot tables, 8 xi stage in moment form, 9 digit planes of the
 * batch operand (split-integer GEMM), 10 rp-only metal matrices; 11..15 unused.
 * vg_profile_read synchronises the device, fills ms_per_tag[16] / count_per_tag[16] with the
 * totals since the last read and resets them. */
int vg_profile_enable(vg_plan* plan, int enable);
int vg_profile_read(vg_plan* plan, double* ms_per_tag, int64_t* count_per_tag);
/* Number of kernels launched by the library since the last call (bench bookkeeping) */
int64_t vg_launch_count(int reset);

/* FP64 FMA throughput probe (8 independent chains per thread): the measured ALU roofline
 * the fused P(k,mu) kernel is graded against.  out_dev: blocks*256 doubles of scratch. */
int vg_probe_dfma(int blocks, int iters, double* out_dev, double* tflops, void* stream);

/* new_metals metal matrices (vega/metals.py:502-752, init time): the weighted histograms over
 * all N1 x N2 pixel pairs of two stacked-weight tables that the reference forms with
 * np.histogram2d / np.histogram (:543-556, :566-584, :625-637, :716-728).  Host pointers.
 * side = [5][n] doubles: comoving distance at the TRUE absorber redshift, at the ASSUMED
 * (main-tracer) redshift, true redshift, assumed redshift, weight (scaling included).
 *   vg_metal_pair_ratio_range  range_out[2] = min, max of d_assumed / d_true over all pairs
 *   vg_metal_pair_histograms   H[nb][nb] by (assumed, true) rp bin; per_true[nb];
 *                              per_assumed[3][nb] = sum w, sum w*rp_assumed, sum w*z_true_pair;
 *                              ratio[nr] (nr = 0: skipped) = sum w / d_true^2 (|rp_true| < 20)
 * Bins are [e_i, e_{i+1}), last one closed, as numpy's; signed_rp = a discrete tracer is in. */
int vg_metal_pair_ratio_range(const double* side1, int32_t n1, const double* side2, int32_t n2,
                              double* range_out);
int vg_metal_pair_histograms(const double* side1, int32_t n1, const double* side2, int32_t n2,
                             int32_t signed_rp, double zmin, double zmax, const double* rp_edges, int32_t nb,
                             const double* ratio_edges, int32_t nr, double* H, double* per_true,
                             double* per_assumed, double* ratio);

/* Stand-alone split-integer GEMM C[m,n] = sum_k A[m,k] B[n,k] on tcgen05 kind::i8 (tests): device
 * pointers, row strides in doubles (K-contiguous operands), `slices` digits.  Synchronous.
 * *error_out != 0: a pipeline wait timed out. */
int vg_ozaki_gemm(const double* A, const double* B, double* C, int M, int N, int K, long long lda,
                  long long ldb, long long ldc, int slices, int32_t* error_out, void* stream);
/* Stand-alone FP64 tensor-core GEMM C[m,n] = alpha sum_k A[m,k] B[n,k] + beta C (tests) */
int vg_dgemm_tn(const double* A, const double* B, double* C, int M, int N, int K,
                long long lda, long long ldb, long long ldc, double alpha, double beta,
                int splits, double* workspace, void* stream);
/* Split-K FP64 tensor-core GEMM C[m,n] = alpha sum_k A[m,k] B[n,k] that allocates NOTHING: no
 * counterpart in the reference -- added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  It is vg_dgemm_tn with splits >= 2 and beta = 0, the split-K partials held in a file-scope
 * device array of VG_STATIC_SCRATCH_DOUBLES doubles (16 MiB, resident on every GPU that loads the
 * library until exit) instead of a caller workspace.  Device pointers, row strides in doubles,
 * K % 16 == 0, lda and ldb even (as vg_dgemm_tn).  VG_ERR_ARG with a message when
 * splits * M * (N rounded up to even) > VG_STATIC_SCRATCH_DOUBLES.  Asynchronous on `stream`; no
 * cudaMalloc / cudaFree / allocating library call.  Every call on a device shares the one array
 * (it keeps the last call's partials; each call overwrites all it reads, nothing needs clearing):
 * two calls that overlap on the GPU -- different streams, different threads -- race. */
#define VG_STATIC_SCRATCH_DOUBLES (1 << 21)
int vg_dgemm_tn_static(const double* A, const double* B, double* C, int M, int N, int K, long long lda,
                       long long ldb, long long ldc, double alpha, int splits, void* stream);
#ifdef __cplusplus
}
#endif
#endif
