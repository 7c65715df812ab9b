// C-ABI of the vega_b200 CUDA library (see include/vega_b200.h): plan construction and
// the kernel sequence of the batched model / chi2 evaluation.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdlib>
#include <cstring>
#include <map>
#include <memory>
#include <string>
#include <mutex>
#include <vector>

#include <nvtx3/nvToolsExt.h>

#include "../../include/vega_b200.h"
#include "vg_common.cuh"
#include "vg_gemm.cuh"
#include "vg_metalmat.cuh"
#include "vg_moments.cuh"
#include "vg_ozaki.cuh"
#include "vg_pk.cuh"
#include "vg_post.cuh"
#include "vg_xi.cuh"

static thread_local char g_error[1024] = "";
static long long g_launches = 0;

void vg_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof(g_error), fmt, ap);
  va_end(ap);
}

#define VG_REQUIRE(cond, ...)            \
  do {                                   \
    if (!(cond)) {                       \
      vg_set_error(__VA_ARGS__);         \
      return VG_ERR_ARG;                 \
    }                                    \
  } while (0)
#define VG_TRY(expr)                     \
  do {                                   \
    int rc__ = (expr);                   \
    if (rc__) return rc__;               \
  } while (0)
#define VG_LAUNCHED()                    \
  do {                                   \
    ++g_launches;                        \
    VG_CUDA_CHECK(cudaGetLastError());   \
  } while (0)

namespace {

inline int pad16(long long n) { return (int)((n + 15) / 16 * 16); }

// bumped by every device allocation: a captured CUDA graph holds raw pointers into the plan's
// buffers, so it is only replayed while no buffer has been (re)allocated since its capture
static long long g_alloc_generation = 0;

struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : p(o.p), bytes(o.bytes) { o.p = nullptr; o.bytes = 0; }
  ~DevBuf() { release(); }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  void swap(DevBuf& o) { std::swap(p, o.p); std::swap(bytes, o.bytes); }
  int alloc(size_t n, bool zero = true) {
    release();
    if (n == 0) n = 16;
    ++g_alloc_generation;
    VG_CUDA_CHECK(cudaMalloc(&p, n));
    bytes = n;
    if (zero) {
      // the fill runs on the legacy default stream: waited for here, because the buffer may be used next
      // on a non-blocking stream (the graph-capture stream, the lanes of a thin batch) that does not
      // order itself after it
      VG_CUDA_CHECK(cudaMemset(p, 0, n));
      VG_CUDA_CHECK(cudaStreamSynchronize(cudaStreamLegacy));
    }
    return VG_OK;
  }
  int ensure(size_t n) { return n <= bytes ? VG_OK : alloc(n); }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// The quadratic form r^T B r only sees the symmetric part of B; the chi2 GEMM reads one triangle
// (vg_gemm.cuh, GemmArgs::sym), so an inverse covariance that is symmetric to rounding only
// (numpy's LU inverse, Q^T C^-1 Q of the marginalisation) is replaced by (B + B^T) / 2 once.
int symmetrize(DevBuf& buf, int n, long long ld) {
  dim3 grid((unsigned)((n + 255) / 256), (unsigned)n);
  vg::symmetrize_kernel<<<grid, 256>>>(buf.as<double>(), n, ld);
  VG_CUDA_CHECK(cudaGetLastError());
  VG_CUDA_CHECK(cudaDeviceSynchronize());
  return VG_OK;
}

template <class T> int upload(DevBuf& buf, const T* host, size_t n) {
  VG_TRY(buf.alloc(n * sizeof(T)));
  VG_CUDA_CHECK(cudaMemcpy(buf.p, host, n * sizeof(T), cudaMemcpyHostToDevice));
  return VG_OK;
}

// host [rows][cols] -> device [rows_pad][ld] zero padded
int upload_padded(DevBuf& buf, const double* host, long long rows, long long cols, long long ld,
                  long long rows_alloc = -1) {
  if (rows_alloc < rows) rows_alloc = rows;
  VG_TRY(buf.alloc((size_t)rows_alloc * ld * sizeof(double)));
  VG_CUDA_CHECK(cudaMemcpy2D(buf.p, ld * sizeof(double), host, cols * sizeof(double),
                             cols * sizeof(double), rows, cudaMemcpyHostToDevice));
  return VG_OK;
}

struct Hankel {
  int n_ell = 0, n_out = 0, n_out_pad = 0, bounds_error = 1;
  int uniform = 0;   // knots log-uniform to 1e-7 of a spacing: the fast gather kernel applies
  double x0[vg::PK_MAX_ELL] = {0, 0, 0, 0}, x1[vg::PK_MAX_ELL] = {0, 0, 0, 0};
  int ells[vg::PK_MAX_ELL] = {0, 0, 0, 0};
  DevBuf xk;    // [n_ell][n_out_pad]
  DevBuf inv_h; // [n_ell][n_out_pad] reciprocal knot spacings
  DevBuf op;    // [n_ell][2*n_out_pad][nk_pad]
};

// one operand of the split-integer GEMM (vg_ozaki.cuh): S digit planes [S][rows][Kp] + row exponents
struct OzakiOperand {
  DevBuf planes, exps;
  long long rows = 0;       // rows the planes were built for (capacity for the workspace operand)
  int Kp = 0, S = 0;
  const double* src = nullptr;   // static operands: the matrix the digits were taken from
};

// pk_ell launch lists of one layout
struct PkLists {
  // evaluations grouped by pk_ell kernel class: 0..19 fast (VMODE*5 + KIND), 20..22 generic (20 + MODE)
  std::vector<int> class_ids[23];
  DevBuf d_class_ids[23];
  // pairs of evaluations that go through the fast kernels together (vg_eval_desc::pk_twin):
  // pair_ids[m][i] with twin pair_twins[m][i]; class_ids[m] then lists the pairs' members LAST
  // (entries [0, n_single[m]) are launched one by one, B < 8 launches them all one by one)
  std::vector<int> pair_ids[23], pair_twins[23];
  DevBuf d_pair_ids[23], d_pair_twins[23];
  int n_single[23] = {0};
};

struct EvalLayout {
  PkLists pk;                    // every evaluation that owns per-point tables
  PkLists pk_hot;                // the same without the ones served by velocity-dispersion tables
  std::vector<int> mom_ids;      // evaluations in moment form
  // velocity-dispersion tables (vg_moments.cuh): owners served by them when the tables are active;
  // their per-point slots are numbered LAST (n_full_hot .. n_full-1), so with the tables active the
  // per-point pass and the Hankel GEMM run on n_full_hot slots and these evaluations read rows
  // n_mom .. n_mom + n_vt - 1 of the moment-table region (d_tabslot_vt)
  std::vector<int> vt_ids, vt_nm;              // vt_nm[i] = terms of evaluation vt_ids[i] (vg::vt_terms)
  std::vector<vg::VtTerms> vt_terms;
  DevBuf d_vt_ids, d_vt_terms, d_tabslot_vt;
  int n_full_hot = 0;
  std::vector<int> gather_ids;   // evaluations K3 runs for: [0, n_gather_fast) through the fast kernel
  int n_gather_fast = 0;
  std::vector<int> mxi_ids;      // LAY_MOM: evaluations whose xi stage is in moment form (static_xi)
  std::vector<int> mxi_owner;    //          evaluation whose moment tables each of them reads
  DevBuf d_mxi_ids, d_mxi_owner;
  // LAY_FOLD: row of each mxi evaluation in LAY_MOM's d_mxi_tab; metal evaluations folded into
  // static vectors (fold_rows = their rows in d_mxi_tab)
  std::vector<int> mxi_rows, fold_ids, fold_owner, fold_rows;
  DevBuf d_mxi_rows, d_fold_ids, d_fold_owner;
  int n_full = 0;                // evaluations that go through pk_ell + the Hankel GEMM
  int n_metal = 0;               // metal column blocks of xmet / mcat
  std::vector<int> metal_map;    // metal slot -> column block (-1: not in this layout)
  DevBuf d_mom_ids, d_gather_ids, d_slot, d_tabslot, d_metal_map;
};
enum { LAY_MOM = 0, LAY_FULL = 1, LAY_ALL = 2, LAY_DIRECT = 3, LAY_FOLD = 4, LAY_COUNT = 5 };
constexpr int FOLD_MAX = 64;   // folded metal terms per correlation (shared-memory coefficient table)

struct Corr {
  int hankel_id = 0, n_model = 0, n_dist = 0, ld_model = 0, ld_dist = 0;
  std::vector<vg_eval_desc> evals;
  DevBuf d_evals;
  // Three ways of laying the evaluations out over the kernels of the front end:
  //   LAY_MOM  hot path, moment form on: per-point pass for the evaluations that need it,
  //            moment form for the rest;   LAY_FULL  hot path, every evaluation per point;
  //   LAY_ALL  per-stage taps and the finalize pass of the frozen terms: every evaluation
  //            incl. the frozen ones, per point, identity slots;
  //   LAY_DIRECT  Model.compute_direct: only the evaluations flagged `direct`, per point.
  //   LAY_FOLD  LAY_MOM with VG_OPT_FOLD_METALS: metal evaluations whose xi stage is in moment
  //            form leave the gather and the matrix product (static vectors Y = M_c X instead).
  // Evaluations with table_owner >= 0 never own tables; frozen ones leave the hot layouts.
  EvalLayout lay[LAY_COUNT];
  DevBuf d_mom_ktab;             // LAY_MOM: [n_mom][9][n_ell][2*n_out_pad]
  DevBuf d_mxi_tab;              // LAY_MOM: [n_mxi][9][ld_model] xi stage applied to the moment tables
  // velocity-dispersion tables (vg_moments.cuh): vt_flag[e] marks the owners served by them, all
  // with the same sampled dispersion (full-vector column vt_col = theta column vt_tcol, limits in
  // vt_map); vt_tab[i] = Chebyshev coefficients of evaluation LAY_MOM.vt_ids[i], K-major
  // [row_len][VT_NODES * nm]; vt_n = Chebyshev orders kept (0: off), vt_err = error measured at finalize
  std::vector<char> vt_flag;
  int vt_col = -1, vt_tcol = -1, vt_n = 0;
  vg::VtMap vt_map{};
  double vt_err = 0.0;
  std::vector<DevBuf> vt_tab;
  bool allow_mxi = true;         // cleared when the plan-time gather of the static tables fails
  std::vector<DevBuf> coords;
  std::vector<double> opz_eff2;
  DevBuf d_coord_ptrs, d_opz_eff2;
  DevBuf d_evolfac;              // [n_eval][ld_model] static bias-evolution x growth rows (vg_xi.cuh)
  int n_metal = 0;               // metal slots declared by the evaluations (all layouts)
  bool metal_dense = false;
  std::vector<std::vector<double>> metal_host;
  // rp-only metal matrices (new_metals, vega/metals.py:354-358): [rp_np][rp_np] per slot
  bool metal_rp = false;
  int rp_np = 0, rp_nt = 0;
  std::vector<std::vector<double>> metal_rp_host;
  DevBuf d_mrp;                  // hot ones, [n_metal_hot][rp_np][rp_np] in column-block order
  std::vector<int> frozen_ids;   // frozen evaluations, in order of d_frozen rows
  DevBuf d_frozen_ids, d_frozen; // d_frozen [n_frozen][ld_model] = weight_const * M_c xi_c
  DevBuf d_mcat;                 // hot metal matrices [n_model][n_metal_hot*ld_model]
  DevBuf d_fold_tab;             // LAY_FOLD: [n_fold][9][ld_model] Y[f][m] = M_c X[row(f)][m]
  DevBuf d_mcat_fold, d_mrp_fold;   // LAY_FOLD: matrices of the metal terms that stay in the product
  OzakiOperand oz_mcat, oz_mcat_fold;   // digit planes of d_mcat / d_mcat_fold (built on first use)
  int oz_auto_slices = 0;        // digits that meet the truncation budget at the plan parameters (0: not dense)
  double oz_bound[vg::OZ_MAX_S + 1] = {0};   // the bound per digit count, relative to max|xi| (diagnostics)
  bool has_dmat = false;
  OzakiOperand oz_dmat, oz_dmat_masked;  // digit planes of d_dmat / d_dmat_masked (built on first use)
  int oz_dm_slices = 0;          // digits of the distortion product that meet the budget (0: DMMA)
  double oz_dm_bound[vg::OZ_MAX_S + 1] = {0};
  // Whitened chi2 path (VG_OPT_WHITEN): with C^-1 = W^T W the chi2 is |W (d - y)|^2, and W folds into
  // everything linear that produces y on the masked bins: d_wd = W x (masked rows of D), the fused
  // operator is built on d_wd, d_wdata = W d, d_wmocks = mocks W^T.  The chi2 GEMM disappears.
  bool has_white = false;        // a square root of the inverse covariance was given (vg_corr_set_whitening)
  bool white_dirty = false;      // the whitened operands are out of date (rebuilt at the next chi2 call)
  bool wdata_dirty = false;      // only W d / the whitened mocks are out of date
  bool whitened = false;         // the residual path runs on the whitened operands
  DevBuf d_white, d_wd, d_wdata, d_wmocks;
  OzakiOperand oz_wd;
  int oz_wd_slices = 0;
  double oz_wd_bound[vg::OZ_MAX_S + 1] = {0};
  // Fused operator of the chi2 path (VG_OPT_FUSE_DM_METALS): the masked rows of the distortion matrix
  // times the metal matrices, d_dmcat[r][c*ld_model + j] = sum_i D[mask r][i] M_c[i][j], so that the
  // metal product lands on the n_mask distorted bins directly (no pre-distortion multiplicative term)
  DevBuf d_dmcat;
  OzakiOperand oz_dmcat;
  int oz_fused_slices = 0;       // digits that meet the budget against max|D xi| (0: DMMA)
  double oz_fused_bound[vg::OZ_MAX_S + 1] = {0};
  DevBuf d_dmat;                 // [n_dist][ld_model]
  DevBuf d_dmat_masked;          // [n_mask][ld_model]
  bool has_inst = false;
  DevBuf d_inst; int idx_inst_amp = -1; double inst_amp_default = 0.0;
  std::vector<vg_bb_term> bb_terms;
  DevBuf d_bb_terms, d_design_pre, d_design_post, d_sky_pre, d_sky_post;
  int rows_pre = 0, rows_post = 0;
  int n_mask = 0, ld_mask = 0;
  bool has_invcov = false;
  std::vector<int32_t> mask_host;
  DevBuf d_mask_idx, d_data, d_invcov;   // invcov [n_mask][ld_mask]
  DevBuf d_mocks; int n_mocks = 0;       // Monte Carlo mocks [n_mocks][n_mask]
  int sx_n[12] = {0};
  DevBuf d_sx[12];
};

}  // namespace

struct vg_plan {
  int npf = 0, n_theta = 0, nk = 0, nk_pad = 0, nmu = 0, n_ell = 0, idx_bao_amp = -1;
  std::vector<double> base;
  DevBuf d_theta_offset, d_theta_force;   // blinding transform of the theta columns (optional)
  bool has_theta_transform = false;
  DevBuf d_base, d_theta_cols, d_k, d_pk[4], d_pk_fid, d_mu, d_wleg;
  DevBuf d_fv_x, d_fv_f; int fv_n = 0;
  DevBuf d_uv_tau, d_uv_a; int uv_n = 0;
  int n_prior = 0;
  DevBuf d_prior_cols, d_prior_mean, d_prior_sigma;
  std::vector<std::unique_ptr<Hankel>> hankels;
  std::vector<std::unique_ptr<Corr>> corrs;
  std::vector<DevBuf> grids;
  DevBuf d_grid_ptrs;
  bool finalized = false;
  bool use_moments = true;
  bool direct_mode = false;      // VG_OPT_DIRECT_PK
  bool fold_metals = false;      // VG_OPT_FOLD_METALS
  bool fuse_dm_metals = true;    // VG_OPT_FUSE_DM_METALS
  bool whiten = true;            // VG_OPT_WHITEN
  bool velo_tables = true;       // VG_OPT_VELO_TABLES
  bool vt_now = false;           // this call's vectors lie inside the tabulated range: tables active
  std::vector<double> theta_lo, theta_hi;   // [sample] limits per theta column (NaN: unbounded)
  std::vector<int> theta_cols_host;
  DevBuf d_vt_flag, w_vtw;       // range-check flag; weights [n_vt][chunk][ldw]
  bool ozaki_dm = true;          // VG_OPT_OZAKI_DM: the distortion product on the split-integer kernel too
  int ozaki_min_batch = 64;      // VG_OPT_OZAKI_MIN_BATCH: smaller chunks take the FP64 DMMA kernel (measured
                                 // crossover on C2 and C3: profiles/r02_small_batches.json)
  int ozaki_slices = -1;         // VG_OPT_OZAKI: digits of the split-integer metal GEMM (0: DMMA, -1: per
                                 // correlation, the fewest digits that meet the truncation budget)
  OzakiOperand oz_A;             // workspace: digit planes of the batch operand
  DevBuf d_oz_error;             // set by a pipeline time-out inside ozaki_gemm_kernel
  int static_bad = 0;
  int std_ells = 0;
  // global covariance
  bool has_global = false; int n_total = 0, ld_total = 0;
  DevBuf d_global_invcov;
  // workspace (sized for `chunk` vectors)
  long long chunk = 0, chunk_max = 1024;
  DevBuf w_pfull, w_pkell, w_tables, w_core, w_xmet, w_met, w_x, w_y, w_diff, w_partial, w_split, w_fold,
      w_chi2, w_status, w_theta, w_mety;
  void* pinned = nullptr; size_t pinned_bytes = 0;
  // CUDA graphs of the chi2 pass for small batches (B <= GRAPH_MAX_B): the ~50-100 launches of a
  // pass are replayed as one graph on fixed staging buffers; keyed on (B, options); dropped
  // whenever the plan's state changes
  struct GraphEntry { cudaGraphExec_t exec = nullptr; int seen = 0; long long launches = 0, generation = -1; };
  std::map<long long, GraphEntry> graphs;
  cudaStream_t gstream = nullptr;
  DevBuf w_mockidx;
  long long graph_replays = 0;
  void drop_graphs() {
    for (auto& kv : graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    graphs.clear();
  }
  // Extra lanes for thin batches (chi2_chunk): the kernels of a single vector or of a gradient stencil
  // are latency bound, so correlation ci runs on lane ci % (1 + EXTRA_LANES); lanes 1.. have their own
  // stream and their own copy of the per-correlation workspace (swapped into the w_* members while
  // their launches are issued)
  static constexpr int EXTRA_LANES = 3;
  struct Lane {
    DevBuf w_pkell, w_tables, w_core, w_xmet, w_met, w_x, w_y, w_diff, w_partial, w_split, w_fold, w_mety, w_vtw;
    OzakiOperand oz_A;
    DevBuf chi2;
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    long long chunk = 0;
  } lanes[EXTRA_LANES];
  bool two_lanes = true;         // VG_OPT_LANES (VG_LANES=0 in the environment turns the second lane off for good)
  // optional per-stage timing with CUDA events (bench.py roofline figures)
  bool profile = false;
  struct ProfRec { int tag; cudaEvent_t e0, e1; };
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;

  vg::PkGridDev grid_dev() const {
    vg::PkGridDev g{};
    g.nk = nk; g.nk_pad = nk_pad; g.nmu = nmu; g.n_ell = n_ell;
    g.std_ells = std_ells;
    g.k = d_k.as<double>();
    for (int i = 0; i < 4; ++i) g.pk_lin[i] = d_pk[i].as<double>();
    g.pk_fid = d_pk_fid.as<double>();
    g.mu = d_mu.as<double>();
    g.wleg = d_wleg.as<double>();
    g.fv_x = d_fv_x.as<double>(); g.fv_f = d_fv_f.as<double>(); g.fv_n = fv_n;
    g.grids = d_grid_ptrs.as<const double*>();
    return g;
  }
};

namespace {
// Brackets a stage with two events on the launch stream when profiling is on.
// VG_NVTX=1 in the environment wraps every stage in an NVTX range named after its tag (timelines
// of nsys / ncu --nvtx); off by default (a range costs a library call per stage)
bool nvtx_enabled() {
  static const bool on = [] { const char* e = std::getenv("VG_NVTX"); return e && e[0] == '1'; }();
  return on;
}
const char* const PROF_NAMES[16] = {"vg:pk_ell", "vg:hankel_gemm", "vg:xi_gather", "vg:metal_gemm", "vg:dm_gemm",
                                    "vg:chi2_gemm", "vg:elementwise", "vg:moment_tables", "vg:moment_xi",
                                    "vg:ozaki_slice", "vg:metal_rp", "vg:velo_tables", "vg:12", "vg:13", "vg:14", "vg:15"};
struct ProfScope {
  vg_plan* plan; cudaStream_t st; cudaEvent_t e1 = nullptr; bool range = false;
  ProfScope(vg_plan* p, int tag, cudaStream_t s) : plan(p), st(s) {
    if (tag < 0) return;
    if (nvtx_enabled()) { nvtxRangePushA(PROF_NAMES[tag & 15]); range = true; }
    if (!plan->profile) return;
    cudaEvent_t ev[2];
    for (auto& e : ev) {
      if (!plan->prof_pool.empty()) { e = plan->prof_pool.back(); plan->prof_pool.pop_back(); }
      else cudaEventCreate(&e);
    }
    cudaEventRecord(ev[0], st);
    e1 = ev[1];
    plan->prof_recs.push_back({tag, ev[0], ev[1]});
  }
  ~ProfScope() {
    if (e1) cudaEventRecord(e1, st);
    if (range) nvtxRangePop();
  }
};
int gemm_auto(vg_plan* plan, const double* A, long long lda, const double* B, long long ldb,
              double* C, long long ldc, int M, int N, int K, int batch, long long strideA,
              long long strideB, long long strideC, cudaStream_t st);
int build_metal_operands(Corr& c, const EvalLayout& L, DevBuf& mcat, DevBuf& mrp);
int freeze_metal_terms(vg_plan* plan, Corr& c, cudaStream_t st);
int ozaki_budget(vg_plan* plan, Corr& c, cudaStream_t st);
int ozaki_budget_dm(vg_plan* plan, Corr& c, cudaStream_t st);
int build_fused_metal_operator(vg_plan* plan, Corr& c, cudaStream_t st);
int rebuild_whitened(vg_plan* plan, Corr& c, cudaStream_t st);
int rewhiten_data(vg_plan* plan, Corr& c, cudaStream_t st);
void flag_velo_table_candidates(vg_plan* plan, Corr& c);
int build_velo_tables(vg_plan* plan, Corr& c, cudaStream_t st);
int launch_metal_rp(Corr& c, const double* xmet, int n_metal, const double* A, long long B, double* met,
                    cudaStream_t st);
enum { PROF_PK = 0, PROF_HANKEL = 1, PROF_XI = 2, PROF_METAL = 3, PROF_DM = 4, PROF_CHI2 = 5, PROF_ELEM = 6, PROF_MOMENT = 7,
       PROF_MXI = 8, PROF_OZ_SLICE = 9, PROF_METAL_RP = 10, PROF_VTAB = 11, PROF_NTAGS = 16, PROF_NONE = -1 };
}  // namespace

namespace {
constexpr int PK_NCLS = 23;   // 0..19 fast kernels (VMODE*5 + KIND), 20..22 generic (20 + MODE)

// pk_ell kernel class of one evaluation
int pk_class(const vg_plan* plan, const vg_eval_desc& d) {
  const uint32_t m = d.dyn_mask;
  const int mode = m == 0 ? vg::PK_MODE_STATIC : (m == VG_F_VELO ? vg::PK_MODE_VELO : vg::PK_MODE_GENERIC);
  const bool hcd = d.hcd_model && (d.hcd_on[0] || d.hcd_on[1]);
  int kind = -1;
  if (!hcd) kind = d.same_tracer ? 0 : 1;
  else if (d.hcd_grid >= 0) {
    if (d.same_tracer) kind = d.fast_metals ? -1 : 2;
    else if (d.hcd_on[0] != d.hcd_on[1]) kind = d.fast_metals ? 4 : 3;
  }
  int vmode = vg::PK_V_NONE;
  if (mode == vg::PK_MODE_VELO) {
    const int n_velo = (d.idx_velo[0] >= 0) + (d.idx_velo[1] >= 0);
    vmode = d.velo_kind == 1 ? vg::PK_V_GAUSS : (n_velo == 2 ? vg::PK_V_LOR2 : vg::PK_V_LOR1);
  }
  const bool fast = mode != vg::PK_MODE_GENERIC && kind >= 0 && plan->n_ell == vg::PK_MAX_ELL;
  return fast ? vmode * 5 + kind : 20 + mode;
}

// The three evaluation layouts of a correlation (see Corr::lay)
constexpr int NUM_SMS = 148;

// K3 launch (hot path, per-stage taps, and the plan-time pass over the static moment tables)
struct GatherJob {
  const int* ids; int n_ids;            // evaluations gathered (device list)
  const int* tab_slot; const int* metal_map; int n_metal;
  const double* pfull; long long B;
  const double* tables; int n_full;
  const double* tables_mom; int n_mom; long long mom_stride_b, mom_stride_j;
  int unit_bias;
  double* core; double* xmet; double* flat_out;
  int* status;
  int fast;                             // the listed evaluations all qualify for xi_gather_fast_kernel
};
// VG_XI_FAST=0 in the environment sends every evaluation through the generic gather kernel
// (comparison runs; the parity tests cover both)
bool fast_gather_enabled() {
  static const bool on = [] { const char* e = std::getenv("VG_XI_FAST"); return !(e && e[0] == '0'); }();
  return on;
}

int launch_xi_gather(vg_plan* plan, Corr& c, const GatherJob& job, cudaStream_t st) {
  const Hankel& h = *plan->hankels[c.hankel_id];
  const int E = (int)c.evals.size();
  const int rowT = 2 * h.n_out_pad;
  const long long Bc = job.B;
  const int n_gather = job.n_ids;
  vg::XiArgs a{};
  a.h.n_ell = h.n_ell; a.h.n_out = h.n_out; a.h.n_out_pad = h.n_out_pad; a.h.bounds_error = h.bounds_error;
  a.h.uniform = h.uniform;
  a.evolfac = c.d_evolfac.p ? c.d_evolfac.as<double>() : nullptr;
  for (int l = 0; l < vg::PK_MAX_ELL; ++l) {
    a.h.ells[l] = h.ells[l];
    a.h.x0[l] = h.x0[l]; a.h.x1[l] = h.x1[l];
    a.h.inv_dx[l] = l < h.n_ell ? (double)(h.n_out - 1) / (h.x1[l] - h.x0[l]) : 0.0;
  }
  a.h.xk = h.xk.as<double>();
  a.h.inv_h = h.inv_h.as<double>();
  a.evals = c.d_evals.as<vg_eval_desc>(); a.n_eval = E;
  a.pfull = job.pfull; a.npf = plan->npf; a.B = (int)Bc;
  a.tables = job.tables; a.n_full = job.n_full;
  a.tables_mom = job.tables_mom; a.n_mom = job.n_mom;
  a.mom_stride_b = job.mom_stride_b; a.mom_stride_j = job.mom_stride_j;
  a.tab_slot = job.tab_slot;
  a.gather_ids = job.ids;
  a.metal_map = job.metal_map;
  a.unit_bias = job.unit_bias;
  a.coords = c.d_coord_ptrs.as<const double*>();
  a.opz_eff2 = c.d_opz_eff2.as<double>();
  a.n_model = c.n_model;
  for (int i = 0; i < 12; ++i) { a.sx.n[i] = c.sx_n[i]; a.sx.tab[i] = c.d_sx[i].as<double>(); }
  a.uv_tau = plan->d_uv_tau.as<double>(); a.uv_a = plan->d_uv_a.as<double>(); a.uv_n = plan->uv_n;
  a.core = job.core; a.xmet = job.xmet; a.flat_out = job.flat_out;
  a.ld_model = c.ld_model; a.n_metal = std::max(job.n_metal, 1);
  a.status = job.status;
  long long ctas_wanted = 4LL * NUM_SMS;
  int chunks = (int)std::max<long long>(1, std::min<long long>(vg_cdiv(c.ld_model, 256),
                                                           (ctas_wanted + Bc * n_gather - 1) / (Bc * n_gather)));
  a.bins_per_cta = vg_cdiv(c.ld_model, chunks);
  const size_t smem = (size_t)h.n_ell * rowT * 8;
  static size_t smem_set = 0;
  if (smem > smem_set) {
    VG_CUDA_CHECK(cudaFuncSetAttribute(vg::xi_gather_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    VG_CUDA_CHECK(cudaFuncSetAttribute(vg::xi_gather_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  dim3 grid(vg_cdiv(c.ld_model, a.bins_per_cta), n_gather, (unsigned)Bc);
  if (job.fast) vg::xi_gather_fast_kernel<<<grid, 256, smem, st>>>(a);
  else vg::xi_gather_kernel<<<grid, 256, smem, st>>>(a);
  VG_LAUNCHED();
  return VG_OK;
}

int build_layouts(vg_plan* plan, Corr& c) {
  const int E = (int)c.evals.size();
  for (int e = 0; e < E; ++e) {
    const vg_eval_desc& d = c.evals[e];
    if (d.table_owner >= 0) {
      VG_REQUIRE(d.table_owner < E && c.evals[d.table_owner].table_owner < 0 && !c.evals[d.table_owner].frozen,
                 "evaluation %d: bad table_owner %d", e, d.table_owner);
      VG_REQUIRE(!d.frozen, "evaluation %d: frozen and table_owner are exclusive", e);
    }
    if (d.frozen) VG_REQUIRE(d.out_slot >= 2, "evaluation %d: only metal terms can be frozen", e);
    if (d.direct) VG_REQUIRE(d.out_slot == 1 && d.table_owner < 0 && !d.frozen,
                             "evaluation %d: a direct evaluation writes the smooth slot", e);
  }
  c.frozen_ids.clear();
  for (int e = 0; e < E; ++e)
    if (c.evals[e].frozen && !c.evals[e].direct) c.frozen_ids.push_back(e);
  if (!c.frozen_ids.empty()) VG_TRY(upload(c.d_frozen_ids, c.frozen_ids.data(), c.frozen_ids.size()));
  if ((int)c.vt_flag.size() != E) c.vt_flag.assign(E, 0);
  for (int li = 0; li < LAY_COUNT; ++li) {
    EvalLayout& L = c.lay[li];
    const bool momlike = li == LAY_MOM || li == LAY_FOLD;
    for (PkLists* P : {&L.pk, &L.pk_hot}) {
      for (auto& ids : P->class_ids) ids.clear();
      for (int m = 0; m < 23; ++m) { P->pair_ids[m].clear(); P->pair_twins[m].clear(); P->n_single[m] = 0; }
    }
    L.mom_ids.clear(); L.gather_ids.clear(); L.mxi_ids.clear(); L.mxi_owner.clear();
    L.mxi_rows.clear(); L.fold_ids.clear(); L.fold_owner.clear(); L.fold_rows.clear();
    L.vt_ids.clear(); L.vt_nm.clear(); L.vt_terms.clear();
    L.n_full = 0; L.n_full_hot = 0; L.n_metal = 0;
    L.metal_map.assign(std::max(c.n_metal, 1), -1);
    std::vector<int> slot(E, 0), tabslot(E, 0);
    std::vector<int> full_ids;     // owners of per-point tables, in evaluation order
    for (int e = 0; e < E; ++e) {
      const vg_eval_desc& d = c.evals[e];
      const bool hot = li == LAY_DIRECT ? d.direct != 0 : (!d.direct && (li == LAY_ALL || !d.frozen));
      if (!hot) continue;
      const int owner = d.table_owner >= 0 ? d.table_owner : e;
      bool folded = false;
      if (momlike && c.allow_mxi && d.static_xi && vg::moment_eligible(c.evals[owner])) {
        int row = -1;
        if (li == LAY_FOLD) {
          const std::vector<int>& base_ids = c.lay[LAY_MOM].mxi_ids;
          row = (int)(std::find(base_ids.begin(), base_ids.end(), e) - base_ids.begin());
        }
        if (li == LAY_FOLD && d.out_slot >= 2 && (int)L.fold_ids.size() < FOLD_MAX) {
          folded = true;
          L.fold_ids.push_back(e); L.fold_owner.push_back(owner); L.fold_rows.push_back(row);
        } else {
          L.mxi_ids.push_back(e);
          L.mxi_owner.push_back(owner);
          if (li == LAY_FOLD) L.mxi_rows.push_back(row);
        }
      } else {
        L.gather_ids.push_back(e);
      }
      if (d.out_slot >= 2 && !folded) L.metal_map[d.out_slot - 2] = L.n_metal++;
      if (d.table_owner >= 0) continue;
      if (momlike && vg::moment_eligible(d)) {
        tabslot[e] = -((int)L.mom_ids.size() + 1);
        L.mom_ids.push_back(e);
      } else {
        full_ids.push_back(e);
      }
    }
    // per-point slots: the evaluations served by velocity-dispersion tables last
    for (int pass = 0; pass < 2; ++pass)
      for (int e : full_ids) {
        const bool vt = momlike && c.vt_flag[e];
        if (vt != (pass == 1)) continue;
        slot[e] = tabslot[e] = L.n_full++;
        L.pk.class_ids[pk_class(plan, c.evals[e])].push_back(e);
        if (!vt) {
          L.n_full_hot = L.n_full;
          L.pk_hot.class_ids[pk_class(plan, c.evals[e])].push_back(e);
        } else {
          L.vt_ids.push_back(e);
          L.vt_terms.push_back(vg::vt_terms(c.evals[e]));
          L.vt_nm.push_back(L.vt_terms.back().n);
        }
      }
    std::vector<int> tabslot_vt = tabslot;
    for (size_t i = 0; i < L.vt_ids.size(); ++i) tabslot_vt[L.vt_ids[i]] = -((int)(L.mom_ids.size() + i) + 1);
    for (int e = 0; e < E; ++e)
      if (c.evals[e].table_owner >= 0) {
        tabslot[e] = tabslot[c.evals[e].table_owner];
        tabslot_vt[e] = tabslot_vt[c.evals[e].table_owner];
      }
    {   // evaluations the fast gather kernel serves first (vg_xi.cuh: xi_gather_fast_kernel)
      const Hankel& hk = *plan->hankels[c.hankel_id];
      auto fast = [&](int e) {
        const vg_eval_desc& d = c.evals[e];
        return hk.uniform && hk.n_ell == 4 && plan->std_ells && d.single_ell < 0 && !d.rel && !d.asy &&
               !d.uv_shot && d.zevol_kind[0] == 0 && d.zevol_kind[1] == 0;
      };
      std::stable_partition(L.gather_ids.begin(), L.gather_ids.end(), fast);
      L.n_gather_fast = (int)std::count_if(L.gather_ids.begin(), L.gather_ids.end(), fast);
    }
    // twins: both members per point in the same fast class of this layout
    for (PkLists* P : {&L.pk, &L.pk_hot}) {
      for (int m = 0; m < 20; ++m) {
        std::vector<int>& ids = P->class_ids[m];
        std::vector<int> single, paired;
        std::vector<char> taken(E, 0);
        for (int e : ids) {
          const int t = c.evals[e].pk_twin;
          if (t >= 0 && t < E && t != e && !taken[t] && !taken[e] &&
              std::find(ids.begin(), ids.end(), t) != ids.end()) {
            P->pair_ids[m].push_back(t); P->pair_twins[m].push_back(e);
            taken[t] = taken[e] = 1;
          }
        }
        for (int e : ids) (taken[e] ? paired : single).push_back(e);
        P->n_single[m] = (int)single.size();
        ids = single;
        ids.insert(ids.end(), paired.begin(), paired.end());
      }
      for (int m = 20; m < PK_NCLS; ++m) P->n_single[m] = (int)P->class_ids[m].size();
      for (int m = 0; m < PK_NCLS; ++m) {
        if (!P->class_ids[m].empty()) VG_TRY(upload(P->d_class_ids[m], P->class_ids[m].data(), P->class_ids[m].size()));
        if (!P->pair_ids[m].empty()) {
          VG_TRY(upload(P->d_pair_ids[m], P->pair_ids[m].data(), P->pair_ids[m].size()));
          VG_TRY(upload(P->d_pair_twins[m], P->pair_twins[m].data(), P->pair_twins[m].size()));
        }
      }
    }
    if (!L.vt_ids.empty()) {
      VG_TRY(upload(L.d_vt_ids, L.vt_ids.data(), L.vt_ids.size()));
      VG_TRY(upload(L.d_vt_terms, L.vt_terms.data(), L.vt_terms.size()));
    }
    VG_TRY(upload(L.d_tabslot_vt, tabslot_vt.data(), tabslot_vt.size()));
    if (!L.mom_ids.empty()) VG_TRY(upload(L.d_mom_ids, L.mom_ids.data(), L.mom_ids.size()));
    if (!L.gather_ids.empty()) VG_TRY(upload(L.d_gather_ids, L.gather_ids.data(), L.gather_ids.size()));
    if (!L.mxi_ids.empty()) {
      VG_TRY(upload(L.d_mxi_ids, L.mxi_ids.data(), L.mxi_ids.size()));
      VG_TRY(upload(L.d_mxi_owner, L.mxi_owner.data(), L.mxi_owner.size()));
      if (!L.mxi_rows.empty()) VG_TRY(upload(L.d_mxi_rows, L.mxi_rows.data(), L.mxi_rows.size()));
    }
    if (!L.fold_ids.empty()) {
      VG_TRY(upload(L.d_fold_ids, L.fold_ids.data(), L.fold_ids.size()));
      VG_TRY(upload(L.d_fold_owner, L.fold_owner.data(), L.fold_owner.size()));
    }
    VG_TRY(upload(L.d_slot, slot.data(), slot.size()));
    VG_TRY(upload(L.d_tabslot, tabslot.data(), tabslot.size()));
    VG_TRY(upload(L.d_metal_map, L.metal_map.data(), L.metal_map.size()));
  }
  return VG_OK;
}
}  // namespace

extern "C" {

const char* vg_last_error(void) { return g_error; }
int vg_version(void) { return 100; }
int vg_sizeof_eval_desc(void) { return (int)sizeof(vg_eval_desc); }
int vg_sizeof_bb_term(void) { return (int)sizeof(vg_bb_term); }
int64_t vg_launch_count(int reset) {
  long long v = g_launches + vg::gemm_launches();
  if (reset) { g_launches = 0; vg::gemm_launches() = 0; }
  return v;
}

int vg_plan_create(int32_t n_params, const double* base, int32_t n_theta, const int32_t* theta_cols,
                   int32_t nk, const double* k, const double* pk_full, const double* pk_smooth,
                   const double* pk_fid, int32_t nmu, const double* mu, int32_t n_ell,
                   const double* wleg, int32_t idx_bao_amp, vg_plan** out) {
  VG_REQUIRE(n_params > 0 && nk > 0 && nmu > 0 && n_ell > 0 && n_ell <= vg::PK_MAX_ELL && out,
             "vg_plan_create: bad sizes");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    vg_set_error("vg_plan_create: no CUDA device (there is no CPU fallback)");
    return VG_ERR_CUDA;
  }
  std::unique_ptr<vg_plan> p(new vg_plan);
  p->npf = n_params; p->n_theta = n_theta; p->nk = nk; p->nk_pad = pad16(nk);
  p->nmu = nmu; p->n_ell = n_ell; p->idx_bao_amp = idx_bao_amp;
  p->base.assign(base, base + n_params);
  for (int j = 0; j < n_theta; ++j)
    VG_REQUIRE(theta_cols[j] >= 0 && theta_cols[j] < n_params, "theta column out of range");
  VG_TRY(upload(p->d_base, base, n_params));
  VG_TRY(upload(p->d_theta_cols, theta_cols, std::max(n_theta, 1)));
  p->theta_cols_host.assign(theta_cols, theta_cols + n_theta);
  VG_TRY(upload(p->d_k, k, nk));
  std::vector<double> peak(nk);
  for (int i = 0; i < nk; ++i) peak[i] = pk_full[i] - pk_smooth[i];   // vega/model.py:177
  VG_TRY(upload(p->d_pk[0], peak.data(), nk));
  VG_TRY(upload(p->d_pk[1], pk_smooth, nk));
  VG_TRY(upload(p->d_pk[2], pk_full, nk));
  VG_TRY(upload(p->d_pk[3], pk_full, nk));       // direct P(k): the template until set
  VG_TRY(upload(p->d_pk_fid, pk_fid, nk));
  VG_TRY(upload(p->d_mu, mu, nmu));
  VG_TRY(upload(p->d_wleg, wleg, (size_t)n_ell * nmu));
  *out = p.release();
  return VG_OK;
}

int vg_plan_destroy(vg_plan* plan) {
  if (!plan) return VG_OK;
  plan->drop_graphs();
  if (plan->gstream) cudaStreamDestroy(plan->gstream);
  for (auto& l : plan->lanes) {
    if (l.stream) cudaStreamDestroy(l.stream);
    if (l.fork) cudaEventDestroy(l.fork);
    if (l.join) cudaEventDestroy(l.join);
  }
  if (plan->pinned) cudaFreeHost(plan->pinned);
  delete plan;
  return VG_OK;
}

int vg_plan_set_chunk(vg_plan* plan, int64_t chunk) {
  VG_REQUIRE(plan && chunk > 0, "bad chunk");
  plan->drop_graphs();
  plan->chunk_max = chunk;
  return VG_OK;
}

int vg_plan_set_option(vg_plan* plan, int32_t option, int32_t value) {
  VG_REQUIRE(plan, "null plan");
  plan->drop_graphs();
  if (option == VG_OPT_MOMENTS) { plan->use_moments = value != 0; return VG_OK; }
  if (option == VG_OPT_FOLD_METALS) { plan->fold_metals = value != 0; return VG_OK; }
  if (option == VG_OPT_OZAKI) {
    VG_REQUIRE(value == 0 || value == -1 || (value >= 3 && value <= vg::OZ_MAX_S),
               "VG_OPT_OZAKI: 0 (off), -1 (automatic) or 3..%d digits", vg::OZ_MAX_S);
    plan->ozaki_slices = value;
    return VG_OK;
  }
  if (option == VG_OPT_OZAKI_DM) { plan->ozaki_dm = value != 0; return VG_OK; }
  if (option == VG_OPT_LANES) { plan->two_lanes = value != 0; return VG_OK; }
  if (option == VG_OPT_OZAKI_MIN_BATCH) {
    VG_REQUIRE(value >= 1, "VG_OPT_OZAKI_MIN_BATCH: a batch size >= 1");
    plan->ozaki_min_batch = value;
    return VG_OK;
  }
  if (option == VG_OPT_VELO_TABLES) { plan->velo_tables = value != 0; return VG_OK; }
  if (option == VG_OPT_FUSE_DM_METALS) { plan->fuse_dm_metals = value != 0; return VG_OK; }
  if (option == VG_OPT_WHITEN) {
    plan->whiten = value != 0;
    for (auto& cp : plan->corrs) cp->white_dirty = true;
    return VG_OK;
  }
  if (option == VG_OPT_DIRECT_PK) {
    if (value) {
      bool any = false;
      for (auto& cp : plan->corrs)
        for (auto& d : cp->evals) any = any || d.direct;
      VG_REQUIRE(any, "VG_OPT_DIRECT_PK: the plan holds no `direct` evaluation");
    }
    plan->direct_mode = value != 0;
    return VG_OK;
  }
  vg_set_error("vg_plan_set_option: unknown option %d", option);
  return VG_ERR_ARG;
}

int vg_plan_set_direct_pk(vg_plan* plan, const double* pk_host) {
  VG_REQUIRE(plan && pk_host, "vg_plan_set_direct_pk: bad arguments");
  plan->drop_graphs();
  VG_CUDA_CHECK(cudaDeviceSynchronize());
  VG_CUDA_CHECK(cudaMemcpy(plan->d_pk[3].p, pk_host, (size_t)plan->nk * 8, cudaMemcpyHostToDevice));
  return VG_OK;
}

int vg_plan_set_theta_transform(vg_plan* plan, const double* offset, const int32_t* force_one) {
  VG_REQUIRE(plan && offset && force_one, "vg_plan_set_theta_transform: bad arguments");
  if (plan->n_theta == 0) return VG_OK;
  VG_TRY(upload(plan->d_theta_offset, offset, plan->n_theta));
  VG_TRY(upload(plan->d_theta_force, force_one, plan->n_theta));
  plan->has_theta_transform = true;
  return VG_OK;
}

int vg_plan_set_theta_limits(vg_plan* plan, const double* lo, const double* hi) {
  VG_REQUIRE(plan && lo && hi, "vg_plan_set_theta_limits: bad arguments");
  VG_REQUIRE(!plan->finalized, "vg_plan_set_theta_limits: call before vg_plan_finalize");
  plan->theta_lo.assign(lo, lo + plan->n_theta);
  plan->theta_hi.assign(hi, hi + plan->n_theta);
  return VG_OK;
}

int vg_plan_set_fvoigt(vg_plan* plan, int32_t n, const double* x, const double* f) {
  VG_REQUIRE(plan && n > 1, "bad fvoigt table");
  VG_TRY(upload(plan->d_fv_x, x, n));
  VG_TRY(upload(plan->d_fv_f, f, n));
  plan->fv_n = n;
  return VG_OK;
}

int vg_plan_set_uvshot_table(vg_plan* plan, int32_t n, const double* tau, const double* a) {
  VG_REQUIRE(plan && n > 1, "bad uv table");
  VG_TRY(upload(plan->d_uv_tau, tau, n));
  VG_TRY(upload(plan->d_uv_a, a, n));
  plan->uv_n = n;
  return VG_OK;
}

int vg_plan_set_priors(vg_plan* plan, int32_t n, const int32_t* cols, const double* mean,
                       const double* sigma) {
  VG_REQUIRE(plan && n >= 0, "bad priors");
  plan->n_prior = n;
  if (n == 0) return VG_OK;
  VG_TRY(upload(plan->d_prior_cols, cols, n));
  VG_TRY(upload(plan->d_prior_mean, mean, n));
  VG_TRY(upload(plan->d_prior_sigma, sigma, n));
  return VG_OK;
}

int vg_plan_add_hankel(vg_plan* plan, int32_t n_ell, const int32_t* ells, int32_t n_out,
                       const double* xk, const double* op, int32_t bounds_error, int32_t* id) {
  VG_REQUIRE(plan && n_ell == plan->n_ell && n_out > 3, "bad hankel operator");
  std::unique_ptr<Hankel> h(new Hankel);
  h->n_ell = n_ell; h->n_out = n_out; h->n_out_pad = pad16(n_out) ; h->bounds_error = bounds_error;
  for (int l = 0; l < n_ell; ++l) {
    h->ells[l] = ells[l];
    h->x0[l] = xk[(size_t)l * n_out];
    h->x1[l] = xk[(size_t)l * n_out + n_out - 1];
  }
  h->uniform = 1;
  for (int l = 0; l < n_ell; ++l) {
    const double* x = xk + (size_t)l * n_out;
    const double mean = (x[n_out - 1] - x[0]) / (n_out - 1);
    for (int j = 0; j + 1 < n_out; ++j) {
      if (std::fabs((x[j + 1] - x[j]) / mean - 1.0) > 1e-7) h->uniform = 0;
      if (std::fabs((x[j] - (x[0] + j * mean)) / mean) > 0.25) h->uniform = 0;
    }
  }
  VG_TRY(h->xk.alloc((size_t)n_ell * h->n_out_pad * 8));
  VG_CUDA_CHECK(cudaMemcpy2D(h->xk.p, (size_t)h->n_out_pad * 8, xk, (size_t)n_out * 8,
                             (size_t)n_out * 8, n_ell, cudaMemcpyHostToDevice));
  {
    std::vector<double> inv((size_t)n_ell * h->n_out_pad, 0.0);
    for (int l = 0; l < n_ell; ++l)
      for (int j = 0; j + 1 < n_out; ++j)
        inv[(size_t)l * h->n_out_pad + j] = 1.0 / (xk[(size_t)l * n_out + j + 1] - xk[(size_t)l * n_out + j]);
    VG_TRY(upload(h->inv_h, inv.data(), inv.size()));
  }
  // op[l][2][n_out][nk] -> device [l][2*n_out_pad][nk_pad], second derivatives at row n_out_pad
  const size_t rows = 2 * (size_t)h->n_out_pad;
  VG_TRY(h->op.alloc((size_t)n_ell * rows * plan->nk_pad * 8));
  for (int l = 0; l < n_ell; ++l)
    for (int half = 0; half < 2; ++half) {
      const double* src = op + ((size_t)l * 2 + half) * n_out * plan->nk;
      double* dst = h->op.as<double>() + ((size_t)l * rows + (size_t)half * h->n_out_pad) * plan->nk_pad;
      VG_CUDA_CHECK(cudaMemcpy2D(dst, (size_t)plan->nk_pad * 8, src, (size_t)plan->nk * 8,
                                 (size_t)plan->nk * 8, n_out, cudaMemcpyHostToDevice));
    }
  *id = (int)plan->hankels.size();
  plan->hankels.push_back(std::move(h));
  return VG_OK;
}

int vg_plan_add_corr(vg_plan* plan, int32_t hankel_id, int32_t n_model, int32_t n_dist,
                     int32_t* corr_id) {
  VG_REQUIRE(plan && hankel_id >= 0 && hankel_id < (int)plan->hankels.size() && n_model > 0 &&
                 n_dist > 0, "bad correlation");
  std::unique_ptr<Corr> c(new Corr);
  c->hankel_id = hankel_id; c->n_model = n_model; c->n_dist = n_dist;
  c->ld_model = pad16(n_model); c->ld_dist = pad16(n_dist);
  *corr_id = (int)plan->corrs.size();
  plan->corrs.push_back(std::move(c));
  return VG_OK;
}

#define VG_CORR(plan, corr)                                                        \
  VG_REQUIRE(plan && corr >= 0 && corr < (int)plan->corrs.size(), "bad correlation id"); \
  Corr& c = *plan->corrs[corr]

int vg_corr_add_coords(vg_plan* plan, int32_t corr, const double* tab, double opz_eff2,
                       int32_t* coord_set) {
  VG_CORR(plan, corr);
  // device table: the 6 rows of the caller + r mu and r sqrt(1 - mu^2) (the unscaled rp, rt of
  // vega/correlation_func.py:226-230), so the gather does not take a square root per bin for them
  const size_t n = (size_t)c.n_model;
  std::vector<double> host(8 * n);
  std::copy(tab, tab + 6 * n, host.begin());
  for (size_t i = 0; i < n; ++i) {
    const double r = tab[i], mu = tab[n + i];
    host[6 * n + i] = r * mu;
    host[7 * n + i] = r * std::sqrt(1.0 - mu * mu);
  }
  DevBuf buf;
  VG_TRY(upload(buf, host.data(), host.size()));
  *coord_set = (int)c.coords.size();
  c.coords.push_back(std::move(buf));
  c.opz_eff2.push_back(opz_eff2);
  return VG_OK;
}

int vg_corr_add_eval(vg_plan* plan, int32_t corr, const vg_eval_desc* desc, int32_t* eval_id) {
  VG_CORR(plan, corr);
  VG_REQUIRE(desc->coord_set >= 0 && desc->coord_set < (int)c.coords.size(), "eval: bad coord_set");
  VG_REQUIRE(desc->out_slot >= 0, "eval: bad out_slot");
  *eval_id = (int)c.evals.size();
  c.evals.push_back(*desc);
  c.n_metal = std::max(c.n_metal, desc->out_slot - 1);
  return VG_OK;
}

int vg_corr_set_static_xi(vg_plan* plan, int32_t corr, int32_t which, int32_t pk_lin_sel, int32_t n,
                          const double* tab) {
  VG_CORR(plan, corr);
  VG_REQUIRE(which >= 0 && which < 4 && pk_lin_sel >= 0 && pk_lin_sel < 3 && n > 3, "bad static xi table");
  VG_TRY(upload(c.d_sx[which * 3 + pk_lin_sel], tab, (size_t)3 * n));
  c.sx_n[which * 3 + pk_lin_sel] = n;
  return VG_OK;
}

int vg_corr_set_metal_matrix(vg_plan* plan, int32_t corr, int32_t slot, const double* mat) {
  VG_CORR(plan, corr);
  VG_REQUIRE(slot >= 0 && slot < 4096, "bad metal slot");
  if ((int)c.metal_host.size() <= slot) c.metal_host.resize(slot + 1);
  c.metal_host[slot].assign(mat, mat + (size_t)c.n_model * c.n_model);
  c.metal_dense = true;
  return VG_OK;
}

int vg_corr_set_metal_matrix_rp(vg_plan* plan, int32_t corr, int32_t slot, const double* mat,
                                int32_t n_rp, int32_t n_rt) {
  VG_CORR(plan, corr);
  VG_REQUIRE(slot >= 0 && slot < 4096, "bad metal slot");
  VG_REQUIRE(n_rp > 0 && n_rt > 0 && n_rp * n_rt == c.n_model, "rp-only metal matrix: %d x %d bins != %d",
             n_rp, n_rt, c.n_model);
  VG_REQUIRE(n_rp <= vg::MRP_MAX_NP, "rp-only metal matrix: more than %d rp bins", vg::MRP_MAX_NP);
  VG_REQUIRE(!c.metal_rp || (c.rp_np == n_rp && c.rp_nt == n_rt), "rp-only metal matrices of different shapes");
  if ((int)c.metal_rp_host.size() <= slot) c.metal_rp_host.resize(slot + 1);
  c.metal_rp_host[slot].assign(mat, mat + (size_t)n_rp * n_rp);
  c.metal_rp = true; c.rp_np = n_rp; c.rp_nt = n_rt;
  return VG_OK;
}

int vg_corr_set_distortion(vg_plan* plan, int32_t corr, const double* dmat) {
  VG_CORR(plan, corr);
  VG_TRY(upload_padded(c.d_dmat, dmat, c.n_dist, c.n_model, c.ld_model));
  c.has_dmat = true;
  return VG_OK;
}

int vg_corr_set_inst_sys(vg_plan* plan, int32_t corr, const double* table, int32_t idx_amp,
                         double amp_default) {
  VG_CORR(plan, corr);
  VG_TRY(upload(c.d_inst, table, c.n_model));
  c.has_inst = true; c.idx_inst_amp = idx_amp; c.inst_amp_default = amp_default;
  return VG_OK;
}

int vg_corr_set_broadband(vg_plan* plan, int32_t corr, int32_t n_terms, const vg_bb_term* terms,
                          int32_t rows_pre, const double* design_pre, int32_t rows_post,
                          const double* design_post, const double* sky_pre, const double* sky_post) {
  VG_CORR(plan, corr);
  VG_REQUIRE(n_terms >= 0 && n_terms <= VG_MAX_BB_TERMS, "too many broadband terms");
  c.bb_terms.assign(terms, terms + n_terms);
  if (n_terms) VG_TRY(upload(c.d_bb_terms, terms, n_terms));
  c.rows_pre = rows_pre; c.rows_post = rows_post;
  if (rows_pre) VG_TRY(upload(c.d_design_pre, design_pre, (size_t)rows_pre * c.n_model));
  if (rows_post) VG_TRY(upload(c.d_design_post, design_post, (size_t)rows_post * c.n_dist));
  VG_TRY(upload(c.d_sky_pre, sky_pre, (size_t)2 * c.n_model));
  VG_TRY(upload(c.d_sky_post, sky_post, (size_t)2 * c.n_dist));
  return VG_OK;
}

int vg_corr_set_data(vg_plan* plan, int32_t corr, int32_t n_mask, const int32_t* mask_idx,
                     const double* data, const double* invcov) {
  VG_CORR(plan, corr);
  VG_REQUIRE(n_mask > 0, "empty data mask");
  plan->drop_graphs();
  for (int j = 0; j < n_mask; ++j)
    VG_REQUIRE(mask_idx[j] >= 0 && mask_idx[j] < c.n_dist, "mask index out of range");
  const bool new_mask = c.mask_host.size() != (size_t)n_mask ||
                        !std::equal(mask_idx, mask_idx + n_mask, c.mask_host.begin());
  c.n_mask = n_mask; c.ld_mask = pad16(n_mask);
  if (new_mask) {
    c.mask_host.assign(mask_idx, mask_idx + n_mask);
    VG_TRY(upload(c.d_mask_idx, mask_idx, n_mask));
    if (plan->finalized && c.has_dmat) {
      vg_set_error("the data mask cannot change after vg_plan_finalize");
      return VG_ERR_STATE;
    }
  }
  VG_TRY(upload(c.d_data, data, n_mask));
  if (invcov) {
    VG_TRY(upload_padded(c.d_invcov, invcov, n_mask, n_mask, c.ld_mask));
    VG_TRY(symmetrize(c.d_invcov, n_mask, c.ld_mask));
    c.has_invcov = true;
    c.has_white = false;          // a square root of the OLD matrix: vg_corr_set_whitening gives the new one
  } else {
    c.has_invcov = false;
    c.has_white = false;
  }
  c.white_dirty = true;
  return VG_OK;
}

int vg_corr_set_data_vector(vg_plan* plan, int32_t corr, int32_t n_mask, const double* data) {
  VG_CORR(plan, corr);
  VG_REQUIRE(data && n_mask == c.n_mask && c.d_data.p, "vg_corr_set_data_vector: call vg_corr_set_data first (same mask)");
  plan->drop_graphs();
  VG_CUDA_CHECK(cudaDeviceSynchronize());
  VG_CUDA_CHECK(cudaMemcpy(c.d_data.p, data, (size_t)n_mask * 8, cudaMemcpyHostToDevice));
  c.wdata_dirty = true;
  return VG_OK;
}

int vg_corr_set_whitening(vg_plan* plan, int32_t corr, int32_t n_mask, const double* w) {
  VG_CORR(plan, corr);
  plan->drop_graphs();
  c.white_dirty = true;
  if (!w) { c.has_white = false; return VG_OK; }
  VG_REQUIRE(n_mask == c.n_mask && c.has_invcov, "vg_corr_set_whitening: call after vg_corr_set_data with an inverse covariance of the same size");
  VG_TRY(upload_padded(c.d_white, w, n_mask, n_mask, c.ld_mask));
  c.has_white = true;
  return VG_OK;
}

int vg_plan_set_global_invcov(vg_plan* plan, int32_t n_total, const double* invcov) {
  VG_REQUIRE(plan && n_total > 0, "bad global covariance");
  plan->drop_graphs();
  plan->n_total = n_total; plan->ld_total = pad16(n_total);
  VG_TRY(upload_padded(plan->d_global_invcov, invcov, n_total, n_total, plan->ld_total));
  VG_TRY(symmetrize(plan->d_global_invcov, n_total, plan->ld_total));
  plan->has_global = true;
  return VG_OK;
}

int vg_plan_finalize(vg_plan* plan, void* stream) {
  VG_REQUIRE(plan && !plan->corrs.empty(), "plan has no correlation");
  plan->drop_graphs();
  plan->vt_now = false;
  cudaStream_t st = (cudaStream_t)stream;
  // number of static grids referenced
  int n_grids = 0;
  for (auto& cp : plan->corrs)
    for (auto& d : cp->evals) n_grids = std::max({n_grids, d.static_grid + 1, d.hcd_grid + 1});
  plan->grids.clear();
  plan->grids.resize(n_grids);
  std::vector<const double*> ptrs(std::max(n_grids, 1), nullptr);
  for (int g = 0; g < n_grids; ++g) {
    VG_TRY(plan->grids[g].alloc((size_t)plan->nmu * plan->nk * 8));
    ptrs[g] = plan->grids[g].as<double>();
  }
  VG_TRY(upload(plan->d_grid_ptrs, ptrs.data(), ptrs.size()));

  plan->std_ells = 1;
  for (auto& h : plan->hankels)
    for (int l = 0; l < vg::PK_MAX_ELL; ++l)
      if (h->n_ell != 4 || h->ells[l] != 2 * l) plan->std_ells = 0;
  DevBuf bad;
  VG_TRY(bad.alloc(sizeof(int)));
  std::vector<char> built(n_grids, 0);
  plan->static_bad = 0;
  for (auto& cp : plan->corrs) {
    Corr& c = *cp;
    VG_REQUIRE(!c.evals.empty(), "correlation without evaluations");
    c.allow_mxi = true;
    VG_TRY(upload(c.d_evals, c.evals.data(), c.evals.size()));
    flag_velo_table_candidates(plan, c);
    VG_TRY(build_layouts(plan, c));
    std::vector<const double*> cptrs;
    for (auto& b : c.coords) cptrs.push_back(b.as<double>());
    VG_TRY(upload(c.d_coord_ptrs, cptrs.data(), cptrs.size()));
    VG_TRY(upload(c.d_opz_eff2, c.opz_eff2.data(), c.opz_eff2.size()));
    {   // static bias-evolution x growth rows of the evaluations whose exponents are not sampled
      const int E = (int)c.evals.size();
      VG_TRY(c.d_evolfac.alloc((size_t)E * c.ld_model * 8));
      dim3 egrid(vg_cdiv(c.ld_model, 256), (unsigned)E);
      vg::evol_rows_kernel<<<egrid, 256, 0, st>>>(c.d_evals.as<vg_eval_desc>(), E, plan->d_base.as<double>(),
                                                 c.d_coord_ptrs.as<const double*>(), c.d_opz_eff2.as<double>(),
                                                 c.n_model, c.ld_model, c.d_evolfac.as<double>());
      VG_LAUNCHED();
    }
    vg::PkGridDev g = plan->grid_dev();
    dim3 grid(vg_cdiv(plan->nk, 128), plan->nmu);
    for (size_t e = 0; e < c.evals.size(); ++e) {
      const vg_eval_desc& d = c.evals[e];
      if (d.static_grid >= 0 && !built[d.static_grid]) {
        vg::static_grid_kernel<<<grid, 128, 0, st>>>(g, c.d_evals.as<vg_eval_desc>(), (int)e, 0,
                                                   plan->d_base.as<double>(),
                                                   plan->grids[d.static_grid].as<double>(), bad.as<int>());
        VG_LAUNCHED();
        built[d.static_grid] = 1;
      }
      if (d.hcd_grid >= 0 && !built[d.hcd_grid]) {
        vg::static_grid_kernel<<<grid, 128, 0, st>>>(g, c.d_evals.as<vg_eval_desc>(), (int)e, 1,
                                                   plan->d_base.as<double>(),
                                                   plan->grids[d.hcd_grid].as<double>(), bad.as<int>());
        VG_LAUNCHED();
        built[d.hcd_grid] = 1;
      }
    }
    // moment form: static knot tables K[j][m] = (Hankel + spline operator)(pk_lin * moment m)
    EvalLayout& LM = c.lay[LAY_MOM];
    if (!LM.mom_ids.empty()) {
      const Hankel& h = *plan->hankels[c.hankel_id];
      const int rowT = 2 * h.n_out_pad;
      const int rows = (int)LM.mom_ids.size() * vg::MOM_MAX;
      DevBuf mom;
      VG_TRY(mom.alloc((size_t)rows * plan->n_ell * plan->nk_pad * 8));
      VG_TRY(c.d_mom_ktab.alloc((size_t)rows * plan->n_ell * rowT * 8));
      dim3 mgrid(vg_cdiv(plan->nk, 128), (unsigned)LM.mom_ids.size());
      vg::moment_grid_kernel<<<mgrid, 128, 0, st>>>(g, c.d_evals.as<vg_eval_desc>(), LM.d_mom_ids.as<int>(),
                                                  plan->d_base.as<double>(), mom.as<double>());
      VG_LAUNCHED();
      VG_TRY(gemm_auto(plan, mom.as<double>(), (long long)plan->n_ell * plan->nk_pad, h.op.as<double>(),
                       plan->nk_pad, c.d_mom_ktab.as<double>(), (long long)plan->n_ell * rowT, rows, rowT,
                       plan->nk_pad, plan->n_ell, plan->nk_pad, (long long)rowT * plan->nk_pad, rowT, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));   // `mom` is released at the end of this scope
    }
    // xi stage in moment form: X[j][m] = gather of the static table K[owner(j)][m] on evaluation
    // j's coordinates at the base parameters (nothing it reads is sampled), weight_const included
    if (!LM.mxi_ids.empty()) {
      const Hankel& h = *plan->hankels[c.hankel_id];
      const long long row_len = (long long)plan->n_ell * 2 * h.n_out_pad;
      const int n_mxi = (int)LM.mxi_ids.size();
      VG_TRY(c.d_mxi_tab.alloc((size_t)n_mxi * vg::MOM_MAX * c.ld_model * 8));
      std::vector<double> base9((size_t)vg::MOM_MAX * plan->npf);
      for (int m = 0; m < vg::MOM_MAX; ++m) std::copy(plan->base.begin(), plan->base.end(), base9.begin() + (size_t)m * plan->npf);
      DevBuf d_base9, d_stat;
      VG_TRY(upload(d_base9, base9.data(), base9.size()));
      VG_TRY(d_stat.alloc(vg::MOM_MAX * sizeof(int)));
      VG_CUDA_CHECK(cudaMemsetAsync(d_stat.p, 0, vg::MOM_MAX * sizeof(int), st));
      GatherJob job{};
      job.ids = LM.d_mxi_ids.as<int>(); job.n_ids = n_mxi;
      job.tab_slot = LM.d_tabslot.as<int>(); job.metal_map = LM.d_metal_map.as<int>(); job.n_metal = LM.n_metal;
      job.pfull = d_base9.as<double>(); job.B = vg::MOM_MAX;
      job.tables = nullptr; job.n_full = 0;
      job.tables_mom = c.d_mom_ktab.as<double>(); job.n_mom = (int)LM.mom_ids.size();
      job.mom_stride_b = row_len; job.mom_stride_j = (long long)vg::MOM_MAX * row_len;   // K is [j][m]
      job.unit_bias = 1;
      job.flat_out = c.d_mxi_tab.as<double>();
      job.status = d_stat.as<int>();
      VG_TRY(launch_xi_gather(plan, c, job, st));
      int stat[vg::MOM_MAX];
      VG_CUDA_CHECK(cudaMemcpyAsync(stat, d_stat.p, sizeof(stat), cudaMemcpyDeviceToHost, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));
      int any = 0;
      for (int m = 0; m < vg::MOM_MAX; ++m) any |= stat[m];
      if (any) {   // out-of-bounds / non-finite bins must be reported per call: keep the gather
        c.allow_mxi = false;
        c.d_mxi_tab.release();
        VG_TRY(build_layouts(plan, c));
      }
    }
    // velocity-dispersion tables (vg_moments.cuh): built and checked against the per-point pass
    if (!LM.vt_ids.empty()) VG_TRY(build_velo_tables(plan, c, st));
    VG_REQUIRE(!(c.metal_rp && c.metal_dense), "a correlation mixes dense and rp-only metal matrices");
    VG_TRY(build_metal_operands(c, LM, c.d_mcat, c.d_mrp));
    // folded metal terms: Y[f][m] = M_c X[row(f)][m], and the matrices of the terms that stay
    {
      const EvalLayout& LF = c.lay[LAY_FOLD];
      const int nf = (int)LF.fold_ids.size();
      if (nf > 0) {
        const size_t blk = (size_t)vg::MOM_MAX * c.ld_model;
        VG_TRY(c.d_fold_tab.alloc((size_t)nf * blk * 8));
        DevBuf mat;
        for (int f = 0; f < nf; ++f) {
          const int slot = c.evals[LF.fold_ids[f]].out_slot - 2;
          const double* src = c.d_mxi_tab.as<double>() + (size_t)LF.fold_rows[f] * blk;
          double* dst = c.d_fold_tab.as<double>() + (size_t)f * blk;
          if (c.metal_rp && slot < (int)c.metal_rp_host.size() && !c.metal_rp_host[slot].empty()) {
            VG_TRY(upload(mat, c.metal_rp_host[slot].data(), c.metal_rp_host[slot].size()));
            VG_TRY(launch_metal_rp(c, src, 1, mat.as<double>(), vg::MOM_MAX, dst, st));
          } else if (c.metal_dense && slot < (int)c.metal_host.size() && !c.metal_host[slot].empty()) {
            VG_TRY(upload_padded(mat, c.metal_host[slot].data(), c.n_model, c.n_model, c.ld_model));
            VG_TRY(gemm_auto(plan, src, c.ld_model, mat.as<double>(), c.ld_model, dst, c.ld_model,
                             vg::MOM_MAX, c.n_model, c.ld_model, 1, 0, 0, 0, st));
          } else {
            VG_CUDA_CHECK(cudaMemcpyAsync(dst, src, blk * 8, cudaMemcpyDeviceToDevice, st));
          }
          VG_CUDA_CHECK(cudaStreamSynchronize(st));   // `mat` is re-used by the next term
        }
        VG_TRY(build_metal_operands(c, LF, c.d_mcat_fold, c.d_mrp_fold));
      }
    }
    // masked rows of the distortion matrix (chi2 needs only the bins kept by the cuts)
    if (c.has_dmat && c.n_mask > 0) {
      VG_TRY(c.d_dmat_masked.alloc((size_t)c.n_mask * c.ld_model * 8));
      for (int j = 0; j < c.n_mask; ++j)
        VG_CUDA_CHECK(cudaMemcpyAsync(c.d_dmat_masked.as<double>() + (size_t)j * c.ld_model,
                                      c.d_dmat.as<double>() + (size_t)c.mask_host[j] * c.ld_model,
                                      (size_t)c.ld_model * 8, cudaMemcpyDeviceToDevice, st));
    }
  }
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  VG_CUDA_CHECK(cudaMemcpy(&plan->static_bad, bad.p, sizeof(int), cudaMemcpyDeviceToHost));
  for (auto& cp : plan->corrs) {
    VG_TRY(freeze_metal_terms(plan, *cp, st));
    VG_TRY(ozaki_budget(plan, *cp, st));
    VG_TRY(ozaki_budget_dm(plan, *cp, st));
    if (cp->n_mask > 0) VG_TRY(rebuild_whitened(plan, *cp, st));   // whitened operands + the fused operator
    else cp->white_dirty = true;                                    // (data not set yet: at the first chi2 call)
    // (the host copies of the metal matrices stay: vg_plan_set_templates runs finalize again)
  }
  if (plan->has_global) {
    int tot = 0;
    for (auto& cp : plan->corrs) tot += cp->n_mask;
    VG_REQUIRE(tot == plan->n_total, "global covariance size %d != sum of masks %d", plan->n_total, tot);
  }
  plan->finalized = true;
  return VG_OK;
}

int64_t vg_plan_device_bytes(const vg_plan* plan) {
  if (!plan) return 0;
  size_t n = 0;
  for (auto& g : plan->grids) n += g.bytes;
  for (auto& h : plan->hankels) n += h->xk.bytes + h->op.bytes;
  for (auto& cp : plan->corrs) {
    const Corr& c = *cp;
    n += c.d_mcat.bytes + c.d_dmat.bytes + c.d_dmat_masked.bytes + c.d_invcov.bytes + c.d_mom_ktab.bytes + c.d_frozen.bytes +
         c.d_mxi_tab.bytes + c.d_fold_tab.bytes + c.d_mcat_fold.bytes + c.d_mrp.bytes + c.d_mrp_fold.bytes;
    for (auto& b : c.coords) n += b.bytes;
    // whitened / fused operators, Monte Carlo mocks, and the digit planes built so far
    n += c.d_white.bytes + c.d_wd.bytes + c.d_wdata.bytes + c.d_wmocks.bytes + c.d_dmcat.bytes + c.d_mocks.bytes +
         c.d_evolfac.bytes;
    for (const OzakiOperand* op : {&c.oz_mcat, &c.oz_mcat_fold, &c.oz_dmat, &c.oz_dmat_masked, &c.oz_wd, &c.oz_dmcat})
      n += op->planes.bytes + op->exps.bytes;
    for (auto& t : c.vt_tab) n += t.bytes;
  }
  n += plan->d_global_invcov.bytes;
  n += plan->w_pfull.bytes + plan->w_pkell.bytes + plan->w_tables.bytes + plan->w_core.bytes +
       plan->w_xmet.bytes + plan->w_met.bytes + plan->w_x.bytes + plan->w_y.bytes + plan->w_diff.bytes +
       plan->w_partial.bytes + plan->w_split.bytes + plan->w_fold.bytes + plan->w_mety.bytes + plan->w_vtw.bytes +
       plan->oz_A.planes.bytes + plan->oz_A.exps.bytes;
  for (const auto& l : plan->lanes)     // the lanes' copies of the per-correlation workspace (VG_OPT_LANES)
    n += l.w_pkell.bytes + l.w_tables.bytes + l.w_core.bytes + l.w_xmet.bytes + l.w_met.bytes + l.w_x.bytes +
         l.w_y.bytes + l.w_diff.bytes + l.w_partial.bytes + l.w_split.bytes + l.w_fold.bytes + l.w_mety.bytes +
         l.w_vtw.bytes + l.oz_A.planes.bytes + l.oz_A.exps.bytes + l.chi2.bytes;
  return (int64_t)n;
}

}  // extern "C"

namespace {


// split-K so that the CTA count fills whole waves of the 148 SMs (1 CTA/SM): the split count
// with the best wave efficiency, keeping >= 16 K-tiles per split
int pick_splits(long long tiles, int ktiles, int ctas_per_sm = 1) {
  int splits = 1;
  const int NUM_SMS = ::NUM_SMS * ctas_per_sm;   // CTA slots of one wave
  if (tiles < 8 * NUM_SMS) {
    double best = 0.0;
    const int max_splits = (int)std::min<long long>(32, std::max(1, ktiles / 16));
    for (int sp = 1; sp <= max_splits; ++sp) {
      const long long ctas = tiles * sp;
      const long long waves = (ctas + NUM_SMS - 1) / NUM_SMS;
      double eff = (double)ctas / (double)(waves * NUM_SMS);
      eff -= 0.004 * (sp - 1);            // the partial-sum write/read of each extra split
      if (eff > best + 1e-9) { best = eff; splits = sp; }
    }
  }
  return splits;
}

// C[M][ldc] = A[M][K] * B[N][K]^T with an automatic deterministic split-K when the tile
// count is too small to fill the 148 SMs.
int gemm_auto(vg_plan* plan, const double* A, long long lda, const double* B, long long ldb,
              double* C, long long ldc, int M, int N, int K, int batch, long long strideA,
              long long strideB, long long strideC, cudaStream_t st) {
  vg::GemmArgs g{};
  g.A = A; g.B = B; g.M = M; g.N = N; g.K = K; g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.batch = batch; g.strideA = strideA; g.strideB = strideB; g.strideC = strideC;
  g.alpha = 1.0; g.beta = 0.0;
  const int bm = vg::gemm_tile_m(M);
  const long long tiles = (long long)vg_cdiv(M, bm) * vg_cdiv(N, vg::GEMM_BN) * batch;
  const int ktiles = K / vg::GEMM_BK;
  const int splits = pick_splits(tiles, ktiles, bm == vg::GEMM_BM ? 1 : 2);
  g.splits = std::max(1, splits);
  if (g.splits == 1) {
    g.C = C;
    VG_TRY(vg::launch_dgemm(g, vg::GEMM_STORE, st));
    return VG_OK;
  }
  const long long n = (long long)M * ldc;
  VG_TRY(plan->w_split.ensure((size_t)g.splits * n * 8));
  g.C = plan->w_split.as<double>();
  g.split_stride = n;
  VG_TRY(vg::launch_dgemm(g, vg::GEMM_STORE, st));
  vg::reduce_splits_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(plan->w_split.as<double>(), C, n, g.splits, n);
  VG_LAUNCHED();
  return VG_OK;
}

// ---- split-integer GEMM on tcgen05 (vg_ozaki.cuh) ------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    return (EncodeTiledFn)p;
  }();
  return fn;
}

template <int S>
int ozaki_slice_launch(const double* src, long long ld, int K, int Kp, long long rows, const int* exps,
                       int8_t* planes, long long plane_stride, cudaStream_t st) {
  const long long n = rows * (Kp / 4);
  vg::ozaki_slice_kernel<S><<<vg_cdiv(n, 256), 256, 0, st>>>(src, ld, K, Kp, rows, exps, planes, plane_stride);
  VG_LAUNCHED();
  return VG_OK;
}

// digit planes + row exponents of src[rows][ld] (first K columns).  The buffers only grow (room for
// `capacity` rows is taken at once), so a workspace operand re-sliced every call never reallocates.
int ozaki_slice(OzakiOperand& op, const double* src, long long ld, long long rows, long long capacity, int K, int S,
                cudaStream_t st) {
  const int Kp = (K + 15) / 16 * 16;
  capacity = std::max(capacity, rows);
  const size_t need = (size_t)S * rows * Kp;
  if (op.planes.bytes < need || op.exps.bytes < (size_t)rows * sizeof(int)) {
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
    VG_TRY(op.planes.alloc(std::max<size_t>(op.planes.bytes, (size_t)S * (size_t)capacity * (size_t)Kp), false));
    VG_TRY(op.exps.alloc(std::max(op.exps.bytes, (size_t)capacity * sizeof(int)), true));
  }
  op.S = S; op.Kp = Kp; op.rows = rows;
  vg::ozaki_rowexp_kernel<<<(unsigned)rows, 256, 0, st>>>(src, ld, K, op.exps.as<int>());
  VG_LAUNCHED();
  int8_t* planes = op.planes.as<int8_t>();
  const long long ps = op.rows * (long long)Kp;
  switch (S) {
    case 3: return ozaki_slice_launch<3>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
    case 4: return ozaki_slice_launch<4>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
    case 5: return ozaki_slice_launch<5>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
    case 6: return ozaki_slice_launch<6>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
    case 7: return ozaki_slice_launch<7>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
    default: return ozaki_slice_launch<8>(src, ld, K, Kp, rows, op.exps.as<int>(), planes, ps, st);
  }
}

// 3-D tensor map over the digit planes: (k, row, digit), box (OZ_BK, box_rows, box_planes), OZ_BK-byte swizzle
int ozaki_map(CUtensorMap* map, const OzakiOperand& op, int box_rows, int box_planes) {
  EncodeTiledFn fn = encode_tiled_fn();
  VG_REQUIRE(fn, "cuTensorMapEncodeTiled is not available from this driver");
  const cuuint64_t gdim[3] = {(cuuint64_t)op.Kp, (cuuint64_t)op.rows, (cuuint64_t)op.S};
  const cuuint64_t gstride[2] = {(cuuint64_t)op.Kp, (cuuint64_t)op.rows * (cuuint64_t)op.Kp};
  const cuuint32_t box[3] = {(cuuint32_t)vg::OZ_BK, (cuuint32_t)box_rows, (cuuint32_t)box_planes};
  const cuuint32_t estride[3] = {1, 1, 1};
  const CUresult rc = fn(map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, op.planes.p, gdim, gstride, box, estride,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, vg::OZ_BK == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  VG_REQUIRE(rc == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with %d", (int)rc);
  return VG_OK;
}

// CTAs per cluster of the split-integer GEMM (1, 2 or 4; VG_OZ_CLUSTER in the environment): the CTAs
// of a cluster share the batch operand's tile through multicast loads (vg_ozaki.cuh)
int ozaki_cluster() {
  static const int cl = [] {
    const char* e = std::getenv("VG_OZ_CLUSTER");
    const int v = e ? std::atoi(e) : 2;
    return (v == 1 || v == 2 || v == 4) ? v : 2;
  }();
  return cl;
}

template <int S, int CL>
int ozaki_gemm_launch_cl(const CUtensorMap& ma, const CUtensorMap& ma_rest, const CUtensorMap& mb, const vg::OzakiArgs& g,
                         dim3 grid, cudaStream_t st) {
  static bool attr_set = false;
  if (!attr_set) {
    VG_CUDA_CHECK(cudaFuncSetAttribute(vg::ozaki_gemm_kernel<S, CL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       vg::OzakiSmem<S>::BYTES));
    attr_set = true;
  }
  if (CL == 1) {
    vg::ozaki_gemm_kernel<S, CL><<<grid, vg::OZ_THREADS, vg::OzakiSmem<S>::BYTES, st>>>(ma, ma_rest, mb, g);
  } else {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = dim3(vg::OZ_THREADS); cfg.dynamicSmemBytes = vg::OzakiSmem<S>::BYTES; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = CL; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    VG_CUDA_CHECK(cudaLaunchKernelEx(&cfg, vg::ozaki_gemm_kernel<S, CL>, ma, ma_rest, mb, g));
  }
  VG_LAUNCHED();
  return VG_OK;
}
template <int S>
int ozaki_gemm_launch(int cl, const CUtensorMap& ma, const CUtensorMap& ma_rest, const CUtensorMap& mb,
                      const vg::OzakiArgs& g, dim3 grid, cudaStream_t st) {
  if (cl == 4) return ozaki_gemm_launch_cl<S, 4>(ma, ma_rest, mb, g, grid, st);
  if (cl == 2) return ozaki_gemm_launch_cl<S, 2>(ma, ma_rest, mb, g, grid, st);
  return ozaki_gemm_launch_cl<S, 1>(ma, ma_rest, mb, g, grid, st);
}

// C[M][ldc] = A B^T from the digit planes of A ([>= M rows]) and B ([N rows]); deterministic split-K
int ozaki_gemm(const OzakiOperand& A, long long M, const OzakiOperand& B, int N, double* C, long long ldc,
               DevBuf& split_ws, DevBuf& error_flag, cudaStream_t st) {
  VG_REQUIRE(A.S == B.S && A.Kp == B.Kp && A.S >= 3, "ozaki_gemm: operands with different digit planes");
  const int S = A.S;
  if (!error_flag.p) VG_TRY(error_flag.alloc(sizeof(int)));
  const int tiles_n = vg_cdiv(N, vg::OZ_BN);
  const int cl = std::min(ozaki_cluster(), tiles_n >= 4 ? 4 : (tiles_n >= 2 ? 2 : 1));
  const int a_per = (S + cl - 1) / cl, a_rest = S % a_per;      // planes per CTA of a cluster; the last share
  CUtensorMap ma, ma_rest, mb;
  VG_TRY(ozaki_map(&ma, A, vg::OZ_BM, a_per));
  VG_TRY(ozaki_map(&ma_rest, A, vg::OZ_BM, a_rest > 0 ? a_rest : a_per));
  VG_TRY(ozaki_map(&mb, B, vg::OZ_BN, S));
  vg::OzakiArgs g{};
  g.M = (int)M; g.N = N; g.S = S;
  g.kb_total = vg_cdiv(A.Kp, vg::OZ_BK);
  const int grid_y = vg_cdiv(tiles_n, cl) * cl;
  const long long tiles = (long long)vg_cdiv(M, vg::OZ_BM) * grid_y;
  // an s32 accumulator takes S products of |q q'| <= 2^12 per k: K blocks per split are bounded
  const int kb_cap = (int)(((1LL << 31) - 1) / (4096LL * vg::OZ_BK * S));
  int splits = vg_cdiv(g.kb_total, kb_cap);
  {   // Split count by a cost model in units of one K block of one CTA (~1.8 us): waves of the 148 SMs
      // (one CTA each) x (the K blocks of a split + a CTA's fixed cost: TMEM allocation, pipeline fill,
      // the epilogue over S accumulators) + the partial sums written and read back by the reduction.
      // Long K (the metal matrices: 589 blocks) still takes the split that fills the last wave; short K
      // (the distortion matrix: 40-79 blocks) no longer pays 7 short CTAs and a reduction for it
    double best = 1e300;
    const int lo = splits, hi = std::max(lo, std::min(64, g.kb_total / (512 / vg::OZ_BK)));
    const double cta_fixed = 2.5;
    const double reduce_per_split = 2.0 * (double)M * (double)ldc * 8.0 / 6.0e12 / 1.8e-6;
    for (int sp = lo; sp <= hi; ++sp) {
      const int per = vg_cdiv(g.kb_total, sp);
      if ((long long)per * (sp - 1) >= g.kb_total) continue;      // an empty last split
      const long long ctas = tiles * sp;
      const long long waves = (ctas + NUM_SMS - 1) / NUM_SMS;
      const double cost = (double)waves * (cta_fixed + per) + (sp > 1 ? sp * reduce_per_split : 0.0);
      if (cost < best - 1e-9) { best = cost; splits = sp; }
    }
  }
  g.kb_per_split = vg_cdiv(g.kb_total, splits);
  g.expA = A.exps.as<int>(); g.expB = B.exps.as<int>();
  g.error_flag = error_flag.as<int>();
  g.ldc = ldc;
  const long long n = M * ldc;
  if (splits > 1) {
    VG_TRY(split_ws.ensure((size_t)splits * n * 8));
    g.C = split_ws.as<double>(); g.split_stride = n;
  } else {
    g.C = C; g.split_stride = 0;
  }
  dim3 grid(vg_cdiv(M, vg::OZ_BM), grid_y, splits);
  int rc;
  switch (S) {
    case 3: rc = ozaki_gemm_launch<3>(cl, ma, ma_rest, mb, g, grid, st); break;
    case 4: rc = ozaki_gemm_launch<4>(cl, ma, ma_rest, mb, g, grid, st); break;
    case 5: rc = ozaki_gemm_launch<5>(cl, ma, ma_rest, mb, g, grid, st); break;
    case 6: rc = ozaki_gemm_launch<6>(cl, ma, ma_rest, mb, g, grid, st); break;
    case 7: rc = ozaki_gemm_launch<7>(cl, ma, ma_rest, mb, g, grid, st); break;
    default: rc = ozaki_gemm_launch<8>(cl, ma, ma_rest, mb, g, grid, st); break;
  }
  VG_TRY(rc);
  if (splits > 1) {
    vg::reduce_splits_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(split_ws.as<double>(), C, n, splits, n);
    VG_LAUNCHED();
  }
  return VG_OK;
}

int ensure_workspace(vg_plan* plan, long long Bc_) {
  if (Bc_ <= plan->chunk) return VG_OK;
  const size_t Bc = (size_t)Bc_;
  size_t pkell = 0, tables = 0, core = 0, xmet = 0, met = 0, y = 0, diff = 0;
  for (auto& cp : plan->corrs) {
    const Corr& c = *cp;
    const Hankel& h = *plan->hankels[c.hankel_id];
    const size_t E = c.evals.size();
    pkell = std::max(pkell, (size_t)Bc * E * plan->n_ell * plan->nk_pad);
    tables = std::max(tables, (size_t)Bc * E * plan->n_ell * 2 * h.n_out_pad);
    core = std::max(core, (size_t)2 * Bc * c.ld_model);
    xmet = std::max(xmet, (size_t)Bc * std::max(c.n_metal, 1) * c.ld_model);
    met = std::max(met, (size_t)Bc * c.ld_model);
    y = std::max(y, (size_t)Bc * std::max(c.ld_dist, c.ld_mask));
    diff = std::max(diff, (size_t)Bc * std::max(c.ld_mask, c.ld_dist));
  }
  if (plan->has_global) diff = std::max(diff, (size_t)Bc * plan->ld_total);
  VG_TRY(plan->w_pfull.alloc((size_t)Bc * plan->npf * 8));
  VG_TRY(plan->w_pkell.alloc(pkell * 8));
  VG_TRY(plan->w_tables.alloc(tables * 8));
  VG_TRY(plan->w_core.alloc(core * 8));
  VG_TRY(plan->w_xmet.alloc(xmet * 8));
  VG_TRY(plan->w_met.alloc(met * 8));
  VG_TRY(plan->w_fold.alloc(met * 8));
  VG_TRY(plan->w_x.alloc(met * 8));
  VG_TRY(plan->w_y.alloc(y * 8));
  VG_TRY(plan->w_mety.alloc(y * 8));
  VG_TRY(plan->w_diff.alloc(diff * 8));
  VG_TRY(plan->w_partial.alloc((size_t)Bc * 256 * 8));
  VG_TRY(plan->w_chi2.alloc((size_t)Bc * 8));
  VG_TRY(plan->w_status.alloc((size_t)Bc * 4));
  VG_TRY(plan->w_theta.alloc((size_t)Bc * std::max(plan->n_theta, 1) * 8));
  plan->chunk = Bc_;
  return VG_OK;
}

int expand_params(vg_plan* plan, const double* theta_dev, long long Bc, int* status_dev, cudaStream_t st) {
  ProfScope prof(plan, PROF_ELEM, st);
  const long long n = Bc * plan->npf;
  vg::expand_theta_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(plan->d_base.as<double>(), plan->npf, theta_dev,
                                                        plan->n_theta, plan->d_theta_cols.as<int>(), Bc,
                                                        plan->w_pfull.as<double>());
  VG_LAUNCHED();
  if (plan->n_theta > 0) {
    const long long m = Bc * plan->n_theta;
    vg::scatter_theta_kernel<<<vg_cdiv(m, 256), 256, 0, st>>>(theta_dev, plan->n_theta,
                                                           plan->d_theta_cols.as<int>(), Bc, plan->npf,
                                                           plan->has_theta_transform ? plan->d_theta_offset.as<double>() : nullptr,
                                                           plan->has_theta_transform ? plan->d_theta_force.as<int>() : nullptr,
                                                           plan->w_pfull.as<double>());
    VG_LAUNCHED();
  }
  VG_CUDA_CHECK(cudaMemsetAsync(status_dev, 0, (size_t)Bc * 4, st));
  return VG_OK;
}

vg::BroadbandDev bb_dev(const Corr& c) {
  vg::BroadbandDev bb{};
  bb.n_terms = (int)c.bb_terms.size();
  bb.terms = c.d_bb_terms.as<vg_bb_term>();
  bb.design_pre = c.d_design_pre.as<double>(); bb.rows_pre = c.rows_pre;
  bb.design_post = c.d_design_post.as<double>(); bb.rows_post = c.rows_post;
  bb.sky_pre = c.d_sky_pre.as<double>(); bb.sky_post = c.d_sky_post.as<double>();
  return bb;
}

// Dispatch of the templated fast pk_ell kernels: vmode 0..3 (PK_V_*), kind 0..4
template <int VM, int KD, int TW>
int launch_pk_fast_one(dim3 grid, dim3 block, size_t smem, cudaStream_t st, const vg::PkGridDev& g,
                       const vg_eval_desc* ev, int E, const int* ids, const int* twins, const int* out_slot,
                       int n_slots, const double* pf, int npf, int B, double* out) {
  static size_t smem_set = 0;
  if (smem > smem_set) {
    VG_CUDA_CHECK(cudaFuncSetAttribute(vg::pk_ell_fast_kernel<2, VM, KD, TW>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    smem_set = smem;
  }
  vg::pk_ell_fast_kernel<2, VM, KD, TW><<<grid, block, smem, st>>>(g, ev, E, ids, twins, out_slot, n_slots, pf,
                                                                 npf, B, out);
  return VG_OK;
}
#define VG_PK_ARGS grid, block, smem, st, g, ev, E, ids, twins, out_slot, n_slots, pf, npf, B, out
template <int VM, int TW>
int launch_pk_fast_kind(int kind, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const vg::PkGridDev& g,
                        const vg_eval_desc* ev, int E, const int* ids, const int* twins, const int* out_slot,
                        int n_slots, const double* pf, int npf, int B, double* out) {
  switch (kind) {
    case 0: return launch_pk_fast_one<VM, 0, TW>(VG_PK_ARGS);
    case 1: return launch_pk_fast_one<VM, 1, TW>(VG_PK_ARGS);
    case 2: return launch_pk_fast_one<VM, 2, TW>(VG_PK_ARGS);
    case 3: return launch_pk_fast_one<VM, 3, TW>(VG_PK_ARGS);
    default: return launch_pk_fast_one<VM, 4, TW>(VG_PK_ARGS);
  }
}
// twins == nullptr: one evaluation per grid.z entry; else pairs (ids[z], twins[z])
int launch_pk_fast(int vmode, int kind, dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                   const vg::PkGridDev& g, const vg_eval_desc* ev, int E, const int* ids, const int* twins,
                   const int* out_slot, int n_slots, const double* pf, int npf, int B, double* out) {
  if (twins) {
    switch (vmode) {
      case 0: return launch_pk_fast_kind<0, 1>(kind, VG_PK_ARGS);
      case 1: return launch_pk_fast_kind<1, 1>(kind, VG_PK_ARGS);
      case 2: return launch_pk_fast_kind<2, 1>(kind, VG_PK_ARGS);
      default: return launch_pk_fast_kind<3, 1>(kind, VG_PK_ARGS);
    }
  }
  switch (vmode) {
    case 0: return launch_pk_fast_kind<0, 0>(kind, VG_PK_ARGS);
    case 1: return launch_pk_fast_kind<1, 0>(kind, VG_PK_ARGS);
    case 2: return launch_pk_fast_kind<2, 0>(kind, VG_PK_ARGS);
    default: return launch_pk_fast_kind<3, 0>(kind, VG_PK_ARGS);
  }
}
#undef VG_PK_ARGS

// Device operands of the metal product for one layout: rp-only blocks [n_metal][rp_np][rp_np] or
// the dense matrices concatenated along K, mcat[i][col*ld_model + j] = M_c[i][j], both in the
// layout's column-block order.
int build_metal_operands(Corr& c, const EvalLayout& L, DevBuf& mcat, DevBuf& mrp) {
  if (c.metal_rp && L.n_metal > 0) {   // rp-only matrices of the hot evaluations, column-block order
    const size_t nn = (size_t)c.rp_np * c.rp_np;
    std::vector<double> host((size_t)L.n_metal * nn, 0.0);
    for (int s = 0; s < c.n_metal; ++s) {
      const int col = L.metal_map[s];
      if (col < 0) continue;
      double* dst = host.data() + (size_t)col * nn;
      if (s < (int)c.metal_rp_host.size() && !c.metal_rp_host[s].empty())
        std::copy(c.metal_rp_host[s].begin(), c.metal_rp_host[s].end(), dst);
      else
        for (int i = 0; i < c.rp_np; ++i) dst[(size_t)i * c.rp_np + i] = 1.0;
    }
    VG_TRY(upload(mrp, host.data(), host.size()));
  }
  // metal matrices of the hot evaluations, concatenated along K: mcat[i][c*ld_model + j] = M_c[i][j]
  if (c.metal_dense && L.n_metal > 0) {
    const size_t ldk = (size_t)L.n_metal * c.ld_model;
    VG_TRY(mcat.alloc((size_t)c.n_model * ldk * 8));
    std::vector<double> eye;
    for (int s = 0; s < c.n_metal; ++s) {
      const int col = L.metal_map[s];
      if (col < 0) continue;
      const double* src;
      if (s < (int)c.metal_host.size() && !c.metal_host[s].empty()) {
        src = c.metal_host[s].data();
      } else {
        if (eye.empty()) {
          eye.assign((size_t)c.n_model * c.n_model, 0.0);
          for (int i = 0; i < c.n_model; ++i) eye[(size_t)i * c.n_model + i] = 1.0;
        }
        src = eye.data();
      }
      VG_CUDA_CHECK(cudaMemcpy2D(mcat.as<double>() + (size_t)col * c.ld_model, ldk * 8, src,
                                 (size_t)c.n_model * 8, (size_t)c.n_model * 8, c.n_model,
                                 cudaMemcpyHostToDevice));
    }
  }
  return VG_OK;
}

int launch_metal_rp(Corr& c, const double* xmet, int n_metal, const double* A, long long B, double* met,
                    cudaStream_t st) {
  // the DMMA tile at every batch size (measured on C3: one vector 0.30 -> 0.17 ms, 31 vectors
  // 0.31 -> 0.18 ms, 451 vectors 2.27 -> 0.89 ms against the FMA tile, which stays reachable
  // through VG_RP_DMMA_MIN for comparison)
  static const long long dmma_min = [] {
    const char* e = std::getenv("VG_RP_DMMA_MIN");
    return e ? std::atoll(e) : 1LL;
  }();
  if (B >= dmma_min) {
    dim3 grid(vg_cdiv(c.rp_nt, vg::MRD_TN), vg_cdiv(c.rp_np, vg::MRD_TM), (unsigned)B);
    vg::metal_rp_dmma_kernel<<<grid, vg::MRD_THREADS, 0, st>>>(xmet, n_metal, c.ld_model, A, c.rp_np, c.rp_nt, met);
  } else {
    dim3 grid(vg_cdiv(c.rp_nt, vg::MRP_TN), vg_cdiv(c.rp_np, vg::MRP_TM), (unsigned)B);
    vg::metal_rp_kernel<<<grid, 256, 0, st>>>(xmet, n_metal, c.ld_model, A, c.rp_np, c.rp_nt, met);
  }
  VG_LAUNCHED();
  return VG_OK;
}

// Stages K1..K3 + metals + pre-distortion assembly of one correlation: leaves
// x[Bc][ld_model] in plan->w_x.  `upto` stops early for the per-stage taps.
// `fused` (nullable): the caller wants the chi2 path's residual; when the correlation has the fused
// distortion x metal operator the metal product is left in plan->w_mety [Bc][ld_mask] on the MASKED
// DISTORTED bins (not added to x) and *fused is set -- run_corr_back adds it after the distortion.
int run_corr_front(vg_plan* plan, Corr& c, long long Bc, int* status_dev, cudaStream_t st, int upto = 99,
                   int layout = -1, int unit_bias = 0, bool* fused = nullptr) {
  const Hankel& h = *plan->hankels[c.hankel_id];
  const int E = (int)c.evals.size();
  vg::PkGridDev g = plan->grid_dev();
  // the per-stage taps (upto <= 2) expose every evaluation's multipoles / knot tables, so
  // they run all evaluations through the per-point pass
  if (layout < 0)
    layout = plan->direct_mode ? LAY_DIRECT : (upto <= 2 ? LAY_ALL
                                                 : (plan->use_moments ? ((plan->fold_metals && upto > 3) ? LAY_FOLD : LAY_MOM)
                                                                      : LAY_FULL));
  const bool direct = layout == LAY_DIRECT;
  const EvalLayout& L = c.lay[layout];
  // velocity-dispersion tables active: their evaluations leave the per-point pass and the Hankel
  // GEMM (slots n_full_hot..) and get rows Em.. of the moment-table region
  const bool vt = plan->vt_now && c.vt_n > 0 && (layout == LAY_MOM || layout == LAY_FOLD) && !L.vt_ids.empty();
  const PkLists& PL = vt ? L.pk_hot : L.pk;
  const int Ef = vt ? L.n_full_hot : L.n_full;
  const int Em = (int)L.mom_ids.size();
  const int Ev = vt ? (int)L.vt_ids.size() : 0;
  const int rowT = 2 * h.n_out_pad;
  const long long row_len = (long long)plan->n_ell * rowT;
  double* tables_full = plan->w_tables.as<double>();
  double* tables_mom = tables_full + Bc * Ef * row_len;
  const int* out_slot = L.d_slot.as<int>();
  if (Ef > 0) {  // K1 + K2a
    ProfScope prof(plan, PROF_PK, st);
    const vg_eval_desc* ev = c.d_evals.as<vg_eval_desc>();
    double* out = plan->w_pkell.as<double>();
    const double* pf = plan->w_pfull.as<double>();
    // fast kernels: fixed CTA shape (2 mu groups x 4 vector groups x 2 vectors)
    const bool fast_ok = Bc >= 8 && plan->nmu % (vg::PK_FAST_MG * vg::PK_FAST_U) == 0;
    const int NB = Bc >= 2 ? 2 : 1;
    const int groups = vg_cdiv(Bc, NB);
    const int EG = groups >= 4 ? 4 : (groups >= 2 ? 2 : 1);
    const int MG = vg::PK_GROUPS / EG;
    const size_t fast_smem = vg::pk_fast_smem_bytes(plan->nmu);
    for (int m = 0; m < PK_NCLS; ++m) {
      const std::vector<int>& class_ids = PL.class_ids[m];
      if (class_ids.empty()) continue;
      const int* ids = PL.d_class_ids[m].as<int>();
      if (m < 20 && fast_ok) {
        dim3 block(vg::PK_KLANES, vg::PK_FAST_MG, vg::PK_FAST_EG);
        if (PL.n_single[m] > 0) {
          dim3 grid(vg_cdiv(Bc, 2 * vg::PK_FAST_EG), vg_cdiv(plan->nk, vg::PK_KLANES), (unsigned)PL.n_single[m]);
          VG_TRY(launch_pk_fast(m / 5, m % 5, grid, block, fast_smem, st, g, ev, E, ids, nullptr, out_slot, Ef,
                                pf, plan->npf, (int)Bc, out));
        }
        if (!PL.pair_ids[m].empty()) {      // one vector, two evaluations per thread
          if (PL.n_single[m] > 0) VG_LAUNCHED();
          dim3 grid(vg_cdiv(Bc, vg::PK_FAST_EG), vg_cdiv(plan->nk, vg::PK_KLANES), (unsigned)PL.pair_ids[m].size());
          VG_TRY(launch_pk_fast(m / 5, m % 5, grid, block, fast_smem, st, g, ev, E, PL.d_pair_ids[m].as<int>(),
                                PL.d_pair_twins[m].as<int>(), out_slot, Ef, pf, plan->npf, (int)Bc, out));
        }
      } else {
        dim3 block(vg::PK_KLANES, MG, EG);
        dim3 grid(vg_cdiv(Bc, NB * EG), vg_cdiv(plan->nk, vg::PK_KLANES), (unsigned)class_ids.size());
        const int mode = m < 20 ? (m < 5 ? vg::PK_MODE_STATIC : vg::PK_MODE_VELO) : m - 20;
#define VG_PK_GEN(NBV, MODEV) \
        vg::pk_ell_kernel<NBV, MODEV><<<grid, block, 0, st>>>(g, ev, E, ids, out_slot, Ef, pf, plan->npf, (int)Bc, out, status_dev)
        if (NB == 2) {
          if (mode == 0) VG_PK_GEN(2, vg::PK_MODE_STATIC);
          else if (mode == 1) VG_PK_GEN(2, vg::PK_MODE_VELO);
          else VG_PK_GEN(2, vg::PK_MODE_GENERIC);
        } else {
          if (mode == 0) VG_PK_GEN(1, vg::PK_MODE_STATIC);
          else if (mode == 1) VG_PK_GEN(1, vg::PK_MODE_VELO);
          else VG_PK_GEN(1, vg::PK_MODE_GENERIC);
        }
#undef VG_PK_GEN
      }
      VG_LAUNCHED();
    }
  }
  if (upto <= 1) return VG_OK;
  // K2b: Hankel + spline operator, one GEMM batched over the multipoles
  if (Ef > 0) {
    ProfScope prof(plan, PROF_HANKEL, st);
    VG_TRY(gemm_auto(plan, plan->w_pkell.as<double>(), (long long)plan->n_ell * plan->nk_pad,
                     h.op.as<double>(), plan->nk_pad, tables_full, row_len, (int)(Bc * Ef), rowT,
                     plan->nk_pad, plan->n_ell, plan->nk_pad, (long long)rowT * plan->nk_pad, rowT, st));
  }
  if (Em > 0) {  // moment form: knot tables as a 3/9-term combination of static tables
    ProfScope prof(plan, PROF_MOMENT, st);
    const long long stretches = vg_cdiv(row_len, 256 * vg::MOM_XPT);
    const int x_iters = vg::moment_x_iters(stretches, (long long)Em * vg_cdiv(Bc, vg::MOM_NB));
    dim3 grid((unsigned)vg_cdiv(stretches, x_iters), (unsigned)Em, (unsigned)vg_cdiv(Bc, vg::MOM_NB));
    vg::moment_tables_kernel<<<grid, 256, 0, st>>>(c.d_evals.as<vg_eval_desc>(), L.d_mom_ids.as<int>(), Em,
                                                  plan->w_pfull.as<double>(), plan->npf, (int)Bc,
                                                  c.d_mom_ktab.as<double>(), (int)row_len, Em + Ev, tables_mom, x_iters);
    VG_LAUNCHED();
  }
  if (Ev > 0) {  // velocity-dispersion tables: weights, then one GEMM per evaluation
    ProfScope prof(plan, PROF_VTAB, st);
    int ldw = 0;
    for (int nm : L.vt_nm) ldw = std::max(ldw, pad16((long long)c.vt_n * nm));
    VG_TRY(plan->w_vtw.ensure((size_t)Ev * (size_t)std::max(plan->chunk, Bc) * ldw * 8));
    double* W = plan->w_vtw.as<double>();
    dim3 wgrid((unsigned)Bc, (unsigned)Ev);
    vg::vtab_weights_kernel<<<wgrid, 128, 0, st>>>(c.d_evals.as<vg_eval_desc>(), L.d_vt_ids.as<int>(),
                                                  L.d_vt_terms.as<vg::VtTerms>(), plan->w_pfull.as<double>(), plan->npf,
                                                  (int)Bc, c.vt_map, c.vt_n, ldw, W);
    VG_LAUNCHED();
    for (int i = 0; i < Ev; ++i) {
      vg::GemmArgs ga{};
      ga.A = W + (size_t)i * Bc * ldw; ga.lda = ldw;
      ga.B = c.vt_tab[i].as<double>(); ga.ldb = (long long)vg::VT_NODES * L.vt_nm[i];
      ga.C = tables_mom + (long long)(Em + i) * row_len; ga.ldc = (long long)(Em + Ev) * row_len;
      ga.M = (int)Bc; ga.N = (int)row_len; ga.K = pad16((long long)c.vt_n * L.vt_nm[i]);
      ga.batch = 1; ga.splits = 1; ga.alpha = 1.0; ga.beta = 0.0;
      VG_TRY(vg::launch_dgemm(ga, vg::GEMM_STORE, st));
    }
  }
  if (upto <= 2) return VG_OK;
  if (!L.gather_ids.empty()) {  // K3
    ProfScope prof(plan, PROF_XI, st);
    GatherJob job{};
    job.ids = L.d_gather_ids.as<int>(); job.n_ids = (int)L.gather_ids.size();
    job.tab_slot = (vt ? L.d_tabslot_vt : L.d_tabslot).as<int>();
    job.metal_map = L.d_metal_map.as<int>(); job.n_metal = L.n_metal;
    job.pfull = plan->w_pfull.as<double>(); job.B = Bc;
    job.tables = tables_full; job.n_full = Ef;
    job.tables_mom = tables_mom; job.n_mom = Em + Ev; job.mom_stride_b = (Em + Ev) * row_len; job.mom_stride_j = row_len;
    job.unit_bias = unit_bias;
    job.core = plan->w_core.as<double>(); job.xmet = plan->w_xmet.as<double>();
    job.status = status_dev;
    const int n_all = job.n_ids, n_fast = fast_gather_enabled() ? L.n_gather_fast : 0;
    if (n_fast > 0) {
      job.n_ids = n_fast; job.fast = 1;
      VG_TRY(launch_xi_gather(plan, c, job, st));
    }
    if (n_all > n_fast) {
      job.ids = L.d_gather_ids.as<int>() + n_fast; job.n_ids = n_all - n_fast; job.fast = 0;
      VG_TRY(launch_xi_gather(plan, c, job, st));
    }
  }
  if (!L.mxi_ids.empty()) {  // xi stage in moment form
    ProfScope prof(plan, PROF_MXI, st);
    const int n_mxi = (int)L.mxi_ids.size();
    const long long stretches = vg_cdiv(c.ld_model, 256);
    const int x_iters = vg::moment_x_iters(stretches, (long long)n_mxi * vg_cdiv(Bc, vg::MXI_NB));
    dim3 grid((unsigned)vg_cdiv(stretches, x_iters), (unsigned)n_mxi, (unsigned)vg_cdiv(Bc, vg::MXI_NB));
    vg::moment_xi_kernel<<<grid, 256, 0, st>>>(c.d_evals.as<vg_eval_desc>(), L.d_mxi_ids.as<int>(),
                                              L.d_mxi_owner.as<int>(), plan->w_pfull.as<double>(), plan->npf,
                                              (int)Bc, c.d_mxi_tab.as<double>(),
                                              L.mxi_rows.empty() ? nullptr : L.d_mxi_rows.as<int>(), c.ld_model,
                                              plan->w_core.as<double>(), plan->w_xmet.as<double>(),
                                              L.d_metal_map.as<int>(), std::max(L.n_metal, 1), x_iters);
    VG_LAUNCHED();
  }
  if (upto <= 3) return VG_OK;
  // metals: sum_c w_c M_c xi_c  (vega/metals.py:258-367)
  const double* met = nullptr;
  const double* met_fold = nullptr;
  if (!L.fold_ids.empty()) {
    ProfScope prof(plan, PROF_ELEM, st);
    const int nf = (int)L.fold_ids.size();
    const size_t smem = (size_t)nf * vg::MOM_MAX * vg::MXI_NB * 8 + (size_t)nf * sizeof(int);
    static size_t fold_smem_set = 48 * 1024;
    if (smem > fold_smem_set) {
      VG_CUDA_CHECK(cudaFuncSetAttribute(vg::moment_fold_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
      fold_smem_set = smem;
    }
    dim3 grid(vg_cdiv(c.ld_model, 256), (unsigned)vg_cdiv(Bc, vg::MXI_NB));
    vg::moment_fold_kernel<<<grid, 256, smem, st>>>(c.d_evals.as<vg_eval_desc>(), L.d_fold_ids.as<int>(),
                                                   L.d_fold_owner.as<int>(), nf, plan->w_pfull.as<double>(),
                                                   plan->npf, (int)Bc, c.d_fold_tab.as<double>(), c.ld_model,
                                                   plan->w_fold.as<double>());
    VG_LAUNCHED();
    met_fold = plan->w_fold.as<double>();
  }
  // a correlation with nothing folded keeps one set of matrices
  const bool fold_layout = layout == LAY_FOLD && !L.fold_ids.empty();
  int oz_S = !c.metal_dense ? 0 : (plan->ozaki_slices < 0 ? c.oz_auto_slices : plan->ozaki_slices);
  // chi2 path with the fused operator: N = n_mask distorted bins instead of n_model model bins
  const bool fuse = fused && plan->fuse_dm_metals && c.d_dmcat.p && !fold_layout && L.n_metal > 0 &&
                    (layout == LAY_MOM || layout == LAY_FULL) && upto > 3 && !direct;
  if (fused) *fused = fuse;
  if (fuse) oz_S = plan->ozaki_slices < 0 ? c.oz_fused_slices : (c.oz_fused_slices > 0 ? plan->ozaki_slices : 0);
  const int metal_rows = fuse ? c.n_mask : c.n_model;            // rows of the operator = columns of the product
  const int metal_ldo = fuse ? c.ld_mask : c.ld_model;
  double* metal_out = fuse ? plan->w_mety.as<double>() : plan->w_met.as<double>();
  if (L.n_metal > 0) {
    const bool use_ozaki = c.metal_dense && oz_S > 0 && Bc >= plan->ozaki_min_batch;
    // (the split-integer branch times its two parts itself: PROF_NONE keeps this scope silent)
    ProfScope prof(plan, use_ozaki ? PROF_NONE : (c.metal_rp ? PROF_METAL_RP : (c.metal_dense ? PROF_METAL : PROF_ELEM)), st);
    if (c.metal_rp) {
      VG_TRY(launch_metal_rp(c, plan->w_xmet.as<double>(), L.n_metal,
                             (fold_layout ? c.d_mrp_fold : c.d_mrp).as<double>(), Bc, plan->w_met.as<double>(), st));
    } else if (c.metal_dense) {
      const long long ldk = (long long)L.n_metal * c.ld_model;
      const DevBuf& mcat = fuse ? c.d_dmcat : (fold_layout ? c.d_mcat_fold : c.d_mcat);
      if (use_ozaki) {
        // split-integer GEMM on tcgen05 (vg_ozaki.cuh): the operator's digit planes are built on
        // first use, the batch operand's on every call
        OzakiOperand& ob = fuse ? c.oz_dmcat : (fold_layout ? c.oz_mcat_fold : c.oz_mcat);
        if (ob.S != oz_S || ob.src != mcat.as<double>()) {
          VG_TRY(ozaki_slice(ob, mcat.as<double>(), ldk, metal_rows, metal_rows, (int)ldk, oz_S, st));
          ob.src = mcat.as<double>();
        }
        {
          ProfScope slice_prof(plan, PROF_OZ_SLICE, st);
          VG_TRY(ozaki_slice(plan->oz_A, plan->w_xmet.as<double>(), ldk, Bc, plan->chunk, (int)ldk, oz_S, st));
        }
        {
          ProfScope gemm_prof(plan, PROF_METAL, st);
          VG_TRY(ozaki_gemm(plan->oz_A, Bc, ob, metal_rows, metal_out, metal_ldo, plan->w_split,
                            plan->d_oz_error, st));
        }
        vg::flag_status_kernel<<<vg_cdiv(Bc, 256), 256, 0, st>>>(plan->d_oz_error.as<int>(), status_dev, Bc);
        VG_LAUNCHED();
      } else {
        VG_TRY(gemm_auto(plan, plan->w_xmet.as<double>(), ldk, mcat.as<double>(), ldk,
                         metal_out, metal_ldo, (int)Bc, metal_rows, (int)ldk, 1, 0, 0, 0, st));
      }
    } else {
      const long long n = Bc * c.ld_model;
      vg::metal_sum_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(plan->w_xmet.as<double>(), L.n_metal, c.ld_model,
                                                         c.n_model, Bc, plan->w_met.as<double>());
      VG_LAUNCHED();
    }
    met = fuse ? nullptr : plan->w_met.as<double>();
  }
  vg::CombineArgs ca{};
  ca.core = plan->w_core.as<double>(); ca.met = met; ca.met_fold = met_fold;
  ca.inst = c.has_inst ? c.d_inst.as<double>() : nullptr;
  ca.idx_inst_amp = c.idx_inst_amp; ca.inst_amp_default = c.inst_amp_default;
  ca.idx_bao_amp = plan->idx_bao_amp;
  ca.bb = bb_dev(c);
  ca.pfull = plan->w_pfull.as<double>(); ca.npf = plan->npf;
  ca.n = c.n_model; ca.ld = c.ld_model; ca.B = Bc;
  ca.x = plan->w_x.as<double>();
  ca.direct = direct ? 1 : 0;
  ca.frozen = c.d_frozen.as<double>(); ca.n_frozen = direct ? 0 : (int)c.frozen_ids.size();
  ca.frozen_ids = c.d_frozen_ids.as<int>(); ca.evals = c.d_evals.as<vg_eval_desc>();
  const long long n = Bc * c.ld_model;
  ProfScope prof(plan, PROF_ELEM, st);
  vg::combine_pre_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(ca);
  VG_LAUNCHED();
  return VG_OK;
}

}  // namespace

namespace {

// ---- velocity-dispersion tables (vg_moments.cuh) ------------------------------------------------
constexpr long long VT_MIN_BATCH = 1;      // every batch size: a single vector reads ~15 MB of coefficients per evaluation
                                           // instead of walking the 814 000-point grid (C3, one vector: pk_ell was 0.41 of 1.02 ms)

// Marks the evaluations of a correlation that velocity-dispersion tables can serve: the moment form
// applies to everything but ONE sampled velocity dispersion, which is a theta column with finite
// [sample] limits 0 <= lo < hi (vg_plan_set_theta_limits); all of them read the same dispersion.
void flag_velo_table_candidates(vg_plan* plan, Corr& c) {
  const int E = (int)c.evals.size();
  c.vt_flag.assign(E, 0);
  c.vt_n = 0; c.vt_col = c.vt_tcol = -1; c.vt_err = 0.0;
  c.vt_tab.clear();
  if ((int)plan->theta_lo.size() != plan->n_theta || plan->n_theta == 0 || plan->has_theta_transform) return;
  if (plan->n_ell != vg::PK_MAX_ELL) return;
  for (int e = 0; e < E; ++e) {
    const vg_eval_desc& d = c.evals[e];
    if (d.table_owner >= 0 || d.frozen || d.direct || !vg::vt_eligible(d)) continue;
    const int col = vg::vt_sigma_column(d);
    if (c.vt_col >= 0 && col != c.vt_col) continue;
    int tcol = -1;
    for (int j = 0; j < plan->n_theta; ++j)
      if (plan->theta_cols_host[j] == col) tcol = j;
    if (tcol < 0) continue;
    const double lo = plan->theta_lo[tcol], hi = plan->theta_hi[tcol];
    if (!(std::isfinite(lo) && std::isfinite(hi) && lo >= 0.0 && hi > lo)) continue;
    if (c.vt_col < 0) {
      c.vt_col = col; c.vt_tcol = tcol;
      c.vt_map.lo = lo; c.vt_map.hi = hi; c.vt_map.eps = vg::VT_EPS;
      c.vt_map.tau_lo = vg::vt_tau(lo, vg::VT_EPS); c.vt_map.tau_hi = vg::vt_tau(hi, vg::VT_EPS);
    }
    c.vt_flag[e] = 1;
  }
  for (int e = 0; e < E; ++e) {   // twins go together or not at all
    const int t = c.evals[e].pk_twin;
    if (t >= 0 && t < E && c.vt_flag[e] != c.vt_flag[t]) c.vt_flag[e] = c.vt_flag[t] = 0;
  }
}

// Chebyshev coefficient tables of the flagged evaluations, then the order count: the smallest of the
// ladder whose xi-stage output matches the per-point pass on VT_NTEST values of sigma to VT_TOL
// (relative to the largest bin of each output slot).  No order count passing clears the flags.
int build_velo_tables(vg_plan* plan, Corr& c, cudaStream_t st) {
  const EvalLayout& LM = c.lay[LAY_MOM];
  const Hankel& h = *plan->hankels[c.hankel_id];
  const int nv = (int)LM.vt_ids.size();
  const int rowT = 2 * h.n_out_pad;
  const long long row_len = (long long)plan->n_ell * rowT;
  const int NQ = vg::VT_NODES;
  {
    std::vector<double> sig(NQ);
    for (int q = 0; q < NQ; ++q) sig[q] = vg::vt_sigma_of(std::cos(M_PI * (q + 0.5) / NQ), c.vt_map);
    DevBuf d_sig, mom, node;
    VG_TRY(upload(d_sig, sig.data(), sig.size()));
    VG_TRY(mom.alloc((size_t)NQ * vg::MOM_MAX * plan->n_ell * plan->nk_pad * 8));
    VG_TRY(node.alloc((size_t)NQ * vg::MOM_MAX * row_len * 8));
    c.vt_tab.clear();
    c.vt_tab.resize(nv);
    vg::PkGridDev g = plan->grid_dev();
    for (int i = 0; i < nv; ++i) {
      const int nm = LM.vt_nm[i];
      dim3 ggrid(vg_cdiv(plan->nk, 128), (unsigned)NQ, 1);
      vg::vtab_grid_kernel<<<ggrid, 128, 0, st>>>(g, c.d_evals.as<vg_eval_desc>(), LM.d_vt_ids.as<int>() + i,
                                                 plan->d_base.as<double>(), d_sig.as<double>(), NQ, mom.as<double>());
      VG_LAUNCHED();
      VG_TRY(gemm_auto(plan, mom.as<double>(), (long long)plan->n_ell * plan->nk_pad, h.op.as<double>(),
                       plan->nk_pad, node.as<double>(), row_len, NQ * vg::MOM_MAX, rowT, plan->nk_pad,
                       plan->n_ell, plan->nk_pad, (long long)rowT * plan->nk_pad, rowT, st));
      VG_TRY(c.vt_tab[i].alloc((size_t)row_len * NQ * nm * 8));
      dim3 cgrid(vg_cdiv(row_len, 128), (unsigned)NQ);
      vg::vtab_cheb_kernel<<<cgrid, 128, 0, st>>>(node.as<double>(), NQ, LM.vt_terms[i], (int)row_len, c.vt_tab[i].as<double>());
      VG_LAUNCHED();
    }
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
  }
  // ---- check against the per-point pass ----
  const int NT = vg::VT_NTEST;
  static const double t_test[vg::VT_NTEST] = {-0.97, -0.71, -0.33, 0.05, 0.41, 0.68, 0.88, 0.985};
  VG_TRY(ensure_workspace(plan, NT));
  std::vector<double> pf((size_t)NT * plan->npf);
  for (int i = 0; i < NT; ++i) {
    std::copy(plan->base.begin(), plan->base.end(), pf.begin() + (size_t)i * plan->npf);
    pf[(size_t)i * plan->npf + c.vt_col] = vg::vt_sigma_of(t_test[i], c.vt_map);
  }
  VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, pf.data(), pf.size() * 8, cudaMemcpyHostToDevice, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, NT * 4, st));
  const bool vt_now0 = plan->vt_now;
  const size_t n_core = (size_t)2 * NT * c.ld_model, n_xmet = (size_t)NT * std::max(LM.n_metal, 1) * c.ld_model;
  DevBuf ref_core, ref_xmet, d_cmp;
  VG_TRY(ref_core.alloc(n_core * 8));
  VG_TRY(ref_xmet.alloc(n_xmet * 8));
  VG_TRY(d_cmp.alloc(2 * sizeof(unsigned long long)));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_core.p, 0, n_core * 8, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_xmet.p, 0, n_xmet * 8, st));
  plan->vt_now = false;
  int rc = run_corr_front(plan, c, NT, plan->w_status.as<int>(), st, 3, LAY_MOM);
  if (rc == VG_OK) {
    cudaMemcpyAsync(ref_core.p, plan->w_core.p, n_core * 8, cudaMemcpyDeviceToDevice, st);
    cudaMemcpyAsync(ref_xmet.p, plan->w_xmet.p, n_xmet * 8, cudaMemcpyDeviceToDevice, st);
  }
  // (slot, first element, row stride, rows) of every output of the xi stage
  struct Out { const double* ref; const double* got; long long stride; };
  std::vector<Out> outs;
  for (int sl = 0; sl < 2; ++sl)
    outs.push_back({ref_core.as<double>() + (size_t)sl * NT * c.ld_model,
                    plan->w_core.as<double>() + (size_t)sl * NT * c.ld_model, c.ld_model});
  for (int m = 0; m < LM.n_metal; ++m)
    outs.push_back({ref_xmet.as<double>() + (size_t)m * c.ld_model, plan->w_xmet.as<double>() + (size_t)m * c.ld_model,
                    (long long)LM.n_metal * c.ld_model});
  static const int ladder[] = {8, 12, 16, 24, 32, 40, 48, 56, 64};
  int chosen = 0;
  double chosen_err = 0.0;
  for (int n : ladder) {
    if (rc != VG_OK) break;
    c.vt_n = n;
    plan->vt_now = true;
    rc = run_corr_front(plan, c, NT, plan->w_status.as<int>(), st, 3, LAY_MOM);
    if (rc != VG_OK) break;
    double err = 0.0;
    for (const Out& o : outs) {
      VG_CUDA_CHECK(cudaMemsetAsync(d_cmp.p, 0, 2 * sizeof(unsigned long long), st));
      dim3 cg(vg_cdiv(c.n_model, 256), NT);
      vg::vtab_compare_kernel<<<cg, 256, 0, st>>>(o.got, o.ref, c.n_model, o.stride, d_cmp.as<unsigned long long>());
      VG_LAUNCHED();
      double r[2];
      VG_CUDA_CHECK(cudaMemcpyAsync(r, d_cmp.p, sizeof(r), cudaMemcpyDeviceToHost, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));
      if (r[1] > 0.0) err = std::max(err, r[0] / r[1]);
      else if (r[0] > 0.0) err = 1.0;
    }
    if (err <= vg::VT_TOL) { chosen = n; chosen_err = err; break; }
  }
  plan->vt_now = vt_now0;
  VG_TRY(rc);
  c.vt_n = chosen;
  c.vt_err = chosen_err;
  if (chosen == 0) {     // the tables do not reach the tolerance: these evaluations keep the per-point pass
    std::fill(c.vt_flag.begin(), c.vt_flag.end(), 0);
    c.vt_tab.clear();
    VG_TRY(build_layouts(plan, c));
  }
  return VG_OK;
}

// Are this call's vectors inside the tabulated range of every correlation?  Sets plan->vt_now.
// theta on the host is checked there; theta on the device costs one small kernel and a 4-byte
// read-back per CALL (not per chunk).
int decide_velo_tables(vg_plan* plan, const double* theta_dev, const double* theta_host, long long B, cudaStream_t st) {
  plan->vt_now = false;
  if (!plan->velo_tables || !plan->use_moments || plan->direct_mode || B < VT_MIN_BATCH) return VG_OK;
  bool any = false;
  for (auto& cp : plan->corrs) any = any || cp->vt_n > 0;
  if (!any) return VG_OK;
  if (theta_host) {
    for (auto& cp : plan->corrs) {
      if (cp->vt_n == 0) continue;
      for (long long b = 0; b < B; ++b) {
        const double v = theta_host[b * plan->n_theta + cp->vt_tcol];
        if (!(v >= cp->vt_map.lo && v <= cp->vt_map.hi)) return VG_OK;
      }
    }
    plan->vt_now = true;
    return VG_OK;
  }
  if (!plan->d_vt_flag.p) VG_TRY(plan->d_vt_flag.alloc(sizeof(int)));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->d_vt_flag.p, 0, sizeof(int), st));
  for (auto& cp : plan->corrs) {
    if (cp->vt_n == 0) continue;
    vg::vtab_range_kernel<<<vg_cdiv(B, 256), 256, 0, st>>>(theta_dev, plan->n_theta, cp->vt_tcol, B, cp->vt_map.lo,
                                                        cp->vt_map.hi, plan->d_vt_flag.as<int>());
    VG_LAUNCHED();
  }
  int flag = 1;
  VG_CUDA_CHECK(cudaMemcpyAsync(&flag, plan->d_vt_flag.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  plan->vt_now = flag == 0;
  return VG_OK;
}

// Digits of the split-integer metal GEMM for one correlation: the fewest S whose truncation bound,
// evaluated with the row statistics of the operator and of the metal xi at the PLAN parameters,
//   2^(-7S) [ max|a| max_j ||b_j||_1 + ||a||_1 max|b| + (S-1) max|a| max|b| max_j nnz(b_j) ]
// (digits beyond S of A, of B, and the dropped digit products -- only where both are non-zero)
// stays under OZAKI_BUDGET * max|xi_smooth|: half of the 1e-10 parity bar.  The bound is a worst
// case (every dropped digit product at its largest magnitude and with the same sign); the errors
// measured on the dense eBOSS-shaped workload are 1e-11 max|xi| at S = 6 and below 1e-12 at S = 7,
// where this bound evaluates to 2e-9 and 2e-11.
constexpr double OZAKI_BUDGET = 5e-11;
int ozaki_budget(vg_plan* plan, Corr& c, cudaStream_t st) {
  const EvalLayout& L = c.lay[LAY_MOM];
  c.oz_auto_slices = 0;
  if (!c.metal_dense || L.n_metal == 0) return VG_OK;
  VG_TRY(ensure_workspace(plan, 1));
  VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, plan->d_base.p, (size_t)plan->npf * 8,
                                cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, 4, st));
  VG_TRY(run_corr_front(plan, c, 1, plan->w_status.as<int>(), st, 3, LAY_MOM, 0));
  const long long ldk = (long long)L.n_metal * c.ld_model;
  DevBuf stats;
  VG_TRY(stats.alloc(9 * sizeof(unsigned long long)));
  unsigned long long* d = stats.as<unsigned long long>();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_xmet.as<double>(), ldk, (int)ldk, d);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<(unsigned)c.n_model, 256, 0, st>>>(c.d_mcat.as<double>(), ldk, (int)ldk, d + 3);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_core.as<double>() + c.ld_model, c.ld_model, c.n_model, d + 6);
  VG_LAUNCHED();
  double h[9];
  VG_CUDA_CHECK(cudaMemcpyAsync(h, stats.p, sizeof(h), cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  const double amax = h[0], a1 = h[1], bmax = h[3], b1 = h[4], bnz = h[5], scale = h[6];
  c.oz_auto_slices = vg::OZ_MAX_S;
  for (int S = vg::OZ_MAX_S; S >= 3; --S) {
    const double bound = std::ldexp(amax * b1 + a1 * bmax + (S - 1) * amax * bmax * bnz, -7 * S);
    c.oz_bound[S] = scale > 0.0 ? bound / scale : 0.0;
    if (scale > 0.0 && bound <= OZAKI_BUDGET * scale) c.oz_auto_slices = S;
  }
  if (!(scale > 0.0) || !(amax > 0.0)) c.oz_auto_slices = 0;    // nothing to size the digits against: DMMA
  return VG_OK;
}

// The same for the distortion matrix product y = D x (vega/model.py:142-143): x the model before
// distortion at the plan parameters, the bound measured against max|y|.
int ozaki_budget_dm(vg_plan* plan, Corr& c, cudaStream_t st) {
  c.oz_dm_slices = 0;
  for (double& b : c.oz_dm_bound) b = 0.0;
  if (!c.has_dmat) return VG_OK;
  VG_TRY(ensure_workspace(plan, 1));
  VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, plan->d_base.p, (size_t)plan->npf * 8,
                                cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, 4, st));
  VG_TRY(run_corr_front(plan, c, 1, plan->w_status.as<int>(), st));
  VG_TRY(gemm_auto(plan, plan->w_x.as<double>(), c.ld_model, c.d_dmat.as<double>(), c.ld_model,
                   plan->w_y.as<double>(), c.ld_dist, 1, c.n_dist, c.ld_model, 1, 0, 0, 0, st));
  DevBuf stats;
  VG_TRY(stats.alloc(9 * sizeof(unsigned long long)));
  unsigned long long* d = stats.as<unsigned long long>();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_x.as<double>(), c.ld_model, c.n_model, d);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<(unsigned)c.n_dist, 256, 0, st>>>(c.d_dmat.as<double>(), c.ld_model, c.n_model, d + 3);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_y.as<double>(), c.ld_dist, c.n_dist, d + 6);
  VG_LAUNCHED();
  double h[9];
  VG_CUDA_CHECK(cudaMemcpyAsync(h, stats.p, sizeof(h), cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  const double amax = h[0], a1 = h[1], bmax = h[3], b1 = h[4], bnz = h[5], scale = h[6];
  if (!(scale > 0.0) || !(amax > 0.0) || !std::isfinite(scale)) return VG_OK;
  c.oz_dm_slices = vg::OZ_MAX_S;
  for (int S = vg::OZ_MAX_S; S >= 3; --S) {
    const double bound = std::ldexp(amax * b1 + a1 * bmax + (S - 1) * amax * bmax * bnz, -7 * S);
    c.oz_dm_bound[S] = bound / scale;
    if (bound <= OZAKI_BUDGET * scale) c.oz_dm_slices = S;
  }
  // the largest digit count that still misses the budget: keep the FP64 kernel
  if (c.oz_dm_bound[vg::OZ_MAX_S] > OZAKI_BUDGET) c.oz_dm_slices = 0;
  return VG_OK;
}

// rows [n][n_mask] (row stride n_mask) -> out[n][n_mask] = rows W^T, i.e. W applied to every row
int whiten_rows(vg_plan* plan, Corr& c, const double* rows, int n, DevBuf& out, cudaStream_t st) {
  DevBuf in_pad, out_pad;
  VG_TRY(in_pad.alloc((size_t)n * c.ld_mask * 8));
  VG_TRY(out_pad.alloc((size_t)n * c.ld_mask * 8));
  VG_CUDA_CHECK(cudaMemcpy2DAsync(in_pad.p, (size_t)c.ld_mask * 8, rows, (size_t)c.n_mask * 8, (size_t)c.n_mask * 8, n,
                                  cudaMemcpyDeviceToDevice, st));
  VG_TRY(gemm_auto(plan, in_pad.as<double>(), c.ld_mask, c.d_white.as<double>(), c.ld_mask, out_pad.as<double>(),
                   c.ld_mask, n, c.n_mask, c.ld_mask, 1, 0, 0, 0, st));
  VG_TRY(out.alloc((size_t)n * c.n_mask * 8));
  VG_CUDA_CHECK(cudaMemcpy2DAsync(out.p, (size_t)c.n_mask * 8, out_pad.p, (size_t)c.ld_mask * 8, (size_t)c.n_mask * 8, n,
                                  cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  return VG_OK;
}

// Whitened chi2 path (VG_OPT_WHITEN).  With C^-1 = W^T W (vg_corr_set_whitening: the Cholesky factor
// the caller took of the matrix it gave to vg_corr_set_data),
//   chi2 = (d - y)^T C^-1 (d - y) = |W d - W y|^2        (vega/vega_interface.py:294-319)
// and y = D_mask x is linear, so W folds into the operators at plan time: d_wd = W D_mask, the fused
// distortion x metal operator is built on d_wd, the data and the mocks are stored as W d.  The chi2
// GEMM (2 n_mask^2 flops per vector) becomes a sum of squares.  Needs a per-correlation covariance
// (under a global one W couples the correlations and W D would have sum(n_mask) rows per
// correlation), a distortion matrix and no post-distortion broadband term (it sits between D and W).
// Called with the correlation's state changed (finalize, new data / covariance / mocks / option);
// rebuilds the fused operator too, which depends on the choice.
int rewhiten_data(vg_plan* plan, Corr& c, cudaStream_t st) {
  c.wdata_dirty = false;
  if (!c.whitened) return VG_OK;
  VG_TRY(whiten_rows(plan, c, c.d_data.as<double>(), 1, c.d_wdata, st));
  if (c.n_mocks > 0) VG_TRY(whiten_rows(plan, c, c.d_mocks.as<double>(), c.n_mocks, c.d_wmocks, st));
  return VG_OK;
}

int rebuild_whitened(vg_plan* plan, Corr& c, cudaStream_t st) {
  c.white_dirty = false;
  c.wdata_dirty = false;
  bool ok = plan->whiten && c.has_white && c.has_invcov && c.has_dmat && !plan->has_global && c.n_mask > 0 &&
            c.d_dmat_masked.p != nullptr;
  for (const vg_bb_term& t : c.bb_terms)
    if (t.position >= 2) ok = false;
  const bool was = c.whitened;
  c.whitened = false;
  c.oz_wd_slices = 0;
  for (double& b : c.oz_wd_bound) b = 0.0;
  if (ok) {
    {   // d_wd = W D_mask: C[r][i] = sum_s W[r][s] D[s][i] -> the TN kernel wants D^T
      DevBuf dt;
      VG_TRY(dt.alloc((size_t)c.ld_model * c.ld_mask * 8));
      dim3 tgrid(vg_cdiv(c.n_model, 32), vg_cdiv(c.n_mask, 32));
      vg::transpose_kernel<<<tgrid, dim3(32, 8), 0, st>>>(c.d_dmat_masked.as<double>(), c.ld_model, c.n_mask, c.n_model,
                                                          dt.as<double>(), c.ld_mask);
      VG_LAUNCHED();
      VG_TRY(c.d_wd.alloc((size_t)c.n_mask * c.ld_model * 8));
      VG_TRY(gemm_auto(plan, c.d_white.as<double>(), c.ld_mask, dt.as<double>(), c.ld_mask, c.d_wd.as<double>(),
                       c.ld_model, c.n_mask, c.n_model, c.ld_mask, 1, 0, 0, 0, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));
    }
    VG_TRY(whiten_rows(plan, c, c.d_data.as<double>(), 1, c.d_wdata, st));
    if (c.n_mocks > 0) VG_TRY(whiten_rows(plan, c, c.d_mocks.as<double>(), c.n_mocks, c.d_wmocks, st));
    c.whitened = true;
    // digits of the split-integer GEMM on d_wd: the bound of ozaki_budget_dm against max|W D x|
    VG_TRY(ensure_workspace(plan, 1));
    VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, plan->d_base.p, (size_t)plan->npf * 8, cudaMemcpyDeviceToDevice, st));
    VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, 4, st));
    const bool vt0 = plan->vt_now;
    plan->vt_now = false;
    VG_TRY(run_corr_front(plan, c, 1, plan->w_status.as<int>(), st));
    plan->vt_now = vt0;
    VG_TRY(gemm_auto(plan, plan->w_x.as<double>(), c.ld_model, c.d_wd.as<double>(), c.ld_model,
                     plan->w_y.as<double>(), c.ld_mask, 1, c.n_mask, c.ld_model, 1, 0, 0, 0, st));
    DevBuf stats;
    VG_TRY(stats.alloc(9 * sizeof(unsigned long long)));
    unsigned long long* d = stats.as<unsigned long long>();
    vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_x.as<double>(), c.ld_model, c.n_model, d);
    VG_LAUNCHED();
    vg::ozaki_stats_kernel<<<(unsigned)c.n_mask, 256, 0, st>>>(c.d_wd.as<double>(), c.ld_model, c.n_model, d + 3);
    VG_LAUNCHED();
    vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_y.as<double>(), c.ld_mask, c.n_mask, d + 6);
    VG_LAUNCHED();
    double h[9];
    VG_CUDA_CHECK(cudaMemcpyAsync(h, stats.p, sizeof(h), cudaMemcpyDeviceToHost, st));
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
    const double amax = h[0], a1 = h[1], bmax = h[3], b1 = h[4], bnz = h[5], scale = h[6];
    if (scale > 0.0 && amax > 0.0 && std::isfinite(scale)) {
      c.oz_wd_slices = vg::OZ_MAX_S;
      for (int S = vg::OZ_MAX_S; S >= 3; --S) {
        const double bound = std::ldexp(amax * b1 + a1 * bmax + (S - 1) * amax * bmax * bnz, -7 * S);
        c.oz_wd_bound[S] = bound / scale;
        if (bound <= OZAKI_BUDGET * scale) c.oz_wd_slices = S;
      }
      if (c.oz_wd_bound[vg::OZ_MAX_S] > OZAKI_BUDGET) c.oz_wd_slices = 0;
    }
  } else {
    c.d_wd.release(); c.d_wdata.release(); c.d_wmocks.release();
  }
  (void)was;
  return build_fused_metal_operator(plan, c, st);
}

// Fused operator of the chi2 path: (masked rows of the distortion matrix) x (metal matrices).
// vega/model.py:142-143 applies D to xi_core + sum_c M_c xi_c; the chi2 only reads the n_mask bins
// the cuts keep, and D (M_c xi_c) = (D M_c) xi_c, so with
//   DM[r][c*ld + j] = sum_i D[mask r][i] M_c[i][j]          (one DGEMM per metal matrix, here)
// the per-vector metal product has n_mask output columns instead of n_model (64 % of them with the
// eBOSS cuts) and the distortion product is left with the core terms.  Exact algebra (rounding of
// one more FP64 product at plan time); needs dense metal matrices, a distortion matrix and NO
// multiplicative pre-distortion broadband term (it sits between the two products).  Digits of the
// split-integer GEMM from the same worst-case bound as ozaki_budget, against max|D xi|.
int build_fused_metal_operator(vg_plan* plan, Corr& c, cudaStream_t st) {
  c.d_dmcat.release();
  c.oz_dmcat.planes.release(); c.oz_dmcat.exps.release();
  c.oz_dmcat.rows = 0; c.oz_dmcat.Kp = 0; c.oz_dmcat.S = 0; c.oz_dmcat.src = nullptr;
  c.oz_fused_slices = 0;
  for (double& b : c.oz_fused_bound) b = 0.0;
  const EvalLayout& L = c.lay[LAY_MOM];
  if (!c.has_dmat || !c.metal_dense || c.metal_rp || c.n_mask <= 0 || L.n_metal == 0 || !c.d_mcat.p) return VG_OK;
  for (const vg_bb_term& t : c.bb_terms)
    if (t.position == 0) return VG_OK;
  const long long ldk = (long long)L.n_metal * c.ld_model;
  const double* dmasked = (c.whitened ? c.d_wd : c.d_dmat_masked).as<double>();   // W folded in (rebuild_whitened)
  VG_TRY(c.d_dmcat.alloc((size_t)c.n_mask * ldk * 8));
  {
    DevBuf mt, prod;   // M_c^T [n_model][ld_model]; D_masked M_c [n_mask][ld_model] (gemm_auto's split-K
                       // reduction owns every column of its output rows: a compact block, copied afterwards)
    VG_TRY(mt.alloc((size_t)c.ld_model * c.ld_model * 8));
    VG_TRY(prod.alloc((size_t)c.n_mask * c.ld_model * 8));
    for (int col = 0; col < L.n_metal; ++col) {
      dim3 tgrid(vg_cdiv(c.n_model, 32), vg_cdiv(c.n_model, 32));
      vg::transpose_kernel<<<tgrid, dim3(32, 8), 0, st>>>(c.d_mcat.as<double>() + (size_t)col * c.ld_model, ldk,
                                                          c.n_model, c.n_model, mt.as<double>(), c.ld_model);
      VG_LAUNCHED();
      VG_TRY(gemm_auto(plan, dmasked, c.ld_model, mt.as<double>(), c.ld_model,
                       prod.as<double>(), c.ld_model, c.n_mask, c.n_model, c.ld_model, 1, 0, 0, 0, st));
      VG_CUDA_CHECK(cudaMemcpy2DAsync(c.d_dmcat.as<double>() + (size_t)col * c.ld_model, (size_t)ldk * 8,
                                      prod.as<double>(), (size_t)c.ld_model * 8, (size_t)c.n_model * 8, c.n_mask,
                                      cudaMemcpyDeviceToDevice, st));
    }
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
  }
  // digits: operator row statistics of DM, the metal xi and max|D x| at the plan parameters
  VG_TRY(ensure_workspace(plan, 1));
  VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, plan->d_base.p, (size_t)plan->npf * 8, cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, 4, st));
  VG_TRY(run_corr_front(plan, c, 1, plan->w_status.as<int>(), st, 99, LAY_MOM, 0));
  VG_TRY(gemm_auto(plan, plan->w_x.as<double>(), c.ld_model, dmasked, c.ld_model,
                   plan->w_y.as<double>(), c.ld_mask, 1, c.n_mask, c.ld_model, 1, 0, 0, 0, st));
  DevBuf stats;
  VG_TRY(stats.alloc(9 * sizeof(unsigned long long)));
  unsigned long long* d = stats.as<unsigned long long>();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_xmet.as<double>(), ldk, (int)ldk, d);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<(unsigned)c.n_mask, 256, 0, st>>>(c.d_dmcat.as<double>(), ldk, (int)ldk, d + 3);
  VG_LAUNCHED();
  vg::ozaki_stats_kernel<<<1, 256, 0, st>>>(plan->w_y.as<double>(), c.ld_mask, c.n_mask, d + 6);
  VG_LAUNCHED();
  double h[9];
  VG_CUDA_CHECK(cudaMemcpyAsync(h, stats.p, sizeof(h), cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  const double amax = h[0], a1 = h[1], bmax = h[3], b1 = h[4], bnz = h[5], scale = h[6];
  if (!(scale > 0.0) || !(amax > 0.0) || !std::isfinite(scale)) return VG_OK;     // DMMA on the fused operator
  c.oz_fused_slices = vg::OZ_MAX_S;
  for (int S = vg::OZ_MAX_S; S >= 3; --S) {
    const double bound = std::ldexp(amax * b1 + a1 * bmax + (S - 1) * amax * bmax * bnz, -7 * S);
    c.oz_fused_bound[S] = bound / scale;
    if (bound <= OZAKI_BUDGET * scale) c.oz_fused_slices = S;
  }
  if (c.oz_fused_bound[vg::OZ_MAX_S] > OZAKI_BUDGET) c.oz_fused_slices = 0;
  return VG_OK;
}

// fast_metals metal x metal terms (vega/metals.py:144-169): M_c xi_c evaluated ONCE at the
// plan's base parameters (the reference caches the first evaluation forever), stored without
// the bias product, which the hot path applies per vector.
int freeze_metal_terms(vg_plan* plan, Corr& c, cudaStream_t st) {
  if (c.frozen_ids.empty()) return VG_OK;
  VG_TRY(ensure_workspace(plan, 1));
  VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_pfull.p, plan->d_base.p, (size_t)plan->npf * 8,
                                cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaMemsetAsync(plan->w_status.p, 0, 4, st));
  VG_TRY(run_corr_front(plan, c, 1, plan->w_status.as<int>(), st, 3, LAY_ALL, 1));
  const EvalLayout& L = c.lay[LAY_ALL];
  const int nf = (int)c.frozen_ids.size();
  VG_TRY(c.d_frozen.alloc((size_t)nf * c.ld_model * 8));
  DevBuf mat;
  for (int f = 0; f < nf; ++f) {
    const int slot = c.evals[c.frozen_ids[f]].out_slot - 2;
    const double* xi = plan->w_xmet.as<double>() + (size_t)L.metal_map[slot] * c.ld_model;
    double* dst = c.d_frozen.as<double>() + (size_t)f * c.ld_model;
    if (c.metal_rp && slot < (int)c.metal_rp_host.size() && !c.metal_rp_host[slot].empty()) {
      VG_TRY(upload(mat, c.metal_rp_host[slot].data(), c.metal_rp_host[slot].size()));
      VG_TRY(launch_metal_rp(c, xi, 1, mat.as<double>(), 1, dst, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));
    } else if (c.metal_dense && slot < (int)c.metal_host.size() && !c.metal_host[slot].empty()) {
      VG_TRY(upload_padded(mat, c.metal_host[slot].data(), c.n_model, c.n_model, c.ld_model));
      VG_TRY(gemm_auto(plan, xi, c.ld_model, mat.as<double>(), c.ld_model, dst, c.ld_model, 1, c.n_model,
                       c.ld_model, 1, 0, 0, 0, st));
      VG_CUDA_CHECK(cudaStreamSynchronize(st));   // `mat` is re-used by the next term
    } else {
      VG_CUDA_CHECK(cudaMemcpyAsync(dst, xi, (size_t)c.ld_model * 8, cudaMemcpyDeviceToDevice, st));
    }
  }
  int bad = 0;
  VG_CUDA_CHECK(cudaMemcpyAsync(&bad, plan->w_status.p, 4, cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  plan->static_bad |= bad;
  return VG_OK;
}

// Distortion matrix + post-distortion broadband of one correlation.
//   data == false: model rows for ALL distorted bins -> out[Bc][ldo = ld_dist]
//   data == true : residual d - m on the masked bins  -> out[Bc][ldo] at column col0
//   y_add (residual only): the metal product already on the masked distorted bins (fused operator,
//   run_corr_front), [Bc][ld_mask], added to D x before the post-distortion broadband
int run_corr_back(vg_plan* plan, Corr& c, long long Bc, bool residual, double* out, int ldo,
                  int ncols, cudaStream_t st, const int* mock_idx = nullptr, int* status_dev = nullptr,
                  const double* y_add = nullptr) {
  const double* y = plan->w_x.as<double>();
  int ldy = c.ld_model;
  int y_is_masked = 0;
  if (c.has_dmat) {
    const bool white = residual && c.whitened;       // W folded into the masked rows (rebuild_whitened)
    const DevBuf& dm = residual ? (white ? c.d_wd : c.d_dmat_masked) : c.d_dmat;
    const int rows = residual ? c.n_mask : c.n_dist;
    ldy = residual ? c.ld_mask : c.ld_dist;
    y_is_masked = residual ? 1 : 0;
    const int auto_S = white ? c.oz_wd_slices : c.oz_dm_slices;
    const int oz_S = plan->ozaki_slices < 0 ? auto_S : (auto_S > 0 ? plan->ozaki_slices : 0);
    if (oz_S > 0 && plan->ozaki_dm && Bc >= plan->ozaki_min_batch) {
      // split-integer GEMM on tcgen05, as for the dense metal matrices (run_corr_front)
      OzakiOperand& ob = residual ? (white ? c.oz_wd : c.oz_dmat_masked) : c.oz_dmat;
      if (ob.S != oz_S || ob.src != dm.as<double>()) {
        VG_TRY(ozaki_slice(ob, dm.as<double>(), c.ld_model, rows, rows, c.n_model, oz_S, st));
        ob.src = dm.as<double>();
      }
      {
        ProfScope slice_prof(plan, PROF_OZ_SLICE, st);
        VG_TRY(ozaki_slice(plan->oz_A, plan->w_x.as<double>(), c.ld_model, Bc, plan->chunk, c.n_model, oz_S, st));
      }
      {
        ProfScope gemm_prof(plan, PROF_DM, st);
        VG_TRY(ozaki_gemm(plan->oz_A, Bc, ob, rows, plan->w_y.as<double>(), ldy, plan->w_split, plan->d_oz_error, st));
      }
      if (status_dev) {
        vg::flag_status_kernel<<<vg_cdiv(Bc, 256), 256, 0, st>>>(plan->d_oz_error.as<int>(), status_dev, Bc);
        VG_LAUNCHED();
      }
    } else {
      ProfScope prof(plan, PROF_DM, st);
      VG_TRY(gemm_auto(plan, plan->w_x.as<double>(), c.ld_model, dm.as<double>(), c.ld_model,
                       plan->w_y.as<double>(), ldy, (int)Bc, rows, c.ld_model, 1, 0, 0, 0, st));
    }
    y = plan->w_y.as<double>();
  }
  vg::PostArgs pa{};
  pa.y = y; pa.ldy = ldy; pa.y_is_masked = y_is_masked;
  pa.y_add = y_add;
  pa.mask_idx = residual ? c.d_mask_idx.as<int>() : nullptr;
  pa.bb = bb_dev(c); pa.n_dist = c.n_dist;
  pa.pfull = plan->w_pfull.as<double>(); pa.npf = plan->npf; pa.idx_bao_amp = plan->idx_bao_amp;
  pa.direct = plan->direct_mode ? 1 : 0;
  pa.data = residual ? (c.whitened ? c.d_wdata : c.d_data).as<double>() : nullptr;
  if (residual && mock_idx) {
    VG_REQUIRE(c.n_mocks > 0, "vg_chi2_batch_mc: correlation without mocks (vg_corr_set_mocks)");
    pa.data = (c.whitened ? c.d_wmocks : c.d_mocks).as<double>(); pa.mock_idx = mock_idx; pa.data_stride = c.n_mask;
  }
  pa.n_out = residual ? c.n_mask : c.n_dist;
  pa.ldo = ldo; pa.ncols = ncols; pa.B = Bc; pa.out = out;
  const long long n = Bc * ncols;
  ProfScope prof(plan, PROF_ELEM, st);
  vg::post_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(pa);
  VG_LAUNCHED();
  return VG_OK;
}

// chi2[b] (+)= diff[b]^T invcov diff[b]
int quad_form(vg_plan* plan, const double* diff, int n, int ld, const double* invcov, long long Bc,
              double* chi2, int accumulate, cudaStream_t st) {
  ProfScope prof(plan, invcov ? PROF_CHI2 : PROF_ELEM, st);
  if (!invcov) {
    vg::chi2_identity_kernel<<<(unsigned)Bc, 256, 0, st>>>(diff, n, ld, chi2, accumulate);
    VG_LAUNCHED();
    return VG_OK;
  }
  vg::GemmArgs g{};
  g.A = diff; g.B = invcov; g.M = (int)Bc; g.N = n; g.K = ld; g.lda = ld; g.ldb = ld;
  g.batch = 1; g.alpha = 1.0; g.beta = 0.0;
  g.sym = 1;      // invcov is symmetric (symmetrize() at upload): one triangle is read
  const int n_tiles0 = vg_cdiv(n, vg::GEMM_BN);
  const int bm = vg::gemm_tile_m((int)Bc);
  g.splits = pick_splits((long long)vg_cdiv(Bc, bm) * n_tiles0, ld / vg::GEMM_BK, bm == vg::GEMM_BM ? 1 : 2);
  const int n_tiles = n_tiles0 * g.splits;
  VG_TRY(plan->w_partial.ensure((size_t)Bc * n_tiles * 8));
  g.C = plan->w_partial.as<double>(); g.ldc = n_tiles;
  g.D = diff; g.ldd = ld;
  VG_TRY(vg::launch_dgemm(g, vg::GEMM_ROWDOT, st));
  vg::chi2_partials_kernel<<<vg_cdiv(Bc, 256), 256, 0, st>>>(plan->w_partial.as<double>(), n_tiles, Bc, chi2, accumulate);
  VG_LAUNCHED();
  return VG_OK;
}

// ---- second lane of the chi2 pass (vg_plan::Lane) ----
long long lane_max_batch() {      // largest chunk that is spread over lanes (VG_LANE_MAX_B in the environment)
  static const long long v = [] {
    const char* e = std::getenv("VG_LANE_MAX_B");
    const long long x = e ? std::atoll(e) : 1024;
    return std::max(1LL, std::min(x, 4096LL));
  }();
  return v;
}
bool lanes_enabled() {
  static const bool on = [] { const char* e = std::getenv("VG_LANES"); return !(e && e[0] == '0'); }();
  return on;
}
void swap_lane(vg_plan* plan, vg_plan::Lane& l) {
  plan->w_pkell.swap(l.w_pkell); plan->w_tables.swap(l.w_tables); plan->w_core.swap(l.w_core);
  plan->w_xmet.swap(l.w_xmet); plan->w_met.swap(l.w_met); plan->w_x.swap(l.w_x); plan->w_y.swap(l.w_y);
  plan->w_partial.swap(l.w_partial); plan->w_split.swap(l.w_split); plan->w_fold.swap(l.w_fold);
  plan->w_mety.swap(l.w_mety); plan->w_vtw.swap(l.w_vtw);
  plan->oz_A.planes.swap(l.oz_A.planes); plan->oz_A.exps.swap(l.oz_A.exps);
  std::swap(plan->oz_A.rows, l.oz_A.rows); std::swap(plan->oz_A.Kp, l.oz_A.Kp);
  std::swap(plan->oz_A.S, l.oz_A.S); std::swap(plan->oz_A.src, l.oz_A.src);
  if (!plan->has_global) plan->w_diff.swap(l.w_diff);     // (a global covariance: one residual row for all)
}
struct LaneSwap {       // a lane's workspace stands in for the plan's while its launches are issued
  vg_plan* plan; vg_plan::Lane* lane;
  LaneSwap(vg_plan* p, vg_plan::Lane* l) : plan(p), lane(l) { if (lane) swap_lane(plan, *lane); }
  ~LaneSwap() { if (lane) swap_lane(plan, *lane); }
};
int ensure_lane(vg_plan* plan, vg_plan::Lane& l, long long Bc_) {
  if (!l.stream) {
    VG_CUDA_CHECK(cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking));
    VG_CUDA_CHECK(cudaEventCreateWithFlags(&l.fork, cudaEventDisableTiming));
    VG_CUDA_CHECK(cudaEventCreateWithFlags(&l.join, cudaEventDisableTiming));
  }
  if (Bc_ <= l.chunk) return VG_OK;
  long long cap = 64;                   // sized in a few steps: 64, 128, ... vectors
  while (cap < Bc_) cap *= 2;
  const size_t Bc = (size_t)cap;
  size_t pkell = 0, tables = 0, core = 0, xmet = 0, met = 0, y = 0;
  for (auto& cp : plan->corrs) {
    const Corr& c = *cp;
    const Hankel& h = *plan->hankels[c.hankel_id];
    const size_t E = c.evals.size();
    pkell = std::max(pkell, (size_t)Bc * E * plan->n_ell * plan->nk_pad);
    tables = std::max(tables, (size_t)Bc * E * plan->n_ell * 2 * h.n_out_pad);
    core = std::max(core, (size_t)2 * Bc * c.ld_model);
    xmet = std::max(xmet, (size_t)Bc * std::max(c.n_metal, 1) * c.ld_model);
    met = std::max(met, (size_t)Bc * c.ld_model);
    y = std::max(y, (size_t)Bc * std::max(c.ld_dist, c.ld_mask));
  }
  VG_TRY(l.w_pkell.alloc(pkell * 8));
  VG_TRY(l.w_tables.alloc(tables * 8));
  VG_TRY(l.w_core.alloc(core * 8));
  VG_TRY(l.w_xmet.alloc(xmet * 8));
  VG_TRY(l.w_met.alloc(met * 8));
  VG_TRY(l.w_fold.alloc(met * 8));
  VG_TRY(l.w_x.alloc(met * 8));
  VG_TRY(l.w_y.alloc(y * 8));
  VG_TRY(l.w_mety.alloc(y * 8));
  VG_TRY(l.w_diff.alloc(y * 8));
  VG_TRY(l.w_partial.alloc((size_t)Bc * 256 * 8));
  VG_TRY(l.chi2.alloc((size_t)Bc * 8));
  l.chunk = cap;
  return VG_OK;
}

int chi2_chunk(vg_plan* plan, const double* theta_dev, long long Bc, double* chi2_dev, int* status_dev,
               cudaStream_t st, const int* mock_idx = nullptr) {
  // operands that follow the data / covariance / mocks / options (never inside a graph capture: a
  // state change drops the graphs and their call counts, so the first call after it runs here)
  for (auto& cp : plan->corrs) {
    if (cp->white_dirty && cp->n_mask > 0) VG_TRY(rebuild_whitened(plan, *cp, st));
    else if (cp->wdata_dirty) VG_TRY(rewhiten_data(plan, *cp, st));
  }
  VG_TRY(expand_params(plan, theta_dev, Bc, status_dev, st));
  // thin batches: the correlations spread over lanes
  const bool laned = plan->two_lanes && lanes_enabled() && !plan->profile && plan->corrs.size() > 1 &&
                     Bc <= lane_max_batch();
  const int n_lanes = laned ? (int)std::min<size_t>(1 + vg_plan::EXTRA_LANES, plan->corrs.size()) : 1;
  for (int l = 1; l < n_lanes; ++l) {
    vg_plan::Lane& lane = plan->lanes[l - 1];
    VG_TRY(ensure_lane(plan, lane, Bc));
    VG_CUDA_CHECK(cudaEventRecord(lane.fork, st));
    VG_CUDA_CHECK(cudaStreamWaitEvent(lane.stream, lane.fork, 0));
  }
  int first[1 + vg_plan::EXTRA_LANES] = {1, 1, 1, 1};
  int col0 = 0;
  int rc = VG_OK;
  for (size_t ci = 0; ci < plan->corrs.size() && rc == VG_OK; ++ci) {
    Corr& c = *plan->corrs[ci];
    const int alt = (int)(ci % n_lanes);
    vg_plan::Lane* lane = alt ? &plan->lanes[alt - 1] : nullptr;
    cudaStream_t s = alt ? lane->stream : st;
    LaneSwap guard(plan, lane);
    rc = [&]() -> int {
      VG_REQUIRE(c.n_mask > 0, "vg_chi2_batch: correlation without data (vg_corr_set_data)");
      bool fused = false;
      VG_TRY(run_corr_front(plan, c, Bc, status_dev, s, 99, -1, 0, &fused));
      const double* y_add = fused ? plan->w_mety.as<double>() : nullptr;
      if (plan->has_global) {
        // residuals of all correlations side by side: [Bc][ld_total]
        VG_TRY(run_corr_back(plan, c, Bc, true, plan->w_diff.as<double>() + col0, plan->ld_total,
                             c.n_mask, s, mock_idx, status_dev, y_add));
        col0 += c.n_mask;
      } else {
        VG_TRY(run_corr_back(plan, c, Bc, true, plan->w_diff.as<double>(), c.ld_mask, c.ld_mask, s, mock_idx,
                             status_dev, y_add));
        VG_TRY(quad_form(plan, plan->w_diff.as<double>(), c.n_mask, c.ld_mask,
                         (c.has_invcov && !c.whitened) ? c.d_invcov.as<double>() : nullptr, Bc,
                         alt ? lane->chi2.as<double>() : chi2_dev, !first[alt], s));
        first[alt] = 0;
      }
      return VG_OK;
    }();
  }
  for (int l = 1; l < n_lanes; ++l) {   // (joined on the error path too: a capture must not end with a stream forked)
    cudaEventRecord(plan->lanes[l - 1].join, plan->lanes[l - 1].stream);
    cudaStreamWaitEvent(st, plan->lanes[l - 1].join, 0);
  }
  VG_TRY(rc);
  for (int l = 1; l < n_lanes && !plan->has_global; ++l) {
    vg::add_vector_kernel<<<vg_cdiv(Bc, 256), 256, 0, st>>>(plan->lanes[l - 1].chi2.as<double>(), Bc, chi2_dev);
    VG_LAUNCHED();
  }
  if (plan->has_global)
    VG_TRY(quad_form(plan, plan->w_diff.as<double>(), plan->n_total, plan->ld_total,
                     plan->d_global_invcov.as<double>(), Bc, chi2_dev, 0, st));
  if (plan->n_prior > 0) {
    vg::prior_kernel<<<vg_cdiv(Bc, 256), 256, 0, st>>>(plan->w_pfull.as<double>(), plan->npf,
                                                    plan->d_prior_cols.as<int>(), plan->d_prior_mean.as<double>(),
                                                    plan->d_prior_sigma.as<double>(), plan->n_prior, Bc, chi2_dev);
    VG_LAUNCHED();
  }
  return VG_OK;
}

// chi2_chunk through a CUDA graph when the batch is small: the pass of a single vector or of a
// gradient stencil is 50-100 launches of a few microseconds each, so launch latency, not the
// device, sets its time.  The first two calls of a (B, options) key run normally (allocations and
// function attributes settle), the third is captured on an internal stream, later ones replay.
constexpr long long GRAPH_MAX_B = 96;
bool graphs_enabled() {
  static const bool on = [] { const char* e = std::getenv("VG_GRAPHS"); return !(e && e[0] == '0'); }();
  return on;
}
int chi2_chunk_small(vg_plan* plan, const double* theta_dev, long long Bc, double* chi2_dev, int* status_dev,
                     cudaStream_t st, const int* mock_idx = nullptr) {
  if (Bc > GRAPH_MAX_B || plan->profile || !graphs_enabled())
    return chi2_chunk(plan, theta_dev, Bc, chi2_dev, status_dev, st, mock_idx);
  const long long key = Bc | ((long long)plan->use_moments << 20) | ((long long)plan->fold_metals << 21) |
                        ((long long)plan->direct_mode << 22) | ((long long)(mock_idx != nullptr) << 23) |
                        ((long long)(plan->ozaki_slices + 1) << 24) | ((long long)plan->ozaki_dm << 30) |
                        ((long long)plan->vt_now << 31) | ((long long)plan->fuse_dm_metals << 32) |
                        ((long long)plan->whiten << 33);
  vg_plan::GraphEntry& g = plan->graphs[key];
  if (g.exec && g.generation != g_alloc_generation) {     // a buffer moved since the capture
    cudaGraphExecDestroy(g.exec);
    g.exec = nullptr;
  }
  if (!g.exec && g.seen < 2) {
    ++g.seen;
    return chi2_chunk(plan, theta_dev, Bc, chi2_dev, status_dev, st, mock_idx);
  }
  // fixed staging buffers: the graph is independent of the caller's pointers
  double* th = plan->w_theta.as<double>();
  double* c2 = plan->w_chi2.as<double>();
  int* stt = plan->w_status.as<int>();
  if (plan->n_theta > 0 && theta_dev != th)
    VG_CUDA_CHECK(cudaMemcpyAsync(th, theta_dev, (size_t)Bc * plan->n_theta * 8, cudaMemcpyDeviceToDevice, st));
  if (mock_idx) {
    VG_TRY(plan->w_mockidx.ensure((size_t)GRAPH_MAX_B * sizeof(int)));
    VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_mockidx.p, mock_idx, (size_t)Bc * sizeof(int), cudaMemcpyDeviceToDevice, st));
  }
  if (!g.exec) {
    if (!plan->gstream) VG_CUDA_CHECK(cudaStreamCreateWithFlags(&plan->gstream, cudaStreamNonBlocking));
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
    const long long before = g_launches + vg::gemm_launches();
    VG_CUDA_CHECK(cudaStreamBeginCapture(plan->gstream, cudaStreamCaptureModeRelaxed));
    const int rc = chi2_chunk(plan, th, Bc, c2, stt, plan->gstream, mock_idx ? plan->w_mockidx.as<int>() : nullptr);
    cudaGraph_t graph = nullptr;
    const cudaError_t end = cudaStreamEndCapture(plan->gstream, &graph);
    if (rc != VG_OK || end != cudaSuccess || !graph) {
      if (graph) cudaGraphDestroy(graph);
      cudaGetLastError();
      g.seen = -1000000;                      // never try this key again
      return chi2_chunk(plan, theta_dev, Bc, chi2_dev, status_dev, st, mock_idx);
    }
    g.launches = g_launches + vg::gemm_launches() - before;
    g.generation = g_alloc_generation;
    const cudaError_t inst = cudaGraphInstantiate(&g.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (inst != cudaSuccess) {
      g.exec = nullptr; g.seen = -1000000;
      cudaGetLastError();
      return chi2_chunk(plan, theta_dev, Bc, chi2_dev, status_dev, st, mock_idx);
    }
  } else {
    g_launches += g.launches;
  }
  VG_CUDA_CHECK(cudaGraphLaunch(g.exec, st));
  ++plan->graph_replays;
  if (chi2_dev != c2) VG_CUDA_CHECK(cudaMemcpyAsync(chi2_dev, c2, (size_t)Bc * 8, cudaMemcpyDeviceToDevice, st));
  if (status_dev != stt) VG_CUDA_CHECK(cudaMemcpyAsync(status_dev, stt, (size_t)Bc * 4, cudaMemcpyDeviceToDevice, st));
  return VG_OK;
}

}  // namespace

extern "C" {

int vg_model_batch(vg_plan* plan, const double* theta_dev, int64_t B, double* const* xi_out_dev,
                   int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && xi_out_dev && status_dev, "vg_model_batch: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const long long chunk = std::min<long long>(B, plan->chunk_max);
  VG_TRY(ensure_workspace(plan, chunk));
  VG_TRY(decide_velo_tables(plan, theta_dev, nullptr, B, st));
  for (long long b0 = 0; b0 < B; b0 += chunk) {
    const long long Bc = std::min<long long>(chunk, B - b0);
    VG_TRY(expand_params(plan, theta_dev + b0 * plan->n_theta, Bc, status_dev + b0, st));
    for (size_t ci = 0; ci < plan->corrs.size(); ++ci) {
      Corr& c = *plan->corrs[ci];
      VG_TRY(run_corr_front(plan, c, Bc, status_dev + b0, st));
      // model rows straight into the caller's [B][n_dist] array
      VG_TRY(run_corr_back(plan, c, Bc, false, xi_out_dev[ci] + b0 * c.n_dist, c.n_dist, c.n_dist, st, nullptr,
                           status_dev + b0));
    }
  }
  return VG_OK;
}

int vg_chi2_batch(vg_plan* plan, const double* theta_dev, int64_t B, double* chi2_dev,
                  int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && chi2_dev && status_dev, "vg_chi2_batch: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const long long chunk = std::min<long long>(B, plan->chunk_max);
  VG_TRY(ensure_workspace(plan, chunk));
  VG_TRY(decide_velo_tables(plan, theta_dev, nullptr, B, st));
  for (long long b0 = 0; b0 < B; b0 += chunk) {
    const long long Bc = std::min<long long>(chunk, B - b0);
    VG_TRY(chi2_chunk_small(plan, theta_dev + b0 * plan->n_theta, Bc, chi2_dev + b0, status_dev + b0, st));
  }
  return VG_OK;
}

int vg_corr_set_mocks(vg_plan* plan, int32_t corr, int32_t n_mocks, const double* mocks) {
  VG_CORR(plan, corr);
  VG_REQUIRE(n_mocks > 0 && mocks && c.n_mask > 0, "vg_corr_set_mocks: set the data mask first");
  plan->drop_graphs();
  VG_TRY(upload(c.d_mocks, mocks, (size_t)n_mocks * c.n_mask));
  c.n_mocks = n_mocks;
  c.wdata_dirty = true;
  return VG_OK;
}

int vg_chi2_batch_mc(vg_plan* plan, const double* theta_dev, const int32_t* mock_idx_dev, int64_t B,
                     double* chi2_dev, int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && chi2_dev && status_dev && mock_idx_dev,
             "vg_chi2_batch_mc: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const long long chunk = std::min<long long>(B, plan->chunk_max);
  VG_TRY(ensure_workspace(plan, chunk));
  VG_TRY(decide_velo_tables(plan, theta_dev, nullptr, B, st));
  for (long long b0 = 0; b0 < B; b0 += chunk) {
    const long long Bc = std::min<long long>(chunk, B - b0);
    VG_TRY(chi2_chunk_small(plan, theta_dev + b0 * plan->n_theta, Bc, chi2_dev + b0, status_dev + b0, st,
                      mock_idx_dev + b0));
  }
  return VG_OK;
}

int vg_chi2_batch_host(vg_plan* plan, const double* theta_host, int64_t B, double* chi2_host,
                       int32_t* status_host, void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && theta_host && chi2_host && status_host,
             "vg_chi2_batch_host: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  const long long chunk = std::min<long long>(B, plan->chunk_max);
  VG_TRY(ensure_workspace(plan, chunk));
  const int nt = std::max(plan->n_theta, 1);
  const size_t need = (size_t)chunk * (nt * 8 + 8 + 4);
  if (need > plan->pinned_bytes) {
    if (plan->pinned) cudaFreeHost(plan->pinned);
    plan->pinned = nullptr; plan->pinned_bytes = 0;
    VG_CUDA_CHECK(cudaMallocHost(&plan->pinned, need));
    plan->pinned_bytes = need;
  }
  double* pin_theta = reinterpret_cast<double*>(plan->pinned);
  double* pin_chi2 = pin_theta + (size_t)chunk * nt;
  int32_t* pin_status = reinterpret_cast<int32_t*>(pin_chi2 + chunk);
  VG_TRY(decide_velo_tables(plan, nullptr, theta_host, B, st));
  for (long long b0 = 0; b0 < B; b0 += chunk) {
    const long long Bc = std::min<long long>(chunk, B - b0);
    if (plan->n_theta > 0) {
      memcpy(pin_theta, theta_host + b0 * plan->n_theta, (size_t)Bc * plan->n_theta * 8);
      VG_CUDA_CHECK(cudaMemcpyAsync(plan->w_theta.p, pin_theta, (size_t)Bc * plan->n_theta * 8,
                                    cudaMemcpyHostToDevice, st));
    }
    VG_TRY(chi2_chunk_small(plan, plan->w_theta.as<double>(), Bc, plan->w_chi2.as<double>(),
                      plan->w_status.as<int>(), st));
    VG_CUDA_CHECK(cudaMemcpyAsync(pin_chi2, plan->w_chi2.p, (size_t)Bc * 8, cudaMemcpyDeviceToHost, st));
    VG_CUDA_CHECK(cudaMemcpyAsync(pin_status, plan->w_status.p, (size_t)Bc * 4, cudaMemcpyDeviceToHost, st));
    VG_CUDA_CHECK(cudaStreamSynchronize(st));
    memcpy(chi2_host + b0, pin_chi2, (size_t)Bc * 8);
    memcpy(status_host + b0, pin_status, (size_t)Bc * 4);
    for (long long b = 0; b < Bc; ++b) status_host[b0 + b] |= plan->static_bad;
  }
  return VG_OK;
}

int vg_corr_n_moment(const vg_plan* plan, int32_t corr) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size() || !plan->use_moments) return 0;
  return (int)plan->corrs[corr]->lay[LAY_MOM].mom_ids.size();
}

// ---- new_metals pair histograms (init time; host pointers in, host pointers out) ----------
namespace {
int upload_pair_side(DevBuf& buf, const double* side, int n, vg::PairSide& out) {
  VG_TRY(upload(buf, side, (size_t)5 * n));
  const double* d = buf.as<double>();
  out.r_true = d; out.r_assumed = d + n; out.z_true = d + 2 * (size_t)n; out.z_assumed = d + 3 * (size_t)n;
  out.weight = d + 4 * (size_t)n; out.n = n;
  return VG_OK;
}
}  // namespace

int vg_metal_pair_ratio_range(const double* side1, int32_t n1, const double* side2, int32_t n2,
                              double* range_out) {
  VG_REQUIRE(side1 && side2 && n1 > 0 && n2 > 0 && range_out, "vg_metal_pair_ratio_range: bad arguments");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    vg_set_error("vg_metal_pair_ratio_range: no CUDA device (there is no CPU fallback)");
    return VG_ERR_CUDA;
  }
  DevBuf b1, b2, mm;
  vg::PairArgs a{};
  VG_TRY(upload_pair_side(b1, side1, n1, a.s1));
  VG_TRY(upload_pair_side(b2, side2, n2, a.s2));
  dim3 grid(vg_cdiv(n2, vg::PAIR_TJ), vg_cdiv(n1, vg::PAIR_TI));
  const size_t ctas = (size_t)grid.x * grid.y;
  VG_TRY(mm.alloc(2 * ctas * 8));
  a.minmax = mm.as<double>();
  vg::pair_ratio_range_kernel<<<grid, vg::PAIR_TJ>>>(a);
  VG_LAUNCHED();
  std::vector<double> host(2 * ctas);
  VG_CUDA_CHECK(cudaMemcpy(host.data(), mm.p, 2 * ctas * 8, cudaMemcpyDeviceToHost));
  double lo = 1e300, hi = -1e300;
  for (size_t c = 0; c < ctas; ++c) { lo = std::min(lo, host[2 * c]); hi = std::max(hi, host[2 * c + 1]); }
  range_out[0] = lo; range_out[1] = hi;
  return VG_OK;
}

int vg_metal_pair_histograms(const double* side1, int32_t n1, const double* side2, int32_t n2,
                             int32_t signed_rp, double zmin, double zmax, const double* rp_edges, int32_t nb,
                             const double* ratio_edges, int32_t nr, double* H, double* per_true,
                             double* per_assumed, double* ratio) {
  VG_REQUIRE(side1 && side2 && n1 > 0 && n2 > 0 && rp_edges && nb > 0 && H && per_true && per_assumed,
             "vg_metal_pair_histograms: bad arguments");
  VG_REQUIRE(nr == 0 || (ratio_edges && ratio), "vg_metal_pair_histograms: ratio histogram without edges");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    vg_set_error("vg_metal_pair_histograms: no CUDA device (there is no CPU fallback)");
    return VG_ERR_CUDA;
  }
  const size_t n_out = (size_t)nb * nb + 4 * (size_t)nb + nr;
  const size_t smem = n_out * 8;
  VG_REQUIRE(smem <= 200 * 1024, "vg_metal_pair_histograms: %d rp bins do not fit shared memory", nb);
  DevBuf b1, b2, be, bq, out;
  vg::PairArgs a{};
  VG_TRY(upload_pair_side(b1, side1, n1, a.s1));
  VG_TRY(upload_pair_side(b2, side2, n2, a.s2));
  VG_TRY(upload(be, rp_edges, (size_t)nb + 1));
  if (nr > 0) VG_TRY(upload(bq, ratio_edges, (size_t)nr + 1));
  VG_TRY(out.alloc(n_out * 8));   // zero filled
  a.signed_rp = signed_rp; a.zmin = zmin; a.zmax = zmax;
  a.rp_edges = be.as<double>(); a.nb = nb;
  a.ratio_edges = nr > 0 ? bq.as<double>() : nullptr; a.nr = nr;
  a.H = out.as<double>(); a.per_true = a.H + (size_t)nb * nb; a.per_assumed = a.per_true + nb;
  a.ratio = a.per_assumed + 3 * (size_t)nb;
  VG_CUDA_CHECK(cudaFuncSetAttribute(vg::pair_histogram_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(vg_cdiv(n2, vg::PAIR_TJ), vg_cdiv(n1, vg::PAIR_TI));
  vg::pair_histogram_kernel<<<grid, vg::PAIR_TJ, smem>>>(a);
  VG_LAUNCHED();
  std::vector<double> host(n_out);
  VG_CUDA_CHECK(cudaMemcpy(host.data(), out.p, n_out * 8, cudaMemcpyDeviceToHost));
  std::copy(host.begin(), host.begin() + (size_t)nb * nb, H);
  std::copy(host.begin() + (size_t)nb * nb, host.begin() + (size_t)nb * nb + nb, per_true);
  std::copy(host.begin() + (size_t)nb * nb + nb, host.begin() + (size_t)nb * nb + 4 * (size_t)nb, per_assumed);
  if (nr > 0) std::copy(host.begin() + (size_t)nb * nb + 4 * (size_t)nb, host.end(), ratio);
  return VG_OK;
}

int vg_corr_velo_tables(const vg_plan* plan, int32_t corr, int32_t* orders, double* error) {
  if (orders) *orders = 0;
  if (error) *error = 0.0;
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size() || !plan->use_moments || !plan->velo_tables) return 0;
  const Corr& c = *plan->corrs[corr];
  if (c.vt_n == 0) return 0;
  if (orders) *orders = c.vt_n;
  if (error) *error = c.vt_err;
  return (int)c.lay[LAY_MOM].vt_ids.size();
}

int vg_corr_fused_metal_operator(const vg_plan* plan, int32_t corr, int32_t* digits, double* bound_per_digits) {
  if (digits) *digits = 0;
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size() || !plan->fuse_dm_metals) return 0;
  const Corr& c = *plan->corrs[corr];
  if (!c.d_dmcat.p) return 0;
  if (digits) *digits = plan->ozaki_slices < 0 ? c.oz_fused_slices : (c.oz_fused_slices > 0 ? plan->ozaki_slices : 0);
  if (bound_per_digits)
    for (int S = 0; S <= vg::OZ_MAX_S; ++S) bound_per_digits[S] = c.oz_fused_bound[S];
  return c.n_mask;
}

int vg_corr_whitened(const vg_plan* plan, int32_t corr) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size()) return 0;
  const Corr& c = *plan->corrs[corr];
  return (c.whitened && !c.white_dirty) ? 1 : 0;     // (a dirty data vector alone keeps the operators)
}

int vg_plan_velo_tables_active(const vg_plan* plan) { return plan && plan->vt_now ? 1 : 0; }

int vg_corr_n_static_xi(const vg_plan* plan, int32_t corr) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size() || !plan->use_moments) return 0;
  return (int)plan->corrs[corr]->lay[LAY_MOM].mxi_ids.size();
}

int vg_corr_n_folded(const vg_plan* plan, int32_t corr) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size() || !plan->use_moments) return 0;
  return (int)plan->corrs[corr]->lay[LAY_FOLD].fold_ids.size();
}

int vg_corr_ozaki_dm_slices(const vg_plan* plan, int32_t corr, double* bound_per_digits) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size()) return 0;
  const Corr& c = *plan->corrs[corr];
  if (bound_per_digits)
    for (int S = 0; S <= vg::OZ_MAX_S; ++S) bound_per_digits[S] = c.oz_dm_bound[S];
  if (!plan->ozaki_dm) return 0;
  return plan->ozaki_slices < 0 ? c.oz_dm_slices : (c.oz_dm_slices > 0 ? plan->ozaki_slices : 0);
}
int vg_plan_static_status(const vg_plan* plan) { return plan ? plan->static_bad : 0; }
int64_t vg_plan_graph_replays(const vg_plan* plan) { return plan ? plan->graph_replays : 0; }

int vg_corr_ozaki_slices(const vg_plan* plan, int32_t corr, double* bound_per_digits) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size()) return 0;
  const Corr& c = *plan->corrs[corr];
  if (bound_per_digits)
    for (int S = 0; S <= vg::OZ_MAX_S; ++S) bound_per_digits[S] = c.oz_bound[S];
  return plan->ozaki_slices < 0 ? c.oz_auto_slices : (c.oz_auto_slices > 0 ? plan->ozaki_slices : 0);
}

int vg_profile_enable(vg_plan* plan, int enable) {
  VG_REQUIRE(plan, "null plan");
  plan->drop_graphs();
  plan->profile = enable != 0;
  return VG_OK;
}

int vg_profile_read(vg_plan* plan, double* ms_per_tag, int64_t* count_per_tag) {
  VG_REQUIRE(plan && ms_per_tag && count_per_tag, "vg_profile_read: bad arguments");
  VG_CUDA_CHECK(cudaDeviceSynchronize());
  for (int t = 0; t < PROF_NTAGS; ++t) { ms_per_tag[t] = 0.0; count_per_tag[t] = 0; }
  for (auto& r : plan->prof_recs) {
    float ms = 0.f;
    VG_CUDA_CHECK(cudaEventElapsedTime(&ms, r.e0, r.e1));
    ms_per_tag[r.tag] += ms;
    count_per_tag[r.tag] += 1;
    plan->prof_pool.push_back(r.e0);
    plan->prof_pool.push_back(r.e1);
  }
  plan->prof_recs.clear();
  return VG_OK;
}

int64_t vg_stage_size(const vg_plan* plan, int32_t corr, int32_t stage) {
  if (!plan || corr < 0 || corr >= (int)plan->corrs.size()) return -1;
  const Corr& c = *plan->corrs[corr];
  const Hankel& h = *plan->hankels[c.hankel_id];
  const int64_t E = (int64_t)c.evals.size();
  switch (stage) {
    case 0: return (int64_t)plan->nmu * plan->nk;
    case 1: return E * plan->n_ell * plan->nk_pad;
    case 2: return E * plan->n_ell * 2 * h.n_out_pad;
    case 3: return (int64_t)(2 + std::max(c.lay[LAY_MOM].n_metal, 1)) * c.ld_model;
    case 4: return c.ld_model;
    default: return -1;
  }
}

int vg_stage_batch(vg_plan* plan, int32_t corr, int32_t eval, int32_t stage,
                   const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev,
                   void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && out_dev && status_dev, "vg_stage_batch: bad arguments");
  VG_CORR(plan, corr);
  cudaStream_t st = (cudaStream_t)stream;
  VG_REQUIRE(B <= plan->chunk_max, "vg_stage_batch: B exceeds the chunk size");
  VG_TRY(ensure_workspace(plan, B));
  VG_TRY(expand_params(plan, theta_dev, B, status_dev, st));
  const int64_t per = vg_stage_size(plan, corr, stage);
  VG_REQUIRE(per > 0, "vg_stage_batch: bad stage");
  if (stage == 0) {
    VG_REQUIRE(eval >= 0 && eval < (int)c.evals.size(), "bad evaluation index");
    dim3 grid(vg_cdiv(plan->nk, 128), plan->nmu, (unsigned)B);
    vg::pk_grid_kernel<<<grid, 128, 0, st>>>(plan->grid_dev(), c.d_evals.as<vg_eval_desc>(), eval,
                                           plan->w_pfull.as<double>(), plan->npf, out_dev, status_dev);
    VG_LAUNCHED();
    return VG_OK;
  }
  plan->vt_now = false;      // the taps expose the per-point pass
  VG_TRY(run_corr_front(plan, c, B, status_dev, st, stage == 4 ? 99 : stage));
  const int nm = std::max(c.lay[LAY_MOM].n_metal, 1);
  if (stage == 1) {
    VG_CUDA_CHECK(cudaMemcpyAsync(out_dev, plan->w_pkell.p, (size_t)B * per * 8, cudaMemcpyDeviceToDevice, st));
  } else if (stage == 2) {
    VG_CUDA_CHECK(cudaMemcpyAsync(out_dev, plan->w_tables.p, (size_t)B * per * 8, cudaMemcpyDeviceToDevice, st));
  } else if (stage == 3) {
    // [B][2 + n_metal][ld_model]: peak core, smooth core, weighted metal xi's
    for (int s = 0; s < 2; ++s)
      VG_CUDA_CHECK(cudaMemcpy2DAsync(out_dev + (size_t)s * c.ld_model, (size_t)per * 8,
                                      plan->w_core.as<double>() + (size_t)s * B * c.ld_model,
                                      (size_t)c.ld_model * 8, (size_t)c.ld_model * 8, B,
                                      cudaMemcpyDeviceToDevice, st));
    VG_CUDA_CHECK(cudaMemcpy2DAsync(out_dev + (size_t)2 * c.ld_model, (size_t)per * 8, plan->w_xmet.p,
                                    (size_t)nm * c.ld_model * 8, (size_t)nm * c.ld_model * 8, B,
                                    cudaMemcpyDeviceToDevice, st));
  } else {
    VG_CUDA_CHECK(cudaMemcpyAsync(out_dev, plan->w_x.p, (size_t)B * per * 8, cudaMemcpyDeviceToDevice, st));
  }
  return VG_OK;
}

int vg_ozaki_gemm(const double* A, const double* B, double* C, int M, int N, int K, long long lda,
                  long long ldb, long long ldc, int slices, int32_t* error_out, void* stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0 || slices < 3 || slices > vg::OZ_MAX_S) {
    vg_set_error("vg_ozaki_gemm: bad arguments");
    return VG_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  OzakiOperand oa, ob;
  DevBuf ws, err;
  VG_TRY(ozaki_slice(oa, A, lda, M, M, K, slices, st));
  VG_TRY(ozaki_slice(ob, B, ldb, N, N, K, slices, st));
  VG_TRY(ozaki_gemm(oa, M, ob, N, C, ldc, ws, err, st));
  int flag = 0;
  VG_CUDA_CHECK(cudaMemcpyAsync(&flag, err.p, sizeof(int), cudaMemcpyDeviceToHost, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));   // the operands above are released on return
  if (error_out) *error_out = flag;
  return VG_OK;
}

// ---- per-stage entry points on caller-supplied arrays ------------------------------------------
int vg_plan_set_templates(vg_plan* plan, const double* pk_full, const double* pk_smooth, void* stream) {
  VG_REQUIRE(plan && pk_full && pk_smooth, "vg_plan_set_templates: bad arguments");
  VG_CUDA_CHECK(cudaDeviceSynchronize());
  std::vector<double> peak(plan->nk);
  for (int i = 0; i < plan->nk; ++i) peak[i] = pk_full[i] - pk_smooth[i];
  VG_CUDA_CHECK(cudaMemcpy(plan->d_pk[0].p, peak.data(), (size_t)plan->nk * 8, cudaMemcpyHostToDevice));
  VG_CUDA_CHECK(cudaMemcpy(plan->d_pk[1].p, pk_smooth, (size_t)plan->nk * 8, cudaMemcpyHostToDevice));
  VG_CUDA_CHECK(cudaMemcpy(plan->d_pk[2].p, pk_full, (size_t)plan->nk * 8, cudaMemcpyHostToDevice));
  // everything derived from the linear P(k): static moment tables, their xi stage, folded and
  // frozen metal terms -- the same pass vg_plan_finalize runs
  if (plan->finalized) return vg_plan_finalize(plan, stream);
  return VG_OK;
}

int vg_pk_batch(vg_plan* plan, int32_t corr, int32_t eval, const double* pk_lin_host, int32_t fast_metals,
                const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && plan->finalized && B > 0 && out_dev && status_dev, "vg_pk_batch: bad arguments");
  VG_CORR(plan, corr);
  VG_REQUIRE(eval >= 0 && eval < (int)c.evals.size(), "vg_pk_batch: bad evaluation index");
  VG_REQUIRE(B <= plan->chunk_max, "vg_pk_batch: B exceeds the chunk size");
  cudaStream_t st = (cudaStream_t)stream;
  vg_eval_desc d = c.evals[eval];
  if (pk_lin_host) {
    VG_TRY(vg_plan_set_direct_pk(plan, pk_lin_host));
    d.pk_lin_sel = 3;
  }
  if (fast_metals >= 0) d.fast_metals = fast_metals;
  DevBuf tmp;
  VG_TRY(upload(tmp, &d, 1));
  VG_TRY(ensure_workspace(plan, B));
  VG_TRY(expand_params(plan, theta_dev, B, status_dev, st));
  dim3 grid(vg_cdiv(plan->nk, 128), plan->nmu, (unsigned)B);
  vg::pk_grid_kernel<<<grid, 128, 0, st>>>(plan->grid_dev(), tmp.as<vg_eval_desc>(), 0, plan->w_pfull.as<double>(),
                                         plan->npf, out_dev, status_dev);
  VG_LAUNCHED();
  VG_CUDA_CHECK(cudaStreamSynchronize(st));     // `tmp` is released on return
  return VG_OK;
}

namespace {
vg::HankelDev hankel_dev(const Hankel& h) {
  vg::HankelDev d{};
  d.n_ell = h.n_ell; d.n_out = h.n_out; d.n_out_pad = h.n_out_pad; d.bounds_error = h.bounds_error;
  d.uniform = h.uniform;
  for (int l = 0; l < vg::PK_MAX_ELL; ++l) {
    d.ells[l] = h.ells[l]; d.x0[l] = h.x0[l]; d.x1[l] = h.x1[l];
    d.inv_dx[l] = l < h.n_ell ? (double)(h.n_out - 1) / (h.x1[l] - h.x0[l]) : 0.0;
  }
  d.xk = h.xk.as<double>(); d.inv_h = h.inv_h.as<double>();
  return d;
}

// multipoles + knot tables of B caller-supplied P(k,mu) grids: ells [B][n_ell][nk_pad], tables [B][n_ell][2 n_out_pad]
int grid_to_tables(vg_plan* plan, const Hankel& h, const double* pk_grid_dev, long long B, DevBuf& ells, DevBuf& tables,
                   cudaStream_t st) {
  const int rowT = 2 * h.n_out_pad;
  VG_TRY(ells.alloc((size_t)B * plan->n_ell * plan->nk_pad * 8));
  VG_TRY(tables.alloc((size_t)B * plan->n_ell * rowT * 8));
  dim3 grid(vg_cdiv(plan->nk_pad, 128), plan->n_ell, (unsigned)B);
  vg::project_grid_kernel<<<grid, 128, 0, st>>>(pk_grid_dev, plan->nmu, plan->nk, plan->nk_pad, plan->n_ell,
                                              plan->d_wleg.as<double>(), ells.as<double>());
  VG_LAUNCHED();
  VG_TRY(gemm_auto(plan, ells.as<double>(), (long long)plan->n_ell * plan->nk_pad, h.op.as<double>(), plan->nk_pad,
                   tables.as<double>(), (long long)plan->n_ell * rowT, (int)B, rowT, plan->nk_pad, plan->n_ell,
                   plan->nk_pad, (long long)rowT * plan->nk_pad, rowT, st));
  return VG_OK;
}
}  // namespace

int vg_pktoxi_batch(vg_plan* plan, int32_t hankel_id, const double* pk_grid_dev, int64_t B, const double* r_host,
                    const double* mu_host, int32_t n, int32_t single_ell, double* xi_out_dev, double* ells_out_dev,
                    int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && pk_grid_dev && B > 0 && status_dev, "vg_pktoxi_batch: bad arguments");
  VG_REQUIRE(hankel_id >= 0 && hankel_id < (int)plan->hankels.size(), "vg_pktoxi_batch: bad transform id");
  VG_REQUIRE(n == 0 || (r_host && mu_host && xi_out_dev), "vg_pktoxi_batch: coordinates without an output");
  cudaStream_t st = (cudaStream_t)stream;
  const Hankel& h = *plan->hankels[hankel_id];
  DevBuf ells, tables, dr, dm;
  VG_CUDA_CHECK(cudaMemsetAsync(status_dev, 0, (size_t)B * 4, st));
  VG_TRY(grid_to_tables(plan, h, pk_grid_dev, B, ells, tables, st));
  if (ells_out_dev)
    VG_CUDA_CHECK(cudaMemcpyAsync(ells_out_dev, ells.p, (size_t)B * plan->n_ell * plan->nk_pad * 8,
                                  cudaMemcpyDeviceToDevice, st));
  if (n > 0) {
    VG_TRY(upload(dr, r_host, n));
    VG_TRY(upload(dm, mu_host, n));
    dim3 grid(vg_cdiv(n, 256), (unsigned)B);
    vg::xi_plain_kernel<<<grid, 256, 0, st>>>(hankel_dev(h), tables.as<double>(), dr.as<double>(), dm.as<double>(), n,
                                            single_ell, xi_out_dev, status_dev);
    VG_LAUNCHED();
  }
  VG_CUDA_CHECK(cudaStreamSynchronize(st));     // the temporaries are released on return
  return VG_OK;
}

int vg_xi_from_grid_batch(vg_plan* plan, int32_t corr, int32_t eval, const double* pk_grid_dev, int32_t pk_lin_sel,
                          const double* theta_dev, int64_t B, double* out_dev, int32_t* status_dev, void* stream) {
  VG_REQUIRE(plan && plan->finalized && pk_grid_dev && B > 0 && out_dev && status_dev,
             "vg_xi_from_grid_batch: bad arguments");
  VG_CORR(plan, corr);
  VG_REQUIRE(eval >= 0 && eval < (int)c.evals.size(), "vg_xi_from_grid_batch: bad evaluation index");
  VG_REQUIRE(B <= plan->chunk_max, "vg_xi_from_grid_batch: B exceeds the chunk size");
  cudaStream_t st = (cudaStream_t)stream;
  const Hankel& h = *plan->hankels[c.hankel_id];
  // a one-evaluation view of the correlation: the caller's tables in slot 0, weight 1
  std::vector<vg_eval_desc> evs(c.evals);
  if (pk_lin_sel >= 0 && pk_lin_sel < 3) evs[eval].pk_lin_sel = pk_lin_sel;   // rel / asy tables of that linear P(k)
  DevBuf d_evs, d_id, d_slot, d_map, ells, tables, flat;
  VG_TRY(upload(d_evs, evs.data(), evs.size()));
  VG_TRY(upload(d_id, &eval, 1));
  std::vector<int> zeros(std::max<size_t>(evs.size(), (size_t)std::max(c.n_metal, 1)), 0);
  VG_TRY(upload(d_slot, zeros.data(), zeros.size()));
  VG_TRY(upload(d_map, zeros.data(), zeros.size()));
  VG_TRY(flat.alloc((size_t)B * c.ld_model * 8));
  VG_TRY(ensure_workspace(plan, B));
  VG_TRY(expand_params(plan, theta_dev, B, status_dev, st));
  VG_TRY(grid_to_tables(plan, h, pk_grid_dev, B, ells, tables, st));
  GatherJob job{};
  job.ids = d_id.as<int>(); job.n_ids = 1;
  job.tab_slot = d_slot.as<int>(); job.metal_map = d_map.as<int>(); job.n_metal = 1;
  job.pfull = plan->w_pfull.as<double>(); job.B = B;
  job.tables = tables.as<double>(); job.n_full = 1;
  job.unit_bias = 2;
  job.flat_out = flat.as<double>();
  job.status = status_dev;
  c.d_evals.swap(d_evs);                        // launch_xi_gather reads the descriptors of the correlation
  int rc = launch_xi_gather(plan, c, job, st);
  c.d_evals.swap(d_evs);
  VG_TRY(rc);
  VG_CUDA_CHECK(cudaMemcpy2DAsync(out_dev, (size_t)c.n_model * 8, flat.p, (size_t)c.ld_model * 8, (size_t)c.n_model * 8, B,
                                  cudaMemcpyDeviceToDevice, st));
  VG_CUDA_CHECK(cudaStreamSynchronize(st));
  return VG_OK;
}

int vg_dgemm_tn(const double* A, const double* B, double* C, int M, int N, int K,
                long long lda, long long ldb, long long ldc, double alpha, double beta,
                int splits, double* workspace, void* stream) {
  vg::GemmArgs g{};
  g.A = A; g.B = B; g.M = M; g.N = N; g.K = K;
  g.lda = lda; g.ldb = ldb; g.ldc = ldc;
  g.batch = 1; g.splits = splits < 1 ? 1 : splits;
  g.alpha = alpha; g.beta = beta;
  cudaStream_t st = (cudaStream_t)stream;
  if (g.splits == 1) {
    g.C = C;
    return vg::launch_dgemm(g, vg::GEMM_STORE, st);
  }
  if (!workspace || beta != 0.0) { vg_set_error("split-K needs a workspace and beta == 0"); return VG_ERR_ARG; }
  g.C = workspace; g.split_stride = (long long)M * ldc;
  int rc = vg::launch_dgemm(g, vg::GEMM_STORE, st);
  if (rc) return rc;
  long long n = (long long)M * ldc;
  vg::reduce_splits_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(workspace, C, n, g.splits, g.split_stride);
  VG_CUDA_CHECK(cudaGetLastError());
  return VG_OK;
}

}  // extern "C"

// ---- split-K DGEMM without any allocation (no counterpart in the reference; see the header) ----
// The split-K partials live in this file-scope device array: every GPU that loads the library holds
// VG_STATIC_SCRATCH_DOUBLES * 8 bytes (16 MiB) until exit.  All calls share it and it keeps the last
// call's contents; the GEMM of a call writes every partial its reduction reads (empty K ranges store
// zeros), so nothing needs clearing.  In host code the NAME below is a host-side stand-in, not a
// device address: the device address comes from cudaGetSymbolAddress, once per device.
static __device__ __align__(256) double vg_static_scratch[VG_STATIC_SCRATCH_DOUBLES];

namespace {
// C[m, n] = sum_s W[s * stride + m * ldw + n] for n < N only (fixed order => deterministic); the pad
// columns of C and of the partials are neither read nor written
__global__ void reduce_splits_cols_kernel(const double* __restrict__ W, double* __restrict__ C, int M, int N,
                                          long long ldw, long long ldc, int splits, long long stride) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)M * N) return;
  const long long m = i / N, n = i - m * N;
  const double* w = W + m * ldw + n;
  double s = w[0];
  for (int k = 1; k < splits; ++k) s += w[(long long)k * stride];
  C[m * ldc + n] = s;
}

constexpr int STATIC_SCRATCH_MAX_DEVICES = 64;
std::mutex g_static_scratch_mutex;
double* g_static_scratch_ptr[STATIC_SCRATCH_MAX_DEVICES] = {};

// device address of vg_static_scratch on the CURRENT device, looked up once per device
int static_scratch_address(double** out) {
  int dev = -1;
  VG_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= STATIC_SCRATCH_MAX_DEVICES) {
    vg_set_error("vg_dgemm_tn_static: device ordinal %d outside [0, %d)", dev, STATIC_SCRATCH_MAX_DEVICES);
    return VG_ERR_ARG;
  }
  std::lock_guard<std::mutex> lock(g_static_scratch_mutex);
  if (!g_static_scratch_ptr[dev]) {
    double* ptr = nullptr;
    VG_CUDA_CHECK(cudaGetSymbolAddress((void**)&ptr, vg_static_scratch));
    g_static_scratch_ptr[dev] = ptr;
  }
  *out = g_static_scratch_ptr[dev];
  return VG_OK;
}
}  // namespace

extern "C" int vg_dgemm_tn_static(const double* A, const double* B, double* C, int M, int N, int K, long long lda,
                                  long long ldb, long long ldc, double alpha, int splits, void* stream) {
  if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0 || splits < 2 || ldc < N) {
    vg_set_error("vg_dgemm_tn_static: bad arguments (splits >= 2, ldc >= N, positive sizes, non-null pointers)");
    return VG_ERR_ARG;
  }
  const long long ldw = N + (N & 1);                    // even row stride: the kernel stores pairs
  const long long need = (long long)splits * M * ldw;
  if (need > (long long)VG_STATIC_SCRATCH_DOUBLES) {
    vg_set_error("vg_dgemm_tn_static: splits * M * (N rounded up to even) = %lld doubles exceeds the static "
                 "scratch of %lld doubles (VG_STATIC_SCRATCH_DOUBLES); use vg_dgemm_tn with a caller workspace",
                 need, (long long)VG_STATIC_SCRATCH_DOUBLES);
    return VG_ERR_ARG;
  }
  double* scratch = nullptr;
  VG_TRY(static_scratch_address(&scratch));
  cudaStream_t st = (cudaStream_t)stream;
  vg::GemmArgs g{};
  g.A = A; g.B = B; g.M = M; g.N = N; g.K = K;
  g.lda = lda; g.ldb = ldb; g.ldc = ldw;
  g.batch = 1; g.splits = splits;
  g.alpha = alpha; g.beta = 0.0;
  g.C = scratch; g.split_stride = (long long)M * ldw;
  VG_TRY(vg::launch_dgemm(g, vg::GEMM_STORE, st));
  const long long n = (long long)M * N;
  reduce_splits_cols_kernel<<<vg_cdiv(n, 256), 256, 0, st>>>(scratch, C, M, N, ldw, ldc, splits, g.split_stride);
  VG_CUDA_CHECK(cudaGetLastError());
  return VG_OK;
}

// ---- FP64 FMA peak probe (roofline denominator of the ALU-bound fused P(k,mu) kernel) ----
namespace {
__global__ void __launch_bounds__(256) dfma_probe_kernel(double* out, int iters, double seed) {
  double a[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) a[j] = seed + j + threadIdx.x * 1e-3;
  const double x = 1.0000001, y = 1e-9;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = fma(a[j], x, y);
  }
  double s = 0.0;
#pragma unroll
  for (int j = 0; j < 8; ++j) s += a[j];
  if (s == 12345.678) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
}  // namespace

extern "C" int vg_probe_dfma(int blocks, int iters, double* out_dev, double* tflops, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaEvent_t e0, e1;
  VG_CUDA_CHECK(cudaEventCreate(&e0));
  VG_CUDA_CHECK(cudaEventCreate(&e1));
  dfma_probe_kernel<<<blocks, 256, 0, st>>>(out_dev, iters, 1.0);   // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    VG_CUDA_CHECK(cudaEventRecord(e0, st));
    dfma_probe_kernel<<<blocks, 256, 0, st>>>(out_dev, iters, 1.0 + rep);
    VG_CUDA_CHECK(cudaEventRecord(e1, st));
    VG_CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0.f;
    VG_CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    best = std::min(best, ms);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *tflops = 2.0 * 8.0 * (double)iters * 256.0 * blocks / (best * 1e-3) / 1e12;
  return VG_OK;
}
