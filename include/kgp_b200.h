/* kgp_b200.h -- C ABI of the B200-native GP inference engine.
 *
 * This is the drop-in boundary for the ONE hot path of alshedivat/keras-gp: the GP
 * engine behind kgp.backend (nlZ, dnlZ, dnlZ/dX, predict for infExact / infGrid).
 * In the reference that path is a chain  Python -> oct2py -> Octave -> GPML .m code
 * (kgp/backend/gpml.py:13-31 command strings, kgp/backend/engines.py:72-99).  There
 * is no FFI in the reference to mirror, so each entry point cites the command
 * string / .m function it replaces.  Plain pointers and sizes only; no torch types.
 *
 * Conventions: all reals are FP64 (the reference up-casts everything,
 * kgp/backend/engines.py:84-88).  Matrices are row-major (NumPy C order).  Pointers
 * are HOST pointers when the function name ends in _host, DEVICE pointers when it
 * ends in _dev.  Every call returns a kgp_status; kgp_last_error() gives text.
 * One context = one CUDA device + one stream + its workspaces; calls on a context
 * are blocking at return and must not be issued concurrently.
 */
#ifndef KGP_B200_H
#define KGP_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct kgp_ctx kgp_ctx;

typedef enum {
  KGP_OK = 0,
  KGP_BAD_ARG = 1,
  KGP_CUDA_ERR = 2,
  KGP_NOT_PD = 3,            /* Octave: "chol: input matrix must be positive definite" */
  KGP_CG_NOT_CONVERGED = 4,  /* GPML/cov/apx.m:277 error('Not converged ...') */
  KGP_NO_GRID = 5,           /* infGrid called before update_grid / xg */
  KGP_NO_DATA = 6
} kgp_status;

/* data slots = the Octave workspace variables X_tr/y_tr, X_tst, X_val, X_tmp
 * (kgp/backend/gpml.py:152-158) */
enum { KGP_SLOT_TR = 0, KGP_SLOT_TST = 1, KGP_SLOT_VAL = 2, KGP_SLOT_TMP = 3, KGP_NUM_SLOTS = 4 };

enum { KGP_INF_EXACT = 0, KGP_INF_GRID = 1 };                 /* infExact | infGrid */
enum { KGP_COV_SEISO = 0, KGP_COV_SEARD = 1, KGP_COV_SM = 2 }; /* covSEiso | covSEard | covSM */
/* mean function: bit mask, summed in this order (meanSum of the set bits) */
enum { KGP_MEAN_ZERO = 0, KGP_MEAN_LINEAR = 1, KGP_MEAN_CONST = 2 };

/* GPML `opt` struct fields that reach this path (GPML/cov/apx.m:118-130) */
typedef struct {
  double cg_tol;          /* default 1e-6 */
  int    cg_maxit;        /* default min(n,20); <=0 selects the default */
  int    ldB2_cheby;      /* 0: Fiedler bound (default), 1: Chebyshev-Hutchinson */
  int    ldB2_cheby_hutch;   /* probes m, default 10 */
  int    ldB2_cheby_degree;  /* degree d, default 15 */
  int    ldB2_cheby_maxit;   /* power/Lanczos iterations for emax, default 50 */
  uint64_t ldB2_cheby_seed;  /* probe seed (intent of apx.m:130) */
  int    ard_reference_form; /* 1: dcovSEard.m:7 quirk (divide every dim by ell_1^2) */
  int    cg_precond;         /* 1: grid-space Woodbury preconditioner for the CG (extension); 0 = reference CG */
  int    pred_var;           /* opt.pred_var (GPML/inf/infGrid.m:94): ns = max(ceil(|pred_var|),20) perturb-and-MAP samples
                              * for the explained variance that kgp_predict_host subtracts; 0 = prior variance only */
  int    ldB2_slq;           /* 1: stochastic Lanczos quadrature log-det + Hutchinson derivatives (extension);
                              * probes as for ldB2_cheby (ldB2_cheby_hutch, ldB2_cheby_seed, kgp_set_probes_host) */
  int    ldB2_slq_steps;     /* Lanczos steps per probe, default 30 */
  int    cg_deterministic;   /* 1 (default): the scatter Mx'(sW d) of the persistent CG accumulates in order-independent fixed point
                              * (integer atomics) -> two runs of a solve are bit-identical; 0: FP64 atomics, ~25 % faster at n = 1M,
                              * last-bit differences from run to run (iteration counts may differ by a few per cent) */
  int    deg;                /* opt.deg (GPML/cov/apx.m:27-32,122; apxGrid.m:495-504): interpolation degree on equispaced axes.
                              * 3 (default) Keys cubic, 1 linear, 0 nearest node; other values (GP interpolation) are refused.
                              * Non-equispaced axes ignore it.  deg != 3 runs the general CG path (no cg_precond / pred_var /
                              * SLQ / Toeplitz factors, like non-equispaced axes) */
  int    deg_axis[4];        /* one degree per grid dimension (GPML/cov/apxGrid.m:448 "deg = deg(:).*ones(p,1)"); an entry < 0 (the
                              * default) means "use deg" */
} kgp_opt;

/* per-call statistics of the last kgp_eval / kgp_dlik_dx / kgp_predict */
typedef struct {
  double ms_total;        /* device time of the last call (CUDA events) */
  double ms_kbuild, ms_potrf, ms_inverse, ms_reduce, ms_solve;   /* exact path phases */
  double ms_gemm;         /* time inside dgemm_nt_kernel launches only (events around them; 0 unless profiling on) */
  double gemm_flops;      /* algorithmic flops issued to dgemm_nt_kernel in the last call */
  long   gemm_launches, kernel_launches;
  int    cg_iters; double cg_relres;   /* grid path */
  int    potrf_info;      /* 0 or 1-based index of the failing pivot */
  long   n, np;           /* problem size and padded size */
  /* dgemm_nt launches by call site (profiling on): 0 potrf panel update, 1 potrf trailing SYRK,
   * 2 Z-sweep GEMM, 3 trsm block update, 4 Z Z^T, 5 predict */
  double gemm_class_ms[8], gemm_class_flops[8];
} kgp_stats;

/* lifetime ------------------------------------------------------------------- */
int  kgp_ctx_create(int device, kgp_ctx** out);
int  kgp_ctx_destroy(kgp_ctx* ctx);
const char* kgp_last_error(kgp_ctx* ctx);
const char* kgp_version(void);
void kgp_opt_default(kgp_opt* opt);
/* sizeof(kgp_opt) (which = 0) / sizeof(kgp_stats) (which = 1) as the library was compiled: a binding (the ctypes stub of
 * INTEGRATION.md, a cgo / JNI struct) checks its own layout against it before the first call.  No counterpart in the reference
 * (oct2py marshals dictionaries); -1 for an unknown `which`. */
long kgp_abi_sizeof(int which);
int  kgp_set_profiling(kgp_ctx* ctx, int on);   /* on: time GEMM launches separately (adds syncs) */

/* multi-GPU: one context per rank / GPU, NCCL over NVLink.  No reference counterpart (the
 * reference runs one Octave process per GP layer, kgp/layers.py:63).  Rank 0 obtains the 128-byte NCCL id, the
 * host layer ships it to the other ranks (torch.distributed in this repo), every rank then calls kgp_comm_init.
 * After that kgp_eval on the exact path shards ONE n x n problem over the ranks (block columns of the Cholesky,
 * rows of the triangular inverse, tiles of the inverse); every rank must issue the same calls with the same
 * data and receives the same results.  nccl_lib_path may be NULL (libnccl.so.2 on the loader path). */
int kgp_nccl_unique_id(char* out128, const char* nccl_lib_path);
const char* kgp_dist_last_error(void);
int kgp_comm_init(kgp_ctx* ctx, int nranks, int rank, const char* id128, const char* nccl_lib_path);
/* ownership maps of the sharded exact path (host logic, callable without a GPU) */
int  kgp_dist_col_owner(int block, int nranks);              /* rank owning block column `block` (NB = 512) */
long kgp_dist_zrow_begin(int rank, int nranks, long n);     /* first row of Z = L^-T owned by `rank`; rank = nranks gives the padded order */
/* KISS-GP row-sharded CG: first data row owned by `rank` (contiguous shares; rank = nranks gives n); and whether a
 * problem of n points on N grid nodes is row-sharded at all (needs n >= 4096 per rank and N within the peer window) */
long kgp_dist_row_begin(int rank, int nranks, long n);
int  kgp_dist_rows_sharded(int nranks, long n, long N);

/* data: replaces eng.push('X_<set>',X) / push('y_<set>',y)  (kgp/backend/gpml.py:152-158) */
int kgp_set_data_host(kgp_ctx* ctx, int slot, const double* X, long n, int D, const double* y /* may be NULL */);
int kgp_set_data_dev (kgp_ctx* ctx, int slot, const double* X, long n, int D, const double* y /* may be NULL */);

/* model: replaces configure()'s push('hyp',hyp) / push('opt',opt) and the cov/mean/inf
 * strings (kgp/backend/gpml.py:76-150).  hyp_cov in GPML order ([log ell..., log sf]);
 * for infGrid the per-dimension blocks are concatenated (column-major read of the
 * tiled array, GPML/cov/apxGrid.m:230-231). */
int kgp_set_model(kgp_ctx* ctx, int inf_kind, int cov_kind, int sm_Q, int mean_mask,
                  const double* hyp_cov, int n_cov, double hyp_lik,
                  const double* hyp_mean, int n_mean, const kgp_opt* opt /* may be NULL */);

/* evaluate: replaces "[nlZ dnlZ post] = gp(hyp, inf, mean, cov, lik, X, y)"
 * (kgp/backend/gpml.py:16-18; GPML/gp.m:111-126 -> infGaussLik.m:19-33).
 * want: 0 = nlZ only, 1 = + dnlZ, 2 = + dnlZ/dX kept on the device for kgp_dlik_dx.
 * dnlZ_host receives [cov..., lik, mean...] (minimize.m:173-186 unwrap order). */
int kgp_eval(kgp_ctx* ctx, int slot, int want, double* nlZ_host, double* dnlZ_host);

/* input gradient: replaces "dlik_dx = dlik(hyp, mean, cov, lik, dcov, X, y)"
 * (kgp/backend/gpml.py:26-28; kgp/backend/utils/dlikExact.m:9-25, dlikGrid.m:8-40).
 * out is n x D row-major.  Re-uses the factorisation of a preceding kgp_eval(...,2)
 * on the same slot/model, otherwise runs the inference itself like the reference. */
int kgp_dlik_dx_host(kgp_ctx* ctx, int slot, double* out);
int kgp_dlik_dx_dev (kgp_ctx* ctx, int slot, double* out);

/* predict: replaces "[ymu ys2 fmu fs2] = gp(hyp, inf, mean, cov, lik, X_tr, y_tr, X_tst)"
 * (kgp/backend/gpml.py:19-25; GPML/gp.m:128-181; GPML/inf/infGrid.m:131-145).
 * Uses slot TR as training data; any output pointer may be NULL. */
int kgp_predict_host(kgp_ctx* ctx, const double* Xs, long ns,
                     double* ymu, double* ys2, double* fmu, double* fs2);

/* Allocation-free predictive mean.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * ymu_dev[i] = m(x*_i) + k(x*_i, X) alpha, the ymu of kgp_predict_host, for infExact with covSEiso / covSEard.
 * Xs_dev (ns x D row-major) and ymu_dev (ns) are DEVICE pointers supplied by the caller; the posterior must be current
 * (kgp_eval on slot TR since the last data / hyper-parameter change), otherwise KGP_NO_DATA -- it never factorises.  Calls no
 * cudaMalloc / cudaFree or allocating library routine: its scratch (ns x ceil(n/512) partial sums) is a static 2 MiB device
 * array inside the library, held by every GPU that loads it until exit; ns * ceil(n/512) > 262144 is refused with KGP_BAD_ARG
 * (n = 40 000: at most 3318 test points per call).  All calls share the array and it keeps the last call's contents (every
 * element read is written first, nothing needs clearing); calls that overlap on one GPU race on it.  Synchronises
 * the context's stream before returning. */
int kgp_predict_mean_noalloc_dev(kgp_ctx* ctx, const double* Xs_dev, long ns, double* ymu_dev);

/* grid: replaces "xg = covGrid('create', X, eq, k)" (kgp/backend/gpml.py:29-31;
 * GPML/cov/apxGrid.m:389-416) and push('xg', ...) (kgp/backend/gpml.py:125-126). */
int kgp_grid_create(kgp_ctx* ctx, int slot, int eq, double k);
int kgp_grid_set_host(kgp_ctx* ctx, int n_axes, const int* axis_len, const double* axes_concat);
/* Mark grid dimensions as Toeplitz factors -- what a nested cell xg{i} = {axis} selects in GPML (GPML/cov/apxGrid.m:258-281):
 * the MVM is the same Toeplitz product, but the Fiedler log-det takes the factor's eigenvalues from a Whittle circulant
 * embedding (apxGrid.m:576-580,617-627) and differentiates it over the embedding lags (GPML/cov/apx.m:256-266), so nlZ and
 * dnlZ differ from dense-factor mode on the same axes.  flags: n_axes ints (NULL clears); call after kgp_grid_set_host. */
int kgp_grid_set_toeplitz(kgp_ctx* ctx, int n_axes, const int* flags);
/* GPML's equi() test for one axis (GPML/cov/apxGrid.m:684-695): 1 when max|diff(a) - (a[1]-a[0])| < 1e-9 and the axis
 * increases.  Axes that fail it but increase take the linear neqinterp branch (apxGrid.m:541-557).  Host logic, no GPU. */
int kgp_grid_axis_equispaced(const double* axis, int len);
/* The clustering step of covGrid('create', x, 0, k) (GPML/cov/apxGrid.m:411 "[idx,xgj] = kmeans(u, min(numel(u),ngj-2))") for
 * one axis: u = the sorted unique values (ascending, nu of them), K clusters, centres written in increasing order.  Octave's
 * kmeans() starts from random centres and cannot be reproduced; this is Lloyd's algorithm from an equal-count partition,
 * deterministic.  kgp_grid_create(eq = 0) calls it per axis and adds the boundary nodes of apxGrid.m:412-415.  Host logic,
 * no GPU.  Returns 0, or -1 on bad arguments (K < 1 or K > nu). */
int kgp_kmeans1d_host(const double* u, long nu, long K, double* centres);
int kgp_grid_get_host(kgp_ctx* ctx, int* n_axes, int* axis_len /* >= 16 ints */, double* axes_concat /* may be NULL */);

/* Learned input projection hyp.proj (GPML/inf/infGrid.m:56-65,82-87,155-167; kgp/backend/utils/dlikGrid.m:13-22,
 * deriv_x.m:5-6,24).  P is the EFFECTIVE projector hP (d x D row-major; the caller has applied opt.proj = 'orth' / 'norm').
 * While set, kgp_set_data_* takes n x D inputs and the engine works on x hP' (n x d); kgp_predict_host projects its test
 * inputs; kgp_dlik_dx_* returns dxP hP (n x D); a new P re-projects the loaded slots.  P = NULL or d = 0 clears.
 * kgp_dnlZ_proj_host: dnlZ/dhP (d x D) = dxP' x -- deriv_proj's dP(i,j) = -alpha' (x(:,j) .* beta_i), with the mean's input
 * derivative analytic instead of the reference's forward difference.  infGrid only. */
int kgp_set_proj_host(kgp_ctx* ctx, const double* P, int d, int D);
int kgp_dnlZ_proj_host(kgp_ctx* ctx, int slot, double* out /* d x D */);

/* Chebyshev-Hutchinson log-det (opt.ldB2_cheby = 1; GPML/cov/apx.m:336-391): inject the probe vectors
 * (V is m x n, one probe per row, entries +-1) and the eigenvalue bound emax.  The reference draws them with
 * randn('seed',.) / eigs inside Octave, which cannot be reproduced, so a parity run passes the
 * same V and emax to the oracle and to the engine.  V = NULL returns to generated probes (splitmix64 of
 * opt.ldB2_cheby_seed) and power-iteration emax; emax <= 0 with V given estimates emax only. */
int kgp_set_probes_host(kgp_ctx* ctx, const double* V, long n, int m, double emax);

/* post.L of the KISS-GP posterior applied to m right-hand sides in ONE batched CG (GPML/cov/apx.m:219-222,
 * 281-302, per-column convergence masks): out = sW .* CG(I + sW Mx Kg Mx' sW, sW .* r).  R and out are m x n,
 * one right-hand side per row.  The building block of opt.pred_var (GPML/inf/infGrid.m:93-104). */
int kgp_grid_solve_host(kgp_ctx* ctx, int slot, const double* R, int m, double* out);

/* Explained variance on the grid by perturb-and-MAP (GPML/inf/infGrid.m:91-101, sample() :147-152):
 *   a ~ N(0,Ku) (eigen square root), A = Mx a + randn/sW, A <- post.L(A), z = Ku Mx' A, vg = mean(z.^2).
 * The reference's unseeded randn(N,ns) / randn(n,ns) are the explicit inputs Rg (ns x N) and Rn (ns x n), one draw per
 * row; both NULL -> generated on the device from `seed`.  The result is cached for kgp_predict_host (used when
 * opt.pred_var != 0) and copied to vg_out (N, first axis fastest) when that is not NULL. */
int kgp_grid_pred_var_host(kgp_ctx* ctx, int slot, const double* Rg, const double* Rn, int ns, uint64_t seed, double* vg_out);

/* posterior pieces for tests / callers that want them (host copies) */
int kgp_get_alpha_host(kgp_ctx* ctx, double* alpha /* n */);

int kgp_last_stats(kgp_ctx* ctx, kgp_stats* out);

/* measurement helpers (bench.py): FP64 DMMA issue-rate micro-benchmark and a timed
 * dgemm_nt launch on synthetic operands; both return TFLOP/s. */
int kgp_bench_dmma_peak(kgp_ctx* ctx, int iters, double* tflops);
int kgp_bench_dgemm(kgp_ctx* ctx, int n, int k, int reps, double* tflops, double* ms);   /* C = A B^T */
int kgp_bench_syrk (kgp_ctx* ctx, int n, int k, int reps, double* tflops, double* ms);   /* C -= A A^T, lower tiles */

#ifdef __cplusplus
}
#endif
#endif
