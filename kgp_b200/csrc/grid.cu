// Host orchestration of the KISS-GP / MSGP (infGrid) path on one B200.
//
// Reference flow (GPML/inf/infGrid.m:53-89 -> infGaussLik.m:19-33 -> apx.m:116-138,218-244):
// interpolation matrix Mx, Kronecker grid covariance Kg, alpha by un-preconditioned CG on
// B = I + sW Mx Kg Mx' sW, Fiedler log-det bound from the factor eigenvalues, hyper-derivatives
// by dirder, input gradient by deriv_x.  Here Mx is never formed (weights recomputed per pass),
// the whole CG iteration is device-resident (scalars, convergence flag and iteration count live
// in device memory; the host only polls the flag every few iterations).
#include "ctx.h"
#include "grid_precond.cuh"
#include "grid_toeplitz.cuh"
#include "exact_dist.cuh"
#include <math.h>
#include <string.h>
#include <algorithm>

namespace kgp {

struct GridState {
  DevBuf axes, Kf, Vf, Af, ee, eoff, de, zv, Qm, Pm, g0, g1, tgrid, beta;
  DevBuf x, r, d, z, ym, partA, partB, S, hyp, hypm, dX, dhyp, out, sortk, sorti, keep, dmean;
  // cell-sorted CG state (grid_cg.cuh)
  DevBuf U, perm, sx, sr, sd, sz, ctl, cnt, KfT, gm0, gm1;
  double* h_ctl = nullptr;                // pinned host copy of the control blocks
  long npad = 0; int sort_slot = -1; unsigned long sort_slot_ver = 0, sort_grid_ver = 0;
  cudaGraphExec_t cg_graph = nullptr; std::vector<char> cg_graph_key;
  DevBuf probes, chT, chU, chY, chZ, chZ2, chS, chC, chP, bva, bvb, bvc, bvs, lzQ, lzP, lzH, chacc;   // Chebyshev log-det (grid_cheby.cuh)
  long probes_n = 0; int probes_m = 0; double probes_emax = 0.0; bool probes_injected = false;
  DevBuf Gm, Ht1, Ht2, VT, se, ps, pw, psel;           // Woodbury preconditioner (grid_precond.cuh); psel: kept modes
  long prank = 0;
  int gram_slot = -1; unsigned long gram_slot_ver = 0, gram_grid_ver = 0; int pcg_iters = 0;
  DevBuf kp0, kp1, gpad, tpad, upad;       // large-factor Kronecker MVM through the DMMA GEMM: zero-padded factors and grid matrices
  DevBuf pdz, prr, bar, part4, g0b, wmaxd, nodew2; double c2max2 = 0.0; int num_sms = 0; size_t smem_optin = 0; int cg_level_used = -1;
  int wmax = 0; int wmax_ctas = 0; unsigned long wmax_slot_ver = 0, wmax_grid_ver = 0; int wmax_slot = -1;
  int p = 0; int ng[GRID_MAXP] = {0}; long N = 0; bool any_neq = false;
  long koff[GRID_MAXP] = {0};            // offsets of factor i inside Kf / Vf
  int eoff_h[GRID_MAXP] = {0};
  unsigned long grid_ver = 0;
  // cached posterior
  unsigned long st_slot_ver = 0, st_model_ver = 0, st_grid_ver = 0; int st_slot = -1, st_level = -1;
  double prior_var = 0.0;
  // row-sharded CG (several ranks): the sorted set of THIS rank's rows lives in alt while the full-data set is
  // active and vice versa (swap_sorted); sub_n < 0 = the active set covers all rows of the slot
  struct SortedSet { DevBuf U, perm; long npad = 0; int sort_slot = -1; unsigned long sort_slot_ver = 0, sort_grid_ver = 0;
                     int wmax = 0, wmax_ctas = 0, wmax_slot = -1; unsigned long wmax_slot_ver = 0, wmax_grid_ver = 0;
                     long sub_lo = 0, sub_n = -1; double c2max2 = 0.0; } alt;
  long sub_lo = 0, sub_n = -1;
  // warm start of the factor eigen-decompositions: eigenvectors of the previous evaluation on the same grid
  DevBuf eigT; unsigned long eig_grid_ver = 0; int eig_evals = 0; bool eig_have = false;
  // explained variance on the grid (opt.pred_var; infGrid.m:91-101)
  DevBuf vg, pvA, pvS, pvG; bool eig_ok = false, vg_nonzero = false;
  unsigned long vg_slot_ver = 0, vg_model_ver = 0, vg_grid_ver = 0; int vg_slot = -1;
};

static GridState* gs(kgp_ctx* c) {
  if (!c->gridstate) c->gridstate = new GridState();
  return (GridState*)c->gridstate;
}
void grid_free(kgp_ctx* c) {
  if (!c->gridstate) return;
  GridState* g = (GridState*)c->gridstate;
  DevBuf* bufs[] = {&g->axes, &g->Kf, &g->Vf, &g->Af, &g->ee, &g->eoff, &g->de, &g->zv, &g->Qm, &g->Pm, &g->g0, &g->g1,
                    &g->tgrid, &g->beta, &g->x, &g->r, &g->d, &g->z, &g->ym, &g->partA, &g->partB, &g->S, &g->hyp,
                    &g->hypm, &g->dX, &g->dhyp, &g->out, &g->sortk, &g->sorti, &g->keep, &g->dmean,
                    &g->U, &g->perm, &g->sx, &g->sr, &g->sd, &g->sz, &g->ctl, &g->cnt, &g->KfT, &g->gm0, &g->gm1, &g->pdz, &g->prr, &g->bar, &g->part4, &g->g0b, &g->wmaxd, &g->nodew2,
                    &g->Gm, &g->Ht1, &g->Ht2, &g->VT, &g->se, &g->ps, &g->pw, &g->psel, &g->probes, &g->chT, &g->chU, &g->chY, &g->chZ, &g->chZ2, &g->bvs, &g->chS, &g->chC, &g->chP, &g->bva, &g->bvb, &g->bvc, &g->lzQ, &g->lzP, &g->lzH, &g->chacc,
                    &g->vg, &g->pvA, &g->pvS, &g->pvG, &g->alt.U, &g->alt.perm, &g->eigT, &g->kp0, &g->kp1, &g->gpad, &g->tpad, &g->upad};
  for (DevBuf* b : bufs) b->release();
  if (g->h_ctl) cudaFreeHost(g->h_ctl);
  if (g->cg_graph) cudaGraphExecDestroy(g->cg_graph);
  delete g; c->gridstate = nullptr;
}

static void swap_sorted(GridState* g) {
  GridState::SortedSet& a = g->alt;
  std::swap(g->U, a.U); std::swap(g->perm, a.perm); std::swap(g->npad, a.npad);
  std::swap(g->sort_slot, a.sort_slot); std::swap(g->sort_slot_ver, a.sort_slot_ver); std::swap(g->sort_grid_ver, a.sort_grid_ver);
  std::swap(g->wmax, a.wmax); std::swap(g->wmax_ctas, a.wmax_ctas); std::swap(g->wmax_slot, a.wmax_slot);
  std::swap(g->wmax_slot_ver, a.wmax_slot_ver); std::swap(g->wmax_grid_ver, a.wmax_grid_ver);
  std::swap(g->sub_lo, a.sub_lo); std::swap(g->sub_n, a.sub_n); std::swap(g->c2max2, a.c2max2);
}
static int nblocks_for(long n) { long b = (n + 255) / 256; return (int)std::min<long>(std::max<long>(b, 1), 1184); }

// upload axes, fill the descriptor
static int grid_setup(kgp_ctx* c, GridState* g, GridDesc& G) {
  if (!c->grid.set) return c->fail(KGP_NO_GRID, "no grid: call update_grid or pass grid_kwargs['xg']");
  const int p = (int)c->grid.axes.size();
  if (p > GRID_MAXP) return c->fail(KGP_BAD_ARG, "at most %d grid dimensions", GRID_MAXP);
  long tot = 0; for (auto& a : c->grid.axes) tot += (long)a.size();
  if (g->grid_ver != c->grid.version || g->p != p) {
    KGP_CUDA(c, g->axes.ensure(tot * 8));
    long o = 0;
    for (int d = 0; d < p; ++d) {
      KGP_CUDA(c, cudaMemcpyAsync(g->axes.d() + o, c->grid.axes[d].data(), c->grid.axes[d].size() * 8,
                                  cudaMemcpyHostToDevice, c->stream));
      o += (long)c->grid.axes[d].size();
    }
    g->grid_ver = c->grid.version;
  }
  memset(&G, 0, sizeof G);
  g->any_neq = false;
  G.p = p; g->p = p; long o = 0, N = 1, ko = 0; int eo = 0;
  for (int d = 0; d < p; ++d) {
    const std::vector<double>& a = c->grid.axes[d];
    G.ng[d] = (int)a.size(); G.stride[d] = N; N *= G.ng[d];
    G.ax[d] = g->axes.d() + o; o += (long)a.size();
    const bool neq = d < (int)c->grid.neq.size() && c->grid.neq[d] != 0;
    G.deg[d] = (d < 4 && c->model.opt.deg_axis[d] >= 0) ? c->model.opt.deg_axis[d] : c->model.opt.deg;   // opt.deg (apx.m:122, apxGrid.m:448): equispaced axes only, neqinterp ignores it
    // the cell-sorted kernels (persistent CG, Gram accumulation, sorted Chebyshev recurrence) hard-wire the cubic stencil:
    // every other interpolation takes the general streaming kernels, like a non-equispaced axis
    g->any_neq |= neq || G.deg[d] != 3;
    G.s0[d] = a[0]; G.ds[d] = neq ? 0.0 : a[1] - a[0];           // ds = 0 marks a non-equispaced axis for the interpolation kernels
    G.range[d] = neq ? 1.0 : *std::max_element(a.begin(), a.end()) - *std::min_element(a.begin(), a.end());   // deriv_x.m:11-15: wi = 1
    g->ng[d] = G.ng[d]; g->koff[d] = ko; ko += (long)G.ng[d] * G.ng[d]; g->eoff_h[d] = eo; eo += G.ng[d];
  }
  G.N = N; g->N = N;
  return KGP_OK;
}

template <typename F> static void for_p(int p, F&& f) {
  switch (p) { case 1: f(std::integral_constant<int, 1>()); break; case 2: f(std::integral_constant<int, 2>()); break;
               case 3: f(std::integral_constant<int, 3>()); break; default: f(std::integral_constant<int, 4>()); }
}

// g_out = Mx' (scale .* v)
static void scatter(kgp_ctx* c, const GridDesc& G, const double* X, long n, int D, const double* v, double scale, double* gout) {
  cudaMemsetAsync(gout, 0, G.N * 8, c->stream);
  const int use_smem = (G.N * 8 <= 96 * 1024) ? 1 : 0;
  const int blocks = use_smem ? (int)std::min<long>((n + 255) / 256, 296) : nblocks_for(n);
  for_p(G.p, [&](auto P) {
    interp_scatter_kernel<decltype(P)::value><<<blocks, 256, use_smem ? G.N * 8 : 0, c->stream>>>(G, X, n, D, v, scale, gout, use_smem);
  });
  c->stats.kernel_launches += 2;
}
// out = (K_p kron ... kron K_1) in ; skip = mode to leave out (-1: none).  Returns the buffer holding the result.
static double* kron_apply(kgp_ctx* c, GridState* g, const GridDesc& G, double* in, double* tmp, int skip, const double* mats = nullptr) {
  double* src = in; double* dst = tmp;
  if (!mats) mats = g->Kf.d();
  for (int d = 0; d < G.p; ++d) {
    if (d == skip) continue;
    long inner = G.stride[d], outer = G.N / (inner * G.ng[d]);
    mode_product_kernel<<<(unsigned)((G.N + 255) / 256), 256, 0, c->stream>>>(mats + g->koff[d], G.ng[d], src, dst, inner, outer);
    c->stats.kernel_launches++;
    std::swap(src, dst);
  }
  return src;
}

static int factor_nhyp(const Model& m) { return m.cov_kind == KGP_COV_SM ? 3 * m.sm_Q : 2; }

static int ensure_grid_buffers(kgp_ctx* c, GridState* g, const GridDesc& G, long n, int D, int nhyp) {
  long ksz = 0; int esz = 0, ngmax = 0;
  for (int d = 0; d < G.p; ++d) { ksz += (long)G.ng[d] * G.ng[d]; esz += G.ng[d]; ngmax = std::max(ngmax, G.ng[d]); }
  KGP_CUDA(c, g->Kf.ensure(ksz * 8)); KGP_CUDA(c, g->Vf.ensure(ksz * 8)); KGP_CUDA(c, g->Af.ensure(ksz * 8));
  KGP_CUDA(c, g->KfT.ensure((long)G.ng[0] * G.ng[0] * 8));
  KGP_CUDA(c, g->ee.ensure(esz * 8)); KGP_CUDA(c, g->eoff.ensure(GRID_MAXP * 4));
  KGP_CUDA(c, g->de.ensure(G.N * 8)); KGP_CUDA(c, g->zv.ensure(ngmax * 8));
  KGP_CUDA(c, g->Qm.ensure((long)ngmax * ngmax * 8)); KGP_CUDA(c, g->Pm.ensure((long)ngmax * ngmax * 8));
  KGP_CUDA(c, g->g0.ensure(G.N * 8)); KGP_CUDA(c, g->g1.ensure(G.N * 8));
  KGP_CUDA(c, g->tgrid.ensure(G.N * 8)); KGP_CUDA(c, g->beta.ensure(G.N * 8));
  DevBuf* vecs[] = {&g->x, &g->r, &g->d, &g->z, &g->ym};
  for (DevBuf* b : vecs) KGP_CUDA(c, b->ensure(n * 8));
  KGP_CUDA(c, g->partA.ensure(2048 * 8)); KGP_CUDA(c, g->partB.ensure(2048 * 8));
  KGP_CUDA(c, g->S.ensure(64 * 8)); KGP_CUDA(c, g->hyp.ensure(std::max(nhyp, 1) * 8));
  KGP_CUDA(c, g->hypm.ensure(128 * 8)); KGP_CUDA(c, g->dmean.ensure(64 * 8));
  KGP_CUDA(c, g->dX.ensure(n * D * 8)); KGP_CUDA(c, g->dhyp.ensure((std::max(nhyp, 1) + 128) * 8));
  KGP_CUDA(c, g->out.ensure(64 * 8));
  return KGP_OK;
}

static double axis_mean(const std::vector<double>& a) { double s = 0; for (double v : a) s += v; return s / (double)a.size(); }

// alpha = sW .* CG(B, sW .* ym)   (apx.m:219-222,281-302); result in g->x (scaled into alpha at the end)
static int cg_solve(kgp_ctx* c, GridState* g, const GridDesc& G, const double* X, long n, int D, double sw,
                    double tol, int maxit) {
  cudaStream_t st = c->stream;
  const int nb = nblocks_for(n);
  cg_init_kernel<<<nb, 256, 0, st>>>(g->ym.d(), sw, n, g->x.d(), g->r.d(), g->d.d(), g->partA.d());
  cg_init_scalars_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, tol, g->S.d());
  c->stats.kernel_launches += 2;
  double hS[CG_NSCAL];
  int iter = 2, parity = 0; bool done = false;
  KGP_CUDA(c, cudaMemcpyAsync(hS, g->S.p, sizeof hS, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  done = hS[CG_DONE] != 0.0;
  const int CHECK = 16;
  while (!done && iter <= maxit) {
    const int upto = std::min(maxit, iter + CHECK - 1);
    for (; iter <= upto; ++iter) {
      scatter(c, G, X, n, D, g->d.d(), sw, g->g0.d());                       // Mx' (sW d)
      double* u = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1);             // Kg ...
      for_p(G.p, [&](auto P) {                                               // z = sW Mx u + d ; d.z
        interp_gather_bz_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, X, n, D, u, sw, g->d.d(), g->z.d(), g->partA.d());
      });
      cg_update1_kernel<<<nb, 256, 0, st>>>(n, g->partA.d(), nb, g->S.d(), parity, g->x.d(), g->r.d(), g->d.d(), g->z.d(), g->partB.d());
      cg_update2_kernel<<<nb, 256, 0, st>>>(n, g->partB.d(), nb, g->S.d(), parity, tol, iter, g->r.d(), g->d.d());
      c->stats.kernel_launches += 3;
      parity ^= 1;
    }
    KGP_CUDA(c, cudaMemcpyAsync(hS, g->S.p, sizeof hS, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    done = hS[CG_DONE] != 0.0;
  }
  KGP_CUDA(c, cudaGetLastError());
  c->stats.cg_iters = (int)hS[CG_ITER]; c->stats.cg_relres = hS[CG_RELRES];
  if (!done)
    return c->fail(KGP_CG_NOT_CONVERGED, "Not converged after %d iterations, r=%1.2e", (int)hS[CG_ITER], hS[CG_RELRES]);
  return KGP_OK;
}

// ---- cell-sorted CG (grid_cg.cuh): 1-D and 2-D grids -------------------------------------------------
constexpr int CG_MAXCOLS = 64;            // right-hand sides per batched solve
constexpr int CG_CHECK = 16;              // iterations per graph launch (the host polls the flags in between)

// permutation of the points by interpolation cell + their pre-divided coordinates; cached per (data, grid)
static int cg_prepare_sorted(kgp_ctx* c, GridState* g, const GridDesc& G, int slot) {
  Slot& S = c->slots[slot];
  if (g->sort_slot == slot && g->sort_slot_ver == S.version && g->sort_grid_ver == c->grid.version && g->npad > 0) return KGP_OK;
  const long n = g->sub_n >= 0 ? g->sub_n : S.n; const int D = S.D;
  const double* Xs = S.X.d() + (g->sub_n >= 0 ? g->sub_lo * D : 0);          // rows of this rank (row-sharded solve)
  cudaStream_t st = c->stream;
  const long npad = ((n + CG_SUB - 1) / CG_SUB) * CG_SUB;
  long Npad = 1; while (Npad < n) Npad <<= 1;
  KGP_CUDA(c, g->sortk.ensure(Npad * 8)); KGP_CUDA(c, g->sorti.ensure(Npad * 4));
  KGP_CUDA(c, g->perm.ensure(n * 4)); KGP_CUDA(c, g->U.ensure(npad * G.p * 8));
  const unsigned nb = (unsigned)((Npad + 255) / 256);
  if (G.p == 1) cg_cellkey_kernel<1><<<nb, 256, 0, st>>>(G, Xs, n, D, g->sortk.d(), (int*)g->sorti.p, Npad);
  else cg_cellkey_kernel<2><<<nb, 256, 0, st>>>(G, Xs, n, D, g->sortk.d(), (int*)g->sorti.p, Npad);
  for (long k = 2; k <= Npad; k <<= 1)
    for (long j = k >> 1; j > 0; j >>= 1)
      bitonic_stage_kernel<<<nb, 256, 0, st>>>(g->sortk.d(), (int*)g->sorti.p, Npad, j, k);
  KGP_CUDA(c, cudaMemcpyAsync(g->perm.p, g->sorti.p, n * 4, cudaMemcpyDeviceToDevice, st));
  const unsigned nb2 = (unsigned)((npad + 255) / 256);
  if (G.p == 1) cg_sorted_coords_kernel<1><<<nb2, 256, 0, st>>>(G, Xs, n, D, (const int*)g->perm.p, npad, g->U.d());
  else cg_sorted_coords_kernel<2><<<nb2, 256, 0, st>>>(G, Xs, n, D, (const int*)g->perm.p, npad, g->U.d());
  KGP_CUDA(c, cudaGetLastError());
  g->npad = npad; g->sort_slot = slot; g->sort_slot_ver = S.version; g->sort_grid_ver = c->grid.version;
  return KGP_OK;
}

struct CgLaunch { GridDesc G; long n, npad, N; int m, sub, ctas; double *U, *sx, *sr, *sd, *sz, *ctl, *gm0, *gm1, *Kf, *KfT; unsigned* cnt; long koff1;
                  int big; int ngp0, ngp1; double *kp0, *kp1, *gpad, *tpad, *upad; };

// dst (rp x cp, zero padded) <- src (r x c), both row-major; and back
__global__ void pad_matrix_kernel(const double* __restrict__ src, int r, int c, double* __restrict__ dst, int rp, int cp) {
  const long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)rp * cp) return;
  const int i = (int)(q / cp), j = (int)(q - (long)i * cp);
  dst[q] = (i < r && j < c) ? src[(long)i * c + j] : 0.0;
}
__global__ void unpad_matrix_kernel(const double* __restrict__ src, int rp, int cp, double* __restrict__ dst, int r, int c) {
  const long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)r * c) return;
  const int i = (int)(q / c), j = (int)(q - (long)i * c);
  dst[q] = src[(long)i * cp + j];
}
// (K2 kron K1) g for LARGE factors (k = 1000 grids): two FP64 tensor-core GEMMs on zero-padded operands, no transposes:
//   T'[a][j1] = sum_j0 K1[a][j0] G[j1][j0]   (A = K1, B = G, both K-contiguous)   then   U[r][a] = sum_j1 K2[r][j1] T'[a][j1]
static void kron2_gemm(const CgLaunch& L, cudaStream_t st) {
  const int n0 = L.G.ng[0], n1 = L.G.ng[1], p0 = L.ngp0, p1 = L.ngp1;
  pad_matrix_kernel<<<(unsigned)(((long)p1 * p0 + 255) / 256), 256, 0, st>>>(L.gm0, n1, n0, L.gpad, p1, p0);
  GemmArgs a; memset(&a, 0, sizeof a);
  a.A = L.kp0; a.lda = p0; a.B = L.gpad; a.ldb = p0; a.C = L.tpad; a.ldc = p1; a.ntm = p0 / 128; a.ntn = p1 / 128;
  a.kbeg = 0; a.kend = p0; a.shape = GEMM_FULL; a.ktri = KTRI_NONE; a.alpha = 1.0; a.beta = 0.0; a.nb = NB;
  gemm_launch(a, st);
  GemmArgs b; memset(&b, 0, sizeof b);
  b.A = L.kp1; b.lda = p1; b.B = L.tpad; b.ldb = p1; b.C = L.upad; b.ldc = p0; b.ntm = p1 / 128; b.ntn = p0 / 128;
  b.kbeg = 0; b.kend = p1; b.shape = GEMM_FULL; b.ktri = KTRI_NONE; b.alpha = 1.0; b.beta = 0.0; b.nb = NB;
  gemm_launch(b, st);
  unpad_matrix_kernel<<<(unsigned)(((long)n1 * n0 + 255) / 256), 256, 0, st>>>(L.upad, p1, p0, L.gm1, n1, n0);
}

// one CG iteration (apx.m:290-300) as kernel launches on st
static void cg_iteration(kgp_ctx* c, GridState* g, const CgLaunch& L, cudaStream_t st) {
  const GridDesc& G = L.G;
  dim3 grid((unsigned)L.ctas, (unsigned)L.m);
  double* u;
  if (G.p == 1) {
    cg_scatter_kernel<1><<<grid, 256, 0, st>>>(G, L.U, L.npad, L.sr, L.sd, L.ctl, L.gm0);
    dim3 kg((unsigned)((G.N + 255) / 256), 1);
    for (int col = 0; col < L.m; ++col)
      mode_product_kernel<<<kg, 256, 0, st>>>(L.Kf, G.ng[0], L.gm0 + col * G.N, L.gm1 + col * G.N, 1, 1);
    u = L.gm1;
    cg_gather_kernel<1><<<grid, 256, 0, st>>>(G, L.U, L.n, L.npad, u, L.sd, L.sz, L.ctl, L.cnt);
  } else {
    cg_scatter_kernel<2><<<grid, 256, 0, st>>>(G, L.U, L.npad, L.sr, L.sd, L.ctl, L.gm0);
    if (L.big) kron2_gemm(L, st);
    else {
    int TA = (G.ng[0] + 31) & ~31; if (TA > 256) TA = 256;
    const size_t ksm = (size_t)(KRON2_THREADS / TA + 1) * G.ng[0] * 8;
    kron2_kernel<<<dim3((unsigned)G.ng[1], (unsigned)L.m), KRON2_THREADS, ksm, st>>>(L.KfT, L.Kf + L.koff1, G.ng[0], G.ng[1], L.gm0, L.gm1, L.ctl, G.N);
    }
    u = L.gm1;
    cg_gather_kernel<2><<<grid, 256, 0, st>>>(G, L.U, L.n, L.npad, u, L.sd, L.sz, L.ctl, L.cnt);
  }
  cg_update_kernel<<<grid, 256, 0, st>>>(L.npad, L.sx, L.sr, L.sd, L.sz, L.ctl, L.cnt, L.gm0, G.N);
}
static int cg_launches_per_iteration(const CgLaunch& L) { return L.G.p == 1 ? 3 + L.m : (L.big ? 7 : 4); }

// whole solve in one cooperative launch (cg_persistent_kernel); returns -1 when the configuration does not fit
template <int P, int LEVEL>
static cudaError_t launch_persistent(const CgPersist& A, int ctas, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(cg_persistent_kernel<P, LEVEL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  void* args[] = {(void*)&A};
  return cudaLaunchCooperativeKernel((const void*)cg_persistent_kernel<P, LEVEL>, dim3((unsigned)ctas), dim3(CGP_THREADS), args, smem, st);
}
static int cg_finish(kgp_ctx* c, GridState* g, int m) {
  int iters = 0, bad = -1; double rel = 0.0;
  for (int col = 0; col < m; ++col) {
    const double* ct = g->h_ctl + col * CT_STRIDE;
    iters = std::max(iters, (int)ct[CT_ITER]); if (ct[CT_RELRES] > rel) rel = ct[CT_RELRES];
    if (ct[CT_DONE] != 1.0 && bad < 0) bad = col;
  }
  c->stats.cg_iters = iters; c->stats.cg_relres = rel;
  if (bad >= 0)
    return c->fail(KGP_CG_NOT_CONVERGED, "Not converged after %d iterations, r=%1.2e", (int)g->h_ctl[bad * CT_STRIDE + CT_ITER],
                   g->h_ctl[bad * CT_STRIDE + CT_RELRES]);
  return KGP_OK;
}
static int cg_run_persistent(kgp_ctx* c, GridState* g, const CgLaunch& L, int maxit) {
  const GridDesc& G = L.G; cudaStream_t st = c->stream;
  if (!g->num_sms) {
    int v = 0; cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, c->device); g->num_sms = v;
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device); g->smem_optin = (size_t)v;
    int coop = 0; cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
    if (!coop) g->num_sms = -1;
  }
  if (getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] persistent CG: sms %d optin %zu\n", g->num_sms, g->smem_optin);
  if (g->num_sms <= 0 || G.p > 2) return -1;
  // one CTA per SM, but never more CTAs than 128-point units: the reference-sized problems (n = 460: 4 units) pay for
  // every participant of the two grid barriers per iteration, and a cooperative launch over all SMs cannot start while
  // the factor eigen-decompositions (side stream) hold two of them -- the two would serialise instead of overlapping
  const long nunits = L.npad / 128;
  static const bool small_grid = !(getenv("KGP_CG_SMALL_GRID") && atoi(getenv("KGP_CG_SMALL_GRID")) == 0);
  const int ctas = (small_grid && g->sub_n < 0 && c->nranks == 1 && nunits < g->num_sms) ? (int)std::max<long>(nunits, 1) : g->num_sms;
  const int upc = (int)((nunits + ctas - 1) / ctas);
  int TA = (G.ng[0] + 31) & ~31; if (TA > CGP_THREADS) TA = CGP_THREADS;
  const size_t kscr = (((size_t)(CGP_THREADS / TA + 1) * G.ng[0] + 1) & ~(size_t)1) * 8 + (size_t)(CGP_THREADS / 32) * CGP_WSCR * 8;
  const size_t budget = g->smem_optin - 6 * 1024;                      // static shared of the kernel (< 6 KB)
  const size_t s1 = kscr + (size_t)upc * 128 * G.p * 8, s2 = s1 + (size_t)upc * 128 * 16;
  static const int force_level = getenv("KGP_CG_LEVEL") ? atoi(getenv("KGP_CG_LEVEL")) : 2;
  int level = 0; size_t smem = kscr;
  if (kscr > budget) return -1;
  if (s1 <= budget && force_level >= 1) { level = 1; smem = s1; }
  if (L.m == 1 && s2 <= budget && force_level >= 2) { level = 2; smem = s2; }
  KGP_CUDA(c, g->pdz.ensure((size_t)L.m * ctas * 8)); KGP_CUDA(c, g->prr.ensure((size_t)L.m * ctas * 8));
  KGP_CUDA(c, g->bar.ensure(256));
  KGP_CUDA(c, cudaMemsetAsync(g->bar.p, 0, 256, st));
  CgPersist A; memset(&A, 0, sizeof A);
  A.G = G; A.U = L.U; A.n = L.n; A.npad = L.npad; A.m = L.m; A.upc = upc;
  A.x = L.sx; A.r = L.sr; A.d = L.sd; A.z = L.sz; A.g0 = L.gm0; A.ug = L.gm1;
  A.K0 = L.Kf; A.K0T = L.KfT; A.K1 = L.Kf + L.koff1;
  A.ctl = L.ctl; A.part_dz = g->pdz.d(); A.part_rr = g->prr.d(); A.bar = (unsigned*)g->bar.p;
  A.dbg = getenv("KGP_DEBUG") ? (long long*)((char*)g->bar.p + 128) : nullptr;
  cudaError_t e;
  if (G.p == 1) e = level == 2 ? launch_persistent<1, 2>(A, ctas, smem, st) : level == 1 ? launch_persistent<1, 1>(A, ctas, smem, st) : launch_persistent<1, 0>(A, ctas, smem, st);
  else e = level == 2 ? launch_persistent<2, 2>(A, ctas, smem, st) : level == 1 ? launch_persistent<2, 1>(A, ctas, smem, st) : launch_persistent<2, 0>(A, ctas, smem, st);
  KGP_CUDA(c, e);
  g->cg_level_used = level;
  c->stats.kernel_launches += 1;
  if (getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] persistent CG: ctas %d upc %d level %d smem %zu m %d\n", ctas, upc, level, smem, L.m);
  KGP_CUDA(c, cudaMemcpyAsync(g->h_ctl, g->ctl.p, (size_t)L.m * CT_STRIDE * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  (void)maxit;
  if (A.dbg) {
    long long tk[8]; cudaMemcpy(tk, A.dbg, sizeof tk, cudaMemcpyDeviceToHost);
    const double it = std::max(1.0, g->h_ctl[CT_ITER] - 1.0), f = 1e-3 / 1.965 / it;   // us per iteration at 1.965 GHz
    fprintf(stderr, "[kgp] us/iter: A %.2f bar %.2f | B %.2f bar %.2f | C %.2f bar %.2f | D %.2f bar %.2f\n", tk[0] * f, tk[1] * f,
            tk[2] * f, tk[3] * f, tk[4] * f, tk[5] * f, tk[6] * f, tk[7] * f);
  }
  return cg_finish(c, g, L.m);
}

template <int P, int LEVEL, bool DBG, bool DIST, bool FX>
static cudaError_t launch_persistent2c(const CgPersist2& A, int ctas, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(cg_persistent2_kernel<P, LEVEL, DBG, DIST, FX>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  void* args[] = {(void*)&A};
  return cudaLaunchCooperativeKernel((const void*)cg_persistent2_kernel<P, LEVEL, DBG, DIST, FX>, dim3((unsigned)ctas), dim3(CGP2_THREADS), args, smem, st);
}
template <int P, int LEVEL, bool DBG, bool DIST = false>
static cudaError_t launch_persistent2b(const CgPersist2& A, int ctas, size_t smem, cudaStream_t st, bool fx) {
  return fx ? launch_persistent2c<P, LEVEL, DBG, DIST, true>(A, ctas, smem, st) : launch_persistent2c<P, LEVEL, DBG, DIST, false>(A, ctas, smem, st);
}
template <int P, int LEVEL>
static cudaError_t launch_persistent2(const CgPersist2& A, int ctas, size_t smem, cudaStream_t st, bool fx) {
  if (A.X.nranks > 1) return A.dbg ? launch_persistent2b<P, LEVEL, true, true>(A, ctas, smem, st, fx) : launch_persistent2b<P, LEVEL, false, true>(A, ctas, smem, st, fx);
  return A.dbg ? launch_persistent2b<P, LEVEL, true>(A, ctas, smem, st, fx) : launch_persistent2b<P, LEVEL, false>(A, ctas, smem, st, fx);
}
// two-barrier formulation (cg_persistent2_kernel); -1 when it does not fit (node window / staging too large)
static int cg_run_persistent2(kgp_ctx* c, GridState* g, const CgLaunch& L, int slot) {
  const GridDesc& G = L.G; cudaStream_t st = c->stream;
  if (!g->num_sms) {
    int v = 0; cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, c->device); g->num_sms = v;
    cudaDeviceGetAttribute(&v, cudaDevAttrMaxSharedMemoryPerBlockOptin, c->device); g->smem_optin = (size_t)v;
    int coop = 0; cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device);
    if (!coop) g->num_sms = -1;
  }
  if (g->num_sms <= 0 || G.p > 2 || L.m > CG_MAXCOLS_DEV) return -1;
  // one CTA per SM, but never more CTAs than 128-point units: the reference-sized problems (n = 460: 4 units) pay for
  // every participant of the two grid barriers per iteration, and a cooperative launch over all SMs cannot start while
  // the factor eigen-decompositions (side stream) hold two of them -- the two would serialise instead of overlapping
  const long nunits = L.npad / 128;
  static const bool small_grid = !(getenv("KGP_CG_SMALL_GRID") && atoi(getenv("KGP_CG_SMALL_GRID")) == 0);
  const int ctas = (small_grid && g->sub_n < 0 && c->nranks == 1 && nunits < g->num_sms) ? (int)std::max<long>(nunits, 1) : g->num_sms;
  const int upc = (int)((nunits + ctas - 1) / ctas);
  Slot& S = c->slots[slot];
  if (g->wmax_slot != slot || g->wmax_slot_ver != S.version || g->wmax_grid_ver != c->grid.version || g->wmax_ctas != ctas) {
    KGP_CUDA(c, g->wmaxd.ensure(64));
    KGP_CUDA(c, cudaMemsetAsync(g->wmaxd.p, 0, 64, st));
    if (G.p == 1) cg_window_kernel<1><<<ctas, 32, 0, st>>>(G, L.U, L.npad, upc, (int*)g->wmaxd.p);
    else cg_window_kernel<2><<<ctas, 32, 0, st>>>(G, L.U, L.npad, upc, (int*)g->wmaxd.p);
    // bound of the fixed-point scatter (fx_scale): max over nodes of sum_i c_i^2, integer-accumulated (deterministic)
    KGP_CUDA(c, g->nodew2.ensure((size_t)G.N * 8));
    KGP_CUDA(c, cudaMemsetAsync(g->nodew2.p, 0, (size_t)G.N * 8, st));
    {
      const unsigned nb = (unsigned)std::min<long>((L.npad + 255) / 256, 2048);
      if (G.p == 1) cg_node_weight2_kernel<1><<<nb, 256, 0, st>>>(G, L.U, L.npad, (unsigned long long*)g->nodew2.p);
      else cg_node_weight2_kernel<2><<<nb, 256, 0, st>>>(G, L.U, L.npad, (unsigned long long*)g->nodew2.p);
      cg_umax_kernel<<<(unsigned)std::min<long>((G.N + 255) / 256, 256), 256, 0, st>>>((const unsigned long long*)g->nodew2.p, G.N, (unsigned long long*)((char*)g->wmaxd.p + 8));
    }
    int w = 0; unsigned long long w2max = 0;
    KGP_CUDA(c, cudaMemcpyAsync(&w, g->wmaxd.p, 4, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaMemcpyAsync(&w2max, (char*)g->wmaxd.p + 8, 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    g->c2max2 = std::max((double)w2max / 1073741824.0, 1.0);
    g->wmax = w; g->wmax_slot = slot; g->wmax_slot_ver = S.version; g->wmax_grid_ver = c->grid.version; g->wmax_ctas = ctas;
  }
  const int wcap = std::max(g->wmax, 4);
  const long ng0 = G.ng[0], ng1 = G.p > 1 ? G.ng[1] : 1;
  auto even = [](long v) { return (v + 1) & ~1L; };
  const long uwin_d = even(G.p == 1 ? wcap : wcap * ng0);
  long stg_need = (long)(CGP2_THREADS / 32) * CGP_WSCR2;
  if (G.p == 1) stg_need = std::max(stg_need, even(ng0));
  else stg_need = std::max(stg_need, even(std::max<long>(G.N, ng0 * ng0)) + even(wcap * ng1) + even(wcap * ng0));
  const size_t budget = g->smem_optin - 8704;                         // static shared of the kernel (8192 B, see ptxas.log)
  const long u_d = (long)upc * 128 * G.p, v_d = (long)upc * 128;
  const size_t s0 = (size_t)(uwin_d + stg_need) * 8, s1 = (size_t)(uwin_d + u_d + stg_need) * 8,
               s2 = (size_t)(uwin_d + u_d + v_d + std::max(v_d, stg_need)) * 8;
  if (s0 > budget) return -1;
  static const int force_level = getenv("KGP_CG_LEVEL") ? atoi(getenv("KGP_CG_LEVEL")) : 2;
  int level = 0; size_t smem = s0;
  if (s1 <= budget && force_level >= 1) { level = 1; smem = s1; }
  if (L.m == 1 && s2 <= budget && force_level >= 2) { level = 2; smem = s2; }
  KGP_CUDA(c, g->part4.ensure((size_t)4 * L.m * ctas * 8));
  const size_t fxbytes = (size_t)L.m * 2 * G.N * 8;                  // fixed-point scatter targets: (hi, lo) per column
  KGP_CUDA(c, g->g0b.ensure(2 * fxbytes));
  KGP_CUDA(c, g->bar.ensure(256));
  KGP_CUDA(c, cudaMemsetAsync(g->bar.p, 0, 256, st));
  KGP_CUDA(c, cudaMemsetAsync(g->g0b.p, 0, 2 * fxbytes, st));
  if (level < 2) KGP_CUDA(c, cudaMemsetAsync(L.sz, 0, (size_t)L.m * L.npad * 8, st));
  CgPersist2 A; memset(&A, 0, sizeof A);
  A.G = G; A.U = L.U; A.n = L.n; A.npad = L.npad; A.m = L.m; A.upc = upc; A.wcap = wcap;
  A.x = L.sx; A.r = L.sr; A.d = L.sd; A.z = L.sz;
  A.fxa = (long long*)g->g0b.p; A.fxb = (long long*)((char*)g->g0b.p + fxbytes); A.npts_bound = L.n; A.c2max2 = g->c2max2;
  A.K0 = L.Kf; A.K0T = L.KfT; A.K1 = L.Kf + L.koff1;
  A.ctl = L.ctl; A.part = g->part4.d(); A.bar = (unsigned*)g->bar.p; A.stg_doubles = stg_need;
  A.dbg = getenv("KGP_DEBUG") ? (long long*)((char*)g->bar.p + 128) : nullptr;
  if (g->sub_n >= 0) {
    if (!c->px_ok || G.N > PX_NCAP || L.m != 1) return c->fail(KGP_BAD_ARG, "row-sharded CG needs the peer window, N <= %ld, one right-hand side", PX_NCAP);
    A.X.nranks = c->nranks; A.X.rank = c->rank; A.X.epoch = c->px_epoch; A.X.ncap = PX_NCAP; A.X.loc = (char*)c->px_local;
    for (int r = 0; r < c->nranks; ++r) A.X.peer[r] = (char*)c->px_peer[r];
    A.X.abort = c->px_abort;
  }
  cudaError_t e;
  // opt.cg_deterministic (default 1): order-independent fixed-point scatter -> bit-reproducible solves; 0: FP64 atomics (faster)
  const bool fx = c->model.opt.cg_deterministic != 0;
  if (G.p == 1) e = level == 2 ? launch_persistent2<1, 2>(A, ctas, smem, st, fx) : level == 1 ? launch_persistent2<1, 1>(A, ctas, smem, st, fx) : launch_persistent2<1, 0>(A, ctas, smem, st, fx);
  else e = level == 2 ? launch_persistent2<2, 2>(A, ctas, smem, st, fx) : level == 1 ? launch_persistent2<2, 1>(A, ctas, smem, st, fx) : launch_persistent2<2, 0>(A, ctas, smem, st, fx);
  KGP_CUDA(c, e);
  g->cg_level_used = level;
  c->stats.kernel_launches += 1;
  KGP_CUDA(c, cudaMemcpyAsync(g->h_ctl, g->ctl.p, (size_t)L.m * CT_STRIDE * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  if (g->sub_n >= 0) c->px_epoch += (unsigned long long)g->h_ctl[CT_PASSES];
  if (A.dbg) {
    long long tk[8]; cudaMemcpy(tk, A.dbg, sizeof tk, cudaMemcpyDeviceToHost);
    const double it = std::max(1.0, g->h_ctl[CT_ITER] - 1.0), f = 1e-3 / 1.965 / it;
    fprintf(stderr, "[kgp] persistent2: level %d smem %zu wcap %d upc %d | us/iter: F1 %.2f F2 %.2f bar %.2f | Cwin %.2f Cloop %.2f bar %.2f | red %.2f\n", level, smem, wcap, upc,
            tk[5] * f, tk[0] * f, tk[1] * f, tk[6] * f, tk[2] * f, tk[3] * f, tk[4] * f);
  }
  return cg_finish(c, g, L.m);
}

// x = CG(B, sw * rhs[:, col]) for m columns at once (sorted order, result in g->sx); rhs in ORIGINAL point order
static int cg_solve_sorted(kgp_ctx* c, GridState* g, const GridDesc& G, int slot, const double* rhs, long rhs_stride, int m,
                           double sw, double tol, int maxit) {
  if (m < 1 || m > CG_MAXCOLS) return c->fail(KGP_BAD_ARG, "CG batch of %d columns (max %d)", m, CG_MAXCOLS);
  int rc = cg_prepare_sorted(c, g, G, slot); if (rc) return rc;
  Slot& S = c->slots[slot];
  const bool sharded = g->sub_n >= 0;
  const long n = sharded ? g->sub_n : S.n, npad = g->npad;
  if (sharded) { rhs += g->sub_lo; if (m != 1) return c->fail(KGP_BAD_ARG, "row-sharded CG handles one right-hand side"); }
  cudaStream_t st = c->stream;
  DevBuf* vecs[] = {&g->sx, &g->sr, &g->sd, &g->sz};
  for (DevBuf* b : vecs) KGP_CUDA(c, b->ensure((size_t)m * npad * 8));
  KGP_CUDA(c, g->gm0.ensure((size_t)m * G.N * 8)); KGP_CUDA(c, g->gm1.ensure((size_t)m * G.N * 8));
  KGP_CUDA(c, g->ctl.ensure(CG_MAXCOLS * CT_STRIDE * 8)); KGP_CUDA(c, g->cnt.ensure(CG_MAXCOLS * 4 * 4));
  if (!g->h_ctl) KGP_CUDA(c, cudaMallocHost((void**)&g->h_ctl, CG_MAXCOLS * CT_STRIDE * 8));
  CgLaunch L; memset(&L, 0, sizeof L);
  L.G = G; L.n = n; L.npad = npad; L.N = G.N; L.m = m;
  L.sub = 0; L.ctas = (int)std::min<long>(npad / CG_SUB, 148 * 3);     // warps stride over 128-point units
  L.U = g->U.d(); L.sx = g->sx.d(); L.sr = g->sr.d(); L.sd = g->sd.d(); L.sz = g->sz.d(); L.ctl = g->ctl.d();
  L.gm0 = g->gm0.d(); L.gm1 = g->gm1.d(); L.Kf = g->Kf.d(); L.KfT = g->KfT.d(); L.cnt = (unsigned*)g->cnt.p;
  L.koff1 = G.p > 1 ? g->koff[1] : 0;

  KGP_CUDA(c, cudaMemsetAsync(g->cnt.p, 0, CG_MAXCOLS * 4 * 4, st));
  KGP_CUDA(c, cudaMemsetAsync(g->ctl.p, 0, CG_MAXCOLS * CT_STRIDE * 8, st));
  KGP_CUDA(c, cudaMemsetAsync(g->gm0.p, 0, (size_t)m * G.N * 8, st));
  cg_init_sorted_kernel<<<dim3((unsigned)std::min<long>(npad / 256, 1184), (unsigned)m), 256, 0, st>>>(
      rhs, rhs_stride, (const int*)g->perm.p, n, npad, sw, tol, maxit, L.sx, L.sr, L.sd, L.ctl, L.cnt);
  c->stats.kernel_launches += 1;
  if (sharded) {                                               // |b|^2 over all ranks (also the per-solve rendezvous)
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(L.ctl + CT_RRCUR, L.ctl + CT_RRCUR, 1, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
    cg_dist_fix_ctl_kernel<<<1, 1, 0, st>>>(L.ctl);
    rc = cg_run_persistent2(c, g, L, slot);
    if (rc == -1) return c->fail(KGP_BAD_ARG, "row-sharded CG: the persistent kernel does not fit this grid");
    return rc;
  }

  // large Kronecker factors (k = 1000 grids): the dense mode products are 4 ng^3 flops per MVM -- they go to the FP64
  // tensor-core GEMM (streaming iteration replayed as a CUDA graph) instead of the in-kernel scalar products
  static const int big_min = getenv("KGP_KRON_GEMM_MIN") ? atoi(getenv("KGP_KRON_GEMM_MIN")) : 192;
  if (G.p == 2 && m == 1 && G.ng[0] >= big_min && G.ng[1] >= big_min) {
    L.big = 1; L.ngp0 = (G.ng[0] + 127) / 128 * 128; L.ngp1 = (G.ng[1] + 127) / 128 * 128;
    KGP_CUDA(c, g->kp0.ensure((size_t)L.ngp0 * L.ngp0 * 8)); KGP_CUDA(c, g->kp1.ensure((size_t)L.ngp1 * L.ngp1 * 8));
    KGP_CUDA(c, g->gpad.ensure((size_t)L.ngp1 * L.ngp0 * 8)); KGP_CUDA(c, g->tpad.ensure((size_t)L.ngp0 * L.ngp1 * 8));
    KGP_CUDA(c, g->upad.ensure((size_t)L.ngp1 * L.ngp0 * 8));
    L.kp0 = g->kp0.d(); L.kp1 = g->kp1.d(); L.gpad = g->gpad.d(); L.tpad = g->tpad.d(); L.upad = g->upad.d();
    pad_matrix_kernel<<<(unsigned)(((long)L.ngp0 * L.ngp0 + 255) / 256), 256, 0, st>>>(L.Kf, G.ng[0], G.ng[0], L.kp0, L.ngp0, L.ngp0);
    pad_matrix_kernel<<<(unsigned)(((long)L.ngp1 * L.ngp1 + 255) / 256), 256, 0, st>>>(L.Kf + L.koff1, G.ng[1], G.ng[1], L.kp1, L.ngp1, L.ngp1);
    c->stats.kernel_launches += 2;
  }
  static const int cg_mode = getenv("KGP_CG_MODE") ? atoi(getenv("KGP_CG_MODE")) : 3;     // 3 persistent two-barrier, 2 persistent four-phase, 1 streaming kernels
  if (cg_mode >= 3 && !L.big) {
    rc = cg_run_persistent2(c, g, L, slot);
    if (rc != -1) return rc;
  }
  if (cg_mode >= 2 && !L.big) {
    rc = cg_run_persistent(c, g, L, maxit);
    if (rc != -1) return rc;                                  // -1: does not fit -> streaming kernels below
  }
  static const bool use_graph = !(getenv("KGP_CG_GRAPH") && atoi(getenv("KGP_CG_GRAPH")) == 0);
  if (use_graph) {
    std::vector<char> key(sizeof L); memcpy(key.data(), &L, sizeof L);
    if (!g->cg_graph || key != g->cg_graph_key) {
      if (g->cg_graph) { cudaGraphExecDestroy(g->cg_graph); g->cg_graph = nullptr; }
      cudaGraph_t graph = nullptr;
      KGP_CUDA(c, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      for (int it = 0; it < CG_CHECK; ++it) cg_iteration(c, g, L, st);
      KGP_CUDA(c, cudaStreamEndCapture(st, &graph));
      KGP_CUDA(c, cudaGraphInstantiate(&g->cg_graph, graph, 0));
      cudaGraphDestroy(graph);
      g->cg_graph_key = key;
    }
  }
  bool done = false; int launched = 0;
  const int lim = maxit + CG_CHECK;       // the device stops itself at maxit; this only bounds the host loop
  while (!done && launched < lim) {
    if (use_graph) KGP_CUDA(c, cudaGraphLaunch(g->cg_graph, st));
    else for (int it = 0; it < CG_CHECK; ++it) cg_iteration(c, g, L, st);
    launched += CG_CHECK;
    c->stats.kernel_launches += (long)CG_CHECK * cg_launches_per_iteration(L);
    KGP_CUDA(c, cudaMemcpyAsync(g->h_ctl, g->ctl.p, (size_t)m * CT_STRIDE * 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    done = true;
    for (int col = 0; col < m; ++col) if (g->h_ctl[col * CT_STRIDE + CT_DONE] == 0.0) done = false;
  }
  KGP_CUDA(c, cudaGetLastError());
  int iters = 0, bad = -1; double rel = 0.0;
  for (int col = 0; col < m; ++col) {
    const double* ct = g->h_ctl + col * CT_STRIDE;
    iters = std::max(iters, (int)ct[CT_ITER]); if (ct[CT_RELRES] > rel) rel = ct[CT_RELRES];
    if (ct[CT_DONE] != 1.0 && bad < 0) bad = col;
  }
  c->stats.cg_iters = iters; c->stats.cg_relres = rel;
  if (bad >= 0)
    return c->fail(KGP_CG_NOT_CONVERGED, "Not converged after %d iterations, r=%1.2e", (int)g->h_ctl[bad * CT_STRIDE + CT_ITER],
                   g->h_ctl[bad * CT_STRIDE + CT_RELRES]);
  return KGP_OK;
}

// ---- Woodbury-preconditioned CG (grid_precond.cuh) --------------------------------------------
static double dbg_now(cudaStream_t st) { cudaStreamSynchronize(st); timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
static int precond_setup(kgp_ctx* c, GridState* g, const GridDesc& G, int slot, double s2) {
  Slot& S = c->slots[slot]; cudaStream_t st = c->stream;
  const bool dbg = getenv("KGP_DEBUG") != nullptr; double tq[8] = {0};
  if (dbg) tq[0] = dbg_now(st);
  const long N = G.N, n = S.n;
  if (G.p > 2 || N > 16384) return c->fail(KGP_BAD_ARG, "opt.cg_precond needs a 1-D/2-D grid with N <= 16384 (N = %ld)", N);
  int rc = cg_prepare_sorted(c, g, G, slot); if (rc) return rc;
  if (g->gram_slot != slot || g->gram_slot_ver != S.version || g->gram_grid_ver != c->grid.version) {
    KGP_CUDA(c, g->Gm.ensure((size_t)N * N * 8));
    KGP_CUDA(c, cudaMemsetAsync(g->Gm.p, 0, (size_t)N * N * 8, st));
    const unsigned nb = (unsigned)std::min<long>((g->npad / 32 + 7) / 8, 148L * 16);
    if (G.p == 1) gram_accum_kernel<1><<<nb, 256, 0, st>>>(G, g->U.d(), n, g->npad, g->Gm.d());
    else gram_accum_kernel<2><<<nb, 256, 0, st>>>(G, g->U.d(), n, g->npad, g->Gm.d());
    g->gram_slot = slot; g->gram_slot_ver = S.version; g->gram_grid_ver = c->grid.version;
    c->stats.kernel_launches++;
  }
  long ksz = 0; for (int d = 0; d < G.p; ++d) ksz += (long)G.ng[d] * G.ng[d];
  KGP_CUDA(c, g->VT.ensure(ksz * 8)); KGP_CUDA(c, g->se.ensure(N * 8));
  KGP_CUDA(c, g->Ht1.ensure((size_t)N * N * 8)); KGP_CUDA(c, g->Ht2.ensure((size_t)N * N * 8));
  KGP_CUDA(c, g->ps.ensure(n * 8)); KGP_CUDA(c, g->pw.ensure(N * 8));
  if (dbg) tq[1] = dbg_now(st);
  cudaStreamWaitEvent(st, c->ev[11], 0);                         // factor eigen-decompositions (side stream)
  if (dbg) tq[2] = dbg_now(st);
  for (int d = 0; d < G.p; ++d) {
    const long nn = (long)G.ng[d] * G.ng[d];
    transpose_small_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(g->Vf.d() + g->koff[d], G.ng[d], g->VT.d() + g->koff[d]);
  }
  sqrt_kron_eig_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(G, g->ee.d(), (const int*)g->eoff.p, g->se.d());
  // kept modes: s^2 e_k n/N > tau  (host: the factor spectra are ng_0 + ng_1 numbers)
  {
    static const double tau = getenv("KGP_PRECOND_TAU") ? atof(getenv("KGP_PRECOND_TAU")) : 1e-3;
    int etot = 0; for (int d = 0; d < G.p; ++d) etot += G.ng[d];
    std::vector<double> he(etot);
    KGP_CUDA(c, cudaMemcpyAsync(he.data(), g->ee.p, (size_t)etot * 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    std::vector<int> sel; sel.reserve(N);
    const double scale = s2 * (double)n / (double)N;
    for (long q = 0; q < N; ++q) {
      long rr = q; double v = 1.0;
      for (int d = 0; d < G.p; ++d) { v *= std::max(he[g->eoff_h[d] + (int)(rr % G.ng[d])], 0.0); rr /= G.ng[d]; }
      if (scale * v > tau) sel.push_back((int)q);
    }
    if (sel.empty()) sel.push_back(0);
    g->prank = (long)sel.size();
    KGP_CUDA(c, g->psel.ensure((size_t)N * 4));
    KGP_CUDA(c, cudaMemcpyAsync(g->psel.p, sel.data(), sel.size() * 4, cudaMemcpyHostToDevice, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));                      // (sel is a local)
  }
  // H = V' G V : mode products with V_d' along the 2p axes of the N x N array (column axes, then row axes)
  const double* src = g->Gm.d(); double* bufs[2] = {g->Ht1.d(), g->Ht2.d()}; int which = 0;
  const long tot = N * N;
  for (int side = 0; side < 2; ++side)
    for (int d = 0; d < G.p; ++d) {
      const long inner = (side == 0 ? 1 : N) * G.stride[d], outer = tot / (inner * G.ng[d]);
      double* dst = bufs[which];
      constexpr int RB = 10;
      const size_t msm = (size_t)((G.ng[d] + RB - 1) / RB) * RB * G.ng[d] * 8;
      if (msm <= 96 * 1024)
        mode_product_rb_kernel<RB><<<148 * 8, 256, msm, st>>>(g->VT.d() + g->koff[d], G.ng[d], src, dst, inner, outer);
      else
        mode_product_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(g->VT.d() + g->koff[d], G.ng[d], src, dst, inner, outer);
      src = dst; which ^= 1;
      c->stats.kernel_launches++;
    }
  if (dbg) tq[3] = dbg_now(st);
  // C = I + s2 E^1/2 H E^1/2 -> blocked Cholesky and triangular inverse of the exact path
  rc = ensure_exact_buffers(c, g->prank, 1); if (rc) return rc;
  const long np = c->np;
  c->st_level = -1;
  int big = 1 << 30;
  KGP_CUDA(c, cudaMemcpyAsync(c->d_info, &big, sizeof(int), cudaMemcpyHostToDevice, st));
  assemble_c_sel_kernel<<<dim3((unsigned)((np + 255) / 256), (unsigned)np), 256, 0, st>>>(src, g->se.d(), s2, N, (const int*)g->psel.p, g->prank, np, c->M.d());
  potrf(c);
  if (dbg) tq[4] = dbg_now(st);
  zsweep(c);
  if (dbg) { tq[5] = dbg_now(st); fprintf(stderr, "[kgp] precond setup ms: gram %.2f eigwait %.2f modeprod %.2f potrf %.2f zsweep %.2f (kept %ld of %ld modes)\n", tq[1] - tq[0], tq[2] - tq[1], tq[3] - tq[2], tq[4] - tq[3], tq[5] - tq[4], g->prank, N); }
  int info = 0;
  KGP_CUDA(c, cudaMemcpyAsync(&info, c->d_info, sizeof(int), cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  if (info != (1 << 30)) return c->fail(KGP_NOT_PD, "preconditioner: grid-space matrix not positive definite (pivot %d)", info);
  return KGP_OK;
}
// s = B~^-1 r = r - s2 M F C^-1 F' M' r ; partial sums of r.s in partA (nb entries)
static int precond_apply(kgp_ctx* c, GridState* g, const GridDesc& G, const Slot& S, double s2, const double* r, double* sout, int nb) {
  cudaStream_t st = c->stream; const long N = G.N, np = c->np, n = S.n; const int D = S.D;
  scatter(c, G, S.X.d(), n, D, r, 1.0, g->g0.d());                                    // M' r
  double* t = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1, g->VT.d());              // V' .
  gather_sel_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(t, g->se.d(), (const int*)g->psel.p, g->prank, np, c->ym.d());   // E^1/2 . on the kept modes
  const int nch = (int)((np + 511) / 512);
  KGP_CUDA(c, c->parts.ensure((size_t)nch * np * 8));
  gemv_t_upper_kernel<<<dim3((unsigned)(np / 128), (unsigned)nch), 128, 0, st>>>(c->W2.d(), np, np, c->ym.d(), c->parts.d());
  sum_parts_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(c->parts.d(), np, nch, 1.0, c->tvec.d());
  gemv_n_upper_kernel<<<(unsigned)((np + 7) / 8), 256, 0, st>>>(c->W2.d(), np, np, c->tvec.d(), 1.0, c->alpha.d());   // C^-1 = Z Z'
  KGP_CUDA(c, cudaMemsetAsync(g->g0.p, 0, N * 8, st));
  scatter_sel_kernel<<<(unsigned)((g->prank + 255) / 256), 256, 0, st>>>(c->alpha.d(), g->se.d(), (const int*)g->psel.p, g->prank, g->g0.d());   // E^1/2 .
  double* w = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1, g->Vf.d());              // V .
  for_p(G.p, [&](auto P) {
    interp_gather_bz_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, S.X.d(), n, D, w, -s2, r, sout, g->partA.d());   // r - s2 M w ; r.s
  });
  c->stats.kernel_launches += 6;
  return KGP_OK;
}
// alpha = sW CG(B, sW ym) with the Woodbury preconditioner; same stopping rule and iteration count as apx.m:281-302
static int pcg_solve(kgp_ctx* c, GridState* g, const GridDesc& G, int slot, double sw, double tol, int maxit,
                     const double* rhs, double* x, bool setup) {
  Slot& S = c->slots[slot]; cudaStream_t st = c->stream; const long n = S.n; const int D = S.D;
  int rc = setup ? precond_setup(c, g, G, slot, sw * sw) : KGP_OK; if (rc) return rc;
  if (getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] pcg setup done at %.2f ms\n", dbg_now(st));
  const int nb = nblocks_for(n);
  double h[4];
  auto fetch = [&](const double* dev, int cnt) -> int {
    KGP_CUDA(c, cudaMemcpyAsync(h, dev, cnt * 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    return KGP_OK;
  };
  double* r = g->r.d(); double* d = g->d.d(); double* z = g->z.d(); double* s = g->ps.d();
  cg_init_kernel<<<nb, 256, 0, st>>>(rhs, sw, n, x, r, d, g->partA.d());                         // x = 0, r = d = b
  sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->out.d() + 8);
  rc = fetch(g->out.d() + 8, 1); if (rc) return rc;
  const double nbn = sqrt(h[0]);
  int iter = 1; double relres = nbn / nbn; bool done = !(relres >= tol);
  double rz = 0.0;
  if (!done) {
    rc = precond_apply(c, g, G, S, sw * sw, r, s, nb); if (rc) return rc;
    sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->out.d() + 8);
    KGP_CUDA(c, cudaMemcpyAsync(d, s, n * 8, cudaMemcpyDeviceToDevice, st));
    rc = fetch(g->out.d() + 8, 1); if (rc) return rc;
    rz = h[0];
  }
  for (iter = 2; !done && iter <= maxit; ++iter) {
    scatter(c, G, S.X.d(), n, D, d, sw, g->g0.d());                                        // z = B d
    double* u = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1);
    for_p(G.p, [&](auto P) {
      interp_gather_bz_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, S.X.d(), n, D, u, sw, d, z, g->partA.d());
    });
    sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->out.d() + 8);
    rc = fetch(g->out.d() + 8, 1); if (rc) return rc;
    const double a = rz / h[0];
    pcg_update_kernel<<<nb, 256, 0, st>>>(n, a, x, r, d, z, g->partB.d());
    sum_small_kernel<<<1, 256, 0, st>>>(g->partB.d(), nb, g->out.d() + 9);
    rc = fetch(g->out.d() + 9, 1); if (rc) return rc;
    relres = sqrt(h[0]) / nbn;
    c->stats.kernel_launches += 5;
    if (!(relres >= tol)) { done = true; break; }
    rc = precond_apply(c, g, G, S, sw * sw, r, s, nb); if (rc) return rc;
    sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->out.d() + 8);
    rc = fetch(g->out.d() + 8, 1); if (rc) return rc;
    const double rzn = h[0];
    pcg_dir_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, rzn / rz, s, d);
    rz = rzn;
  }
  if (getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] pcg iterations done at %.2f ms\n", dbg_now(st));
  c->stats.cg_iters = std::min(iter, maxit); c->stats.cg_relres = relres;
  if (!done) return c->fail(KGP_CG_NOT_CONVERGED, "Not converged after %d iterations, r=%1.2e", std::min(iter, maxit), relres);
  axpy_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, sw, x, x);              // alpha = sW q
  return KGP_OK;
}

// ---- Chebyshev-Hutchinson log-det and its derivatives (grid_cheby.cuh; apx.m:336-391) --------------------------
__global__ void sum_scaled_acc_kernel(const double* __restrict__ part, int cnt, double coef, double* __restrict__ out) {
  __shared__ double sh[32];
  double t = sum_partials(part, cnt, sh);
  if (threadIdx.x == 0) out[0] += coef * t;
}
// y = Kg x on grid vectors without clobbering x (kron_apply ping-pongs between its two buffers)
static void kron_to(kgp_ctx* c, GridState* g, const GridDesc& G, const double* x, double* y, int skip) {
  cudaMemcpyAsync(g->g0.p, x, G.N * 8, cudaMemcpyDeviceToDevice, c->stream);
  double* r = kron_apply(c, g, G, g->g0.d(), g->g1.d(), skip);
  cudaMemcpyAsync(y, r, G.N * 8, cudaMemcpyDeviceToDevice, c->stream);
}
static int cheby_logdet(kgp_ctx* c, GridState* g, const GridDesc& G, int slot, double sw, bool want_deriv, int kind, int Q, int nh) {
  (void)kind; (void)Q; (void)nh;
  Slot& S = c->slots[slot]; Model& mo = c->model;
  const long n = S.n; const int D = S.D; cudaStream_t st = c->stream;
  const int d = mo.opt.ldB2_cheby_degree > 0 ? mo.opt.ldB2_cheby_degree : 15;
  if (d > 48) return c->fail(KGP_BAD_ARG, "ldB2_cheby_degree %d too large (max 48)", d);
  int m = mo.opt.ldB2_cheby_hutch > 0 ? mo.opt.ldB2_cheby_hutch : 10;
  if (g->probes_injected) {
    if (g->probes_n != n) return c->fail(KGP_BAD_ARG, "injected probes have %ld rows, data has %ld", g->probes_n, n);
    m = g->probes_m;
  } else {
    KGP_CUDA(c, g->probes.ensure((size_t)m * n * 8));
    rademacher_kernel<<<(unsigned)(((long)m * n + 255) / 256), 256, 0, st>>>(mo.opt.ldB2_cheby_seed, n, m, g->probes.d());
    g->probes_n = n; g->probes_m = m;
  }
  long ksz = 0; for (int q = 0; q < G.p; ++q) ksz += (long)G.ng[q] * G.ng[q];

  KGP_CUDA(c, g->chC.ensure((size_t)2 * d * d * 8)); KGP_CUDA(c, g->chP.ensure((size_t)ksz * 8));
  KGP_CUDA(c, g->chacc.ensure(64));
  // the recurrence runs on the cell-sorted point set of the CG path when there is one (1-D / 2-D equispaced grids, all rows
  // of the slot active): sorted_scatter_kernel / cheby_step_sorted_kernel; otherwise on the points as given
  static const bool sorted_ok = !(getenv("KGP_CHEBY_SORTED") && atoi(getenv("KGP_CHEBY_SORTED")) == 0);
  bool sorted = sorted_ok && G.p <= 2 && !g->any_neq && g->sub_n < 0;
  if (sorted) { int rc = cg_prepare_sorted(c, g, G, slot); if (rc) return rc; sorted = g->sub_n < 0 && g->npad >= n; }
  const long nv = sorted ? g->npad : n;                                     // length of a Krylov vector
  // probes of this rank; on the sorted set MB of them advance together (blockIdx.y): the grid-sized middle of a step
  // (memset, Kronecker product, scalar accumulation) is launch latency, shared by the batch
  int lo = 0, hi = m;
  if (c->nranks > 1) { const int base = m / c->nranks, extra = m % c->nranks; lo = c->rank * base + std::min(c->rank, extra); hi = lo + base + (c->rank < extra ? 1 : 0); }
  static const int mb_env = getenv("KGP_CHEBY_BATCH") ? atoi(getenv("KGP_CHEBY_BATCH")) : 8;
  int MB = sorted ? std::max(1, std::min(std::min(mb_env, 16), hi - lo)) : 1;
  while (MB > 1 && (size_t)MB * nv * 8 * 3 > ((size_t)3 << 29)) MB /= 2;     // three vector buffers: keep them under 1.5 GB
  KGP_CUDA(c, g->chT.ensure((size_t)MB * d * G.N * 8)); KGP_CUDA(c, g->chU.ensure((size_t)MB * d * G.N * 8));
  KGP_CUDA(c, g->bva.ensure((size_t)MB * nv * 8)); KGP_CUDA(c, g->bvb.ensure((size_t)MB * nv * 8)); KGP_CUDA(c, g->bvs.ensure((size_t)MB * nv * 8));
  const int nb = nblocks_for(n);
  KGP_CUDA(c, g->partA.ensure((size_t)std::max(2048, MB * nb) * 8));
  const long gst = (long)d * G.N;                                           // grid vectors of one probe: [d][N]
  KGP_CUDA(c, g->chY.ensure((size_t)MB * gst * 8)); KGP_CUDA(c, g->chS.ensure((size_t)MB * d * d * 8));
  KGP_CUDA(c, g->chZ.ensure((size_t)MB * gst * 8)); KGP_CUDA(c, g->chZ2.ensure((size_t)MB * gst * 8));
  KGP_CUDA(c, g->partB.ensure((size_t)2048 * 8));
  auto dot_to = [&](const double* a, const double* b2, long cnt, double* out) {   // out[0] = a.b, two-stage (deterministic)
    const int nbd = (int)std::min<long>(std::max<long>((cnt + 1023) / 1024, 1), 1184);
    dot_part_kernel<<<nbd, 256, 0, st>>>(a, b2, cnt, g->partB.d());
    sum_small_kernel<<<1, 256, 0, st>>>(g->partB.d(), nbd, out);
  };
  const bool k2 = G.p == 2 && std::max(G.ng[0], G.ng[1]) < 192;
  auto apply_step = [&](const double* bv, const double* v, double c1, double c2, double* Tj, double* Uj, double* out, int mb) {
    if (!sorted) {
      scatter(c, G, S.X.d(), n, D, bv, sw, Tj);                             // T = Mx'(sW bv)
      kron_to(c, g, G, Tj, Uj, -1);                                         // U = Kg T
      for_p(G.p, [&](auto P) {
        cheby_step_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, S.X.d(), n, D, Uj, bv, v, c1, c2, out, g->partA.d());
      });
      c->stats.kernel_launches += 1;
      return;
    }
    // (the T_j of the batch were zeroed together before the recurrence; mb = 1 callers zero theirs)
    const dim3 gp((unsigned)nb, (unsigned)mb);
    if (G.p == 1) sorted_scatter_kernel<1><<<gp, 256, 0, st>>>(G, g->U.d(), g->npad, bv, sw, Tj, nv, gst);
    else sorted_scatter_kernel<2><<<gp, 256, 0, st>>>(G, g->U.d(), g->npad, bv, sw, Tj, nv, gst);
    if (G.p == 1) {
      for (int i = 0; i < mb; ++i)
        mode_product_kernel<<<(unsigned)((G.N + 255) / 256), 256, 0, st>>>(g->Kf.d(), G.ng[0], Tj + i * gst, Uj + i * gst, 1, 1);
    } else if (k2) {
      int TA = (G.ng[0] + 31) & ~31; if (TA > 256) TA = 256;
      const size_t ksm = (size_t)(KRON2_THREADS / TA + 1) * G.ng[0] * 8;
      kron2_kernel<<<dim3((unsigned)G.ng[1], (unsigned)mb), KRON2_THREADS, ksm, st>>>(g->KfT.d(), g->Kf.d() + g->koff[1], G.ng[0], G.ng[1], Tj, Uj, nullptr, gst);
    } else {
      for (int i = 0; i < mb; ++i) kron_to(c, g, G, Tj + i * gst, Uj + i * gst, -1);
    }
    if (G.p == 1) cheby_step_sorted_kernel<1><<<gp, 256, 0, st>>>(G, g->U.d(), n, g->npad, Uj, bv, v, c1, c2, out, g->partA.d(), nv, gst);
    else cheby_step_sorted_kernel<2><<<gp, 256, 0, st>>>(G, g->U.d(), n, g->npad, Uj, bv, v, c1, c2, out, g->partA.d(), nv, gst);
    c->stats.kernel_launches += 3;
  };
  // probes i .. i+mb-1 in the order the recurrence works in
  auto probe_vec = [&](int i, int mb) -> const double* {
    const double* v = g->probes.d() + (long)i * n;
    if (!sorted) return v;
    permute_in_kernel<<<dim3((unsigned)nb, (unsigned)mb), 256, 0, st>>>(v, (const int*)g->perm.p, n, g->npad, g->bvs.d());
    return g->bvs.d();
  };
  // ---- emax: injected, or power iteration on B = I + sW K sW   (the reference: eigs(B,1,'lm'), apx.m:344-350)
  double emax = g->probes_injected ? g->probes_emax : 0.0;
  if (!(emax > 0.0)) {
    const int its = mo.opt.ldB2_cheby_maxit > 0 ? mo.opt.ldB2_cheby_maxit : 50;
    double nrm2 = (double)n, lam = 1.0;
    KGP_CUDA(c, cudaMemcpyAsync(g->bva.p, probe_vec(0, 1), nv * 8, cudaMemcpyDeviceToDevice, st));
    double* cur = g->bva.d(); double* nxt = g->bvb.d();
    for (int it = 0; it < its; ++it) {
      if (sorted) cudaMemsetAsync(g->chT.p, 0, G.N * 8, st);
      apply_step(cur, cur, 1.0, sw, g->chT.d(), g->chU.d(), nxt, 1);         // nxt = B cur ; partials of cur.nxt
      sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->chacc.d() + 2);
      dot_to(nxt, nxt, n, g->chacc.d() + 3);
      double h2[2];
      KGP_CUDA(c, cudaMemcpyAsync(h2, g->chacc.d() + 2, 16, cudaMemcpyDeviceToHost, st));
      KGP_CUDA(c, cudaStreamSynchronize(st));
      lam = h2[0] / nrm2;
      vec_scale_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, 1.0 / sqrt(h2[1]), nxt, nxt);
      nrm2 = 1.0; std::swap(cur, nxt);
    }
    emax = lam;
  }
  // ---- polynomial coefficients exactly as apx.m:356-372
  const double emin = 1.0, a = emin + emax, dl = 1.0 - emax / a, sf = 2.0 / a / (1.0 - 2.0 * dl), cI = 1.0 / (1.0 - 2.0 * dl);
  std::vector<double> xk(d + 1), fk(d + 1), cc(d + 1, 0.0), Tk((size_t)(d + 1) * (d + 1));
  for (int i = 0; i <= d; ++i) { xk[i] = cos(M_PI * ((double)i + 0.5) / (double)(d + 1)); fk[i] = log(((1.0 - 2.0 * dl) / 2.0) * xk[i] + 0.5); }
  for (int r = 0; r <= d; ++r) {
    Tk[(size_t)r * (d + 1)] = 1.0; if (d >= 1) Tk[(size_t)r * (d + 1) + 1] = xk[r];
    for (int i = 2; i <= d; ++i) Tk[(size_t)r * (d + 1) + i] = 2.0 * xk[r] * Tk[(size_t)r * (d + 1) + i - 1] - Tk[(size_t)r * (d + 1) + i - 2];
  }
  for (int i = 0; i <= d; ++i) { double sacc = 0.0; for (int r = 0; r <= d; ++r) sacc += Tk[(size_t)r * (d + 1) + i] * fk[r]; cc[i] = 2.0 / (double)(d + 1) * sacc; }
  cc[0] /= 2.0;
  std::vector<double> p1(d + 1, 0.0), p2(d + 1, 0.0), p3(d + 1), pp(d + 1);
  p1[0] = 1.0; p2[1] = 1.0;
  for (int i = 0; i <= d; ++i) pp[i] = cc[0] * p1[i] + cc[1] * p2[i];
  for (int i = 2; i <= d; ++i) {
    p3[0] = -p1[0]; for (int q = 1; q <= d; ++q) p3[q] = 2.0 * p2[q - 1] - p1[q];
    for (int q = 0; q <= d; ++q) pp[q] += cc[i] * p3[q];
    p1 = p2; p2 = p3;
  }
  // coefficient tables of the grid-space derivative forms
  std::vector<double> hC((size_t)2 * d * d, 0.0);             // [0]: C[u][s] for dhyp ; [1]: Wd[a][b] for sum(dW)
  for (int j = 1; j <= d; ++j) {
    const int kmax = (j + 1) / 2;
    for (int k = 1; k <= kmax; ++k) {
      double akj = sf * pp[j] / (double)m;
      if (k == kmax && (j % 2) == 1) akj /= 2.0;
      hC[(size_t)(j - k) * d + (k - 1)] += akj;
    }
    const double cj = sf * pp[j] / (4.0 * (double)m) / (sw * sw);
    for (int l = 0; l < j; ++l) { hC[(size_t)d * d + (size_t)(j - 1 - l) * d + l] += cj; hC[(size_t)d * d + (size_t)l * d + (j - 1 - l)] += cj; }
  }
  KGP_CUDA(c, cudaMemcpyAsync(g->chC.p, hC.data(), hC.size() * 8, cudaMemcpyHostToDevice, st));
  KGP_CUDA(c, cudaMemsetAsync(g->chP.p, 0, ksz * 8, st));
  KGP_CUDA(c, cudaMemsetAsync(g->chacc.p, 0, 64, st));
  // ---- probes of this rank, MB at a time
  for (int i0 = lo; i0 < hi; i0 += MB) {
    const int mb = std::min(MB, hi - i0);
    const double* v = probe_vec(i0, mb);
    dot_to(v, v, sorted ? (long)mb * nv : n, g->chacc.d() + 2);                                          // (padding entries are zero)
    sum_scaled_acc_kernel<<<1, 256, 0, st>>>(g->chacc.d() + 2, 1, pp[0] / (2.0 * m), g->chacc.d());      // j = 0 term: v'v p_0
    if (sorted) KGP_CUDA(c, cudaMemsetAsync(g->chT.p, 0, (size_t)mb * gst * 8, st));
    const double* cur = v; double* bufs[2] = {g->bva.d(), g->bvb.d()};
    for (int j = 0; j < d; ++j) {
      double* nxt = bufs[j & 1];
      apply_step(cur, v, sf - cI, sf * sw, g->chT.d() + (long)j * G.N, g->chU.d() + (long)j * G.N, nxt, mb);   // Bs cur
      sum_scaled_acc_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb * mb, pp[j + 1] / (2.0 * m), g->chacc.d());
      cur = nxt;
    }
    c->stats.kernel_launches += 2 + d;
    if (want_deriv) {
      // the whole batch at once: [mb][d][N] contiguous is mb d times as many outer slices of every mode
      const int nd = mb * d;
      gram_small_kernel<<<dim3(d, d, mb), 256, 0, st>>>(g->chT.d(), g->chU.d(), G.N, d, g->chS.d());
      dot_tiled_kernel<<<1, 1024, 0, st>>>(g->chS.d(), g->chC.d() + (long)d * d, d * d, mb, g->chacc.d() + 2);
      sum_scaled_acc_kernel<<<1, 256, 0, st>>>(g->chacc.d() + 2, 1, 1.0, g->chacc.d() + 1);
      mix_small_kernel<<<dim3((unsigned)((G.N + 255) / 256), d, mb), 256, 0, st>>>(g->chT.d(), g->chC.d(), G.N, d, g->chY.d());
      for (int f = 0; f < G.p; ++f) {
        const double* src = g->chY.d(); double* zb[2] = {g->chZ.d(), g->chZ2.d()}; int bi = 0;
        for (int dd = 0; dd < G.p; ++dd) {
          if (dd == f) continue;
          const long in2 = G.stride[dd], out2 = G.N / (in2 * G.ng[dd]) * nd;
          mode_product_kernel<<<(unsigned)(((long)nd * G.N + 255) / 256), 256, 0, st>>>(g->Kf.d() + g->koff[dd], G.ng[dd], src, zb[bi], in2, out2);
          src = zb[bi]; bi ^= 1; c->stats.kernel_launches++;
        }
        const long inner = G.stride[f], outer = G.N / (inner * G.ng[f]) * nd;
        mode_gram_acc_kernel<<<dim3(G.ng[f], G.ng[f]), 256, 0, st>>>(src, g->chT.d(), G.ng[f], inner, outer, g->chP.d() + g->koff[f]);
      }
      c->stats.kernel_launches += 4 + G.p;
    }
  }
  if (c->nranks > 1) {
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(g->chP.p, g->chP.p, (size_t)ksz, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
    KGP_NCCL(c, api, api.AllReduce(g->chacc.p, g->chacc.p, 2, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
  }
  // out[2] = ldB2 = n log(a)/2 + probe sum ; out[3] = sum(dW)
  double h2[2];
  KGP_CUDA(c, cudaMemcpyAsync(h2, g->chacc.p, 16, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  h2[0] += (double)n * log(a) / 2.0;
  KGP_CUDA(c, cudaMemcpyAsync(g->out.d() + 2, h2, 16, cudaMemcpyHostToDevice, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  return KGP_OK;
}

// ---- stochastic Lanczos quadrature log-det (extension, no GPML counterpart) -----------------------
//   ldB2 = 1/(2m) sum_i |v_i|^2 e1' log(T_i) e1 with T_i the Lanczos tridiagonal of B = I + sW K sW started at v_i;
//   derivatives by Hutchinson: d ldB2 = 1/(2m) sum_i v_i' B^-1 dB v_i, g_i = B^-1 v_i from the (batched or
//   preconditioned) CG; with uniform sW both bilinear forms collapse to grid space like the Chebyshev ones.
static void sym_eig_host(std::vector<double>& A, int k, std::vector<double>& V) {      // cyclic Jacobi, A -> diag, V cols = vectors
  V.assign((size_t)k * k, 0.0); for (int i = 0; i < k; ++i) V[(size_t)i * k + i] = 1.0;
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0, dg = 0.0;
    for (int i = 0; i < k; ++i) for (int j = 0; j < k; ++j) (i == j ? dg : off) += A[(size_t)i * k + j] * A[(size_t)i * k + j];
    if (off <= 1e-30 * dg) break;
    for (int p = 0; p < k - 1; ++p)
      for (int q = p + 1; q < k; ++q) {
        const double apq = A[(size_t)p * k + q];
        if (fabs(apq) < 1e-300) continue;
        const double th = (A[(size_t)q * k + q] - A[(size_t)p * k + p]) / (2.0 * apq);
        const double t = (th >= 0 ? 1.0 : -1.0) / (fabs(th) + sqrt(th * th + 1.0)), cs = 1.0 / sqrt(t * t + 1.0), sn = t * cs;
        for (int r = 0; r < k; ++r) { const double x = A[(size_t)r * k + p], y = A[(size_t)r * k + q]; A[(size_t)r * k + p] = cs * x - sn * y; A[(size_t)r * k + q] = sn * x + cs * y; }
        for (int r = 0; r < k; ++r) { const double x = A[(size_t)p * k + r], y = A[(size_t)q * k + r]; A[(size_t)p * k + r] = cs * x - sn * y; A[(size_t)q * k + r] = sn * x + cs * y; }
        for (int r = 0; r < k; ++r) { const double x = V[(size_t)r * k + p], y = V[(size_t)r * k + q]; V[(size_t)r * k + p] = cs * x - sn * y; V[(size_t)r * k + q] = sn * x + cs * y; }
      }
  }
}
static int slq_logdet(kgp_ctx* c, GridState* g, const GridDesc& G, int slot, double sw, bool want_deriv) {
  Slot& S = c->slots[slot]; Model& mo = c->model;
  const long n = S.n; const int D = S.D; cudaStream_t st = c->stream;
  const int steps = mo.opt.ldB2_slq_steps > 0 ? mo.opt.ldB2_slq_steps : 30;
  if (steps > 512) return c->fail(KGP_BAD_ARG, "ldB2_slq_steps %d too large (max 512)", steps);
  if ((double)steps * (double)n * 8.0 > 64e9) return c->fail(KGP_BAD_ARG, "ldB2_slq_steps x n: Lanczos basis above 64 GB");
  int m = mo.opt.ldB2_cheby_hutch > 0 ? mo.opt.ldB2_cheby_hutch : 10;
  if (g->probes_injected) {
    if (g->probes_n != n) return c->fail(KGP_BAD_ARG, "injected probes have %ld rows, data has %ld", g->probes_n, n);
    m = g->probes_m;
  } else {
    KGP_CUDA(c, g->probes.ensure((size_t)m * n * 8));
    rademacher_kernel<<<(unsigned)(((long)m * n + 255) / 256), 256, 0, st>>>(mo.opt.ldB2_cheby_seed, n, m, g->probes.d());
    g->probes_n = n; g->probes_m = m;
  }
  long ksz = 0; for (int q = 0; q < G.p; ++q) ksz += (long)G.ng[q] * G.ng[q];
  KGP_CUDA(c, g->chT.ensure((size_t)2 * G.N * 8)); KGP_CUDA(c, g->chU.ensure((size_t)2 * G.N * 8));
  KGP_CUDA(c, g->chP.ensure((size_t)ksz * 8)); KGP_CUDA(c, g->chacc.ensure(64));
  KGP_CUDA(c, g->bvc.ensure(n * 8)); KGP_CUDA(c, g->lzQ.ensure((size_t)steps * n * 8));
  KGP_CUDA(c, g->lzP.ensure((size_t)steps * nblocks_for(n) * 8)); KGP_CUDA(c, g->lzH.ensure((size_t)steps * 8));
  KGP_CUDA(c, cudaMemsetAsync(g->chP.p, 0, ksz * 8, st));
  KGP_CUDA(c, cudaMemsetAsync(g->chacc.p, 0, 64, st));
  const int nb = nblocks_for(n);
  int lo = 0, hi = m;
  if (c->nranks > 1) { const int base = m / c->nranks, extra = m % c->nranks; lo = c->rank * base + std::min(c->rank, extra); hi = lo + base + (c->rank < extra ? 1 : 0); }
  double quad = 0.0;
  for (int i = lo; i < hi; ++i) {
    const double* v = g->probes.d() + (long)i * n;
    double h2[2];
    dot_kernel<<<1, 1024, 0, st>>>(v, v, n, g->chacc.d() + 2);
    KGP_CUDA(c, cudaMemcpyAsync(h2, g->chacc.d() + 2, 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    const double nv2 = h2[0];
    // Lanczos with full re-orthogonalisation: the basis is kept (steps x n), so two implementations that round
    // differently stay on the same Krylov sequence and the quadrature is reproducible
    double* Qb = g->lzQ.d(); double* w = g->bvc.d();
    vec_scale_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, 1.0 / sqrt(nv2), v, Qb);
    std::vector<double> al, be; double beta = 0.0;
    for (int j = 0; j < steps; ++j) {
      const double* q = Qb + (long)j * n; const double* qp = j > 0 ? Qb + (long)(j - 1) * n : q;     // (beta = 0 at j = 0)
      scatter(c, G, S.X.d(), n, D, q, sw, g->chT.d());                          // w = B q ; partials of q.w
      kron_to(c, g, G, g->chT.d(), g->chU.d(), -1);
      for_p(G.p, [&](auto P) {
        cheby_step_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, S.X.d(), n, D, g->chU.d(), q, q, 1.0, sw, w, g->partA.d());
      });
      sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), nb, g->chacc.d() + 4);
      lanczos_update_kernel<<<nb, 256, 0, st>>>(n, g->chacc.d() + 4, beta, q, qp, w, g->partB.d());
      multi_dot_kernel<<<nb, 256, 0, st>>>(n, j + 1, Qb, w, g->lzP.d());
      sum_rows_kernel<<<j + 1, 256, 0, st>>>(g->lzP.d(), nb, g->lzH.d());
      multi_axpy_kernel<<<nb, 256, 0, st>>>(n, j + 1, Qb, g->lzH.d(), w, g->partB.d());
      sum_small_kernel<<<1, 256, 0, st>>>(g->partB.d(), nb, g->chacc.d() + 5);
      KGP_CUDA(c, cudaMemcpyAsync(h2, g->chacc.d() + 4, 16, cudaMemcpyDeviceToHost, st));
      KGP_CUDA(c, cudaStreamSynchronize(st));
      c->stats.kernel_launches += 7;
      const double a = h2[0], b = sqrt(h2[1]);
      al.push_back(a);
      if (j + 1 == steps || !(b > 1e-12 * std::max(fabs(a), 1.0))) break;
      be.push_back(b);
      vec_scale_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, 1.0 / b, w, Qb + (long)(j + 1) * n);
      beta = b;
    }
    const int k = (int)al.size();
    std::vector<double> T((size_t)k * k, 0.0), Vv;
    for (int r = 0; r < k; ++r) { T[(size_t)r * k + r] = al[r]; if (r + 1 < k) { T[(size_t)r * k + r + 1] = be[r]; T[(size_t)(r + 1) * k + r] = be[r]; } }
    sym_eig_host(T, k, Vv);
    double qv = 0.0;
    for (int r = 0; r < k; ++r) qv += Vv[r] * Vv[r] * log(fmax(T[(size_t)r * k + r], 1e-300));     // first components: row 0 of V
    quad += nv2 * qv / (2.0 * m);
  }
  if (want_deriv && hi > lo) {
    const bool pre = mo.opt.cg_precond != 0;
    const int MB = pre ? 1 : CG_MAXCOLS;
    const int maxit = mo.opt.cg_maxit > 0 ? mo.opt.cg_maxit : (int)std::min<long>(n, 20);
    KGP_CUDA(c, g->pvS.ensure((size_t)MB * n * 8));
    for (int c0 = lo; c0 < hi; c0 += MB) {
      const int mb = std::min(MB, hi - c0);
      const double* V0 = g->probes.d() + (long)c0 * n;
      double gscale;                                                         // pvS = gscale * B^-1 v
      if (pre) {
        int rc = pcg_solve(c, g, G, slot, sw, mo.opt.cg_tol, maxit, V0, g->pvS.d(), false); if (rc) return rc;
        gscale = sw * sw;
      } else {
        int rc = cg_solve_sorted(c, g, G, slot, V0, n, mb, sw, mo.opt.cg_tol, maxit); if (rc) return rc;
        for (int col = 0; col < mb; ++col)
          cg_unsort_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->sx.d() + (long)col * g->npad, (const int*)g->perm.p, n, 1.0, g->pvS.d() + (long)col * n);
        gscale = sw;
      }
      for (int col = 0; col < mb; ++col) {
        double* Tg = g->chT.d(); double* Tv = g->chT.d() + G.N; double* Uv = g->chU.d();
        scatter(c, G, S.X.d(), n, D, g->pvS.d() + (long)col * n, sw / gscale, Tg);       // Mx'(sW g)
        scatter(c, G, S.X.d(), n, D, V0 + (long)col * n, sw, Tv);                          // Mx'(sW v)
        kron_to(c, g, G, Tv, Uv, -1);
        dot_kernel<<<1, 1024, 0, st>>>(Tg, Uv, G.N, g->chacc.d() + 2);
        sum_scaled_acc_kernel<<<1, 256, 0, st>>>(g->chacc.d() + 2, 1, 1.0 / (2.0 * m * sw * sw), g->chacc.d() + 1);   // sum(dW)
        for (int f = 0; f < G.p; ++f) {
          const long inner = G.stride[f], outer = G.N / (inner * G.ng[f]);
          kron_to(c, g, G, Tg, g->chU.d() + G.N, f);
          mode_gram_acc_kernel<<<dim3(G.ng[f], G.ng[f]), 128, 0, st>>>(g->chU.d() + G.N, Tv, G.ng[f], inner, outer, g->chP.d() + g->koff[f]);
        }
        c->stats.kernel_launches += 5 + 2 * G.p;
      }
    }
    vec_scale_copy_kernel<<<(unsigned)((ksz + 255) / 256), 256, 0, st>>>(ksz, 1.0 / (2.0 * m), g->chP.d(), g->chP.d());
  }
  // chacc[0] = ldB2 share of this rank, chacc[1] = sum(dW) share
  KGP_CUDA(c, cudaMemcpyAsync(g->chacc.p, &quad, 8, cudaMemcpyHostToDevice, st));
  if (c->nranks > 1) {
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(g->chP.p, g->chP.p, (size_t)ksz, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
    KGP_NCCL(c, api, api.AllReduce(g->chacc.p, g->chacc.p, 2, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
  }
  KGP_CUDA(c, cudaMemcpyAsync(g->out.d() + 2, g->chacc.p, 16, cudaMemcpyDeviceToDevice, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  return KGP_OK;
}

int grid_set_probes(kgp_ctx* c, const double* V, long n, int m, double emax) {
  GridState* g = gs(c);
  if (!V || m <= 0) { g->probes_injected = false; return KGP_OK; }
  KGP_CUDA(c, g->probes.ensure((size_t)m * n * 8));
  KGP_CUDA(c, cudaMemcpyAsync(g->probes.p, V, (size_t)m * n * 8, cudaMemcpyHostToDevice, c->stream));
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));
  g->probes_n = n; g->probes_m = m; g->probes_emax = emax; g->probes_injected = true;
  c->model.version++;                                  // cached posteriors no longer match
  return KGP_OK;
}

static bool gstate_valid(kgp_ctx* c, GridState* g, int slot, int level) {
  return g->st_level >= level && g->st_slot == slot && g->st_slot_ver == c->slots[slot].version &&
         g->st_model_ver == c->model.version && g->st_grid_ver == c->grid.version;
}

int grid_eval(kgp_ctx* c, int slot, int want, double* nlZ_out, double* dnlZ_out) {
  GridState* g = gs(c); Slot& S = c->slots[slot]; Model& m = c->model;
  const long n = S.n; const int D = S.D;
  if (m.cov_kind != KGP_COV_SEISO && m.cov_kind != KGP_COV_SM) return c->fail(KGP_BAD_ARG, "infGrid supports covSEiso / covSM factors");
  const bool slq = m.opt.ldB2_slq != 0;
  const bool cheby = m.opt.ldB2_cheby != 0 || slq;           // stochastic estimators share the derivative plumbing
  GridDesc G; int rc = grid_setup(c, g, G); if (rc) return rc;
  if (G.p != D) return c->fail(KGP_BAD_ARG, "Grid and data dimension are different.");
  const int nh = factor_nhyp(m), nhyp = nh * G.p;
  if ((int)m.hyp_cov.size() != nhyp) return c->fail(KGP_BAD_ARG, "Number of cov function hyperparameters disagree with cov function");
  const int nmean = ((m.mean_mask & 1) ? D : 0) + ((m.mean_mask & 2) ? 1 : 0);
  if ((int)m.hyp_mean.size() != nmean) return c->fail(KGP_BAD_ARG, "Number of mean function hyperparameters disagree with mean function");
  rc = ensure_grid_buffers(c, g, G, n, D, nhyp); if (rc) return rc;
  cudaStream_t st = c->stream;
  cudaStreamSynchronize(c->stream2);        // a previous call may have returned early with side-stream work pending
  memset(&c->stats, 0, sizeof c->stats); c->stats.n = n; c->stats.np = G.N;
  g->st_level = -1;
  const double sn2 = exp(2.0 * m.hyp_lik), sw = 1.0 / sqrt(sn2), w = 1.0 / sn2;
  const int kind = m.cov_kind == KGP_COV_SM ? 2 : 0, Q = m.sm_Q;

  cudaEventRecord(c->ev[0], st);
  KGP_CUDA(c, cudaMemcpyAsync(g->hyp.p, m.hyp_cov.data(), nhyp * 8, cudaMemcpyHostToDevice, st));
  if (nmean) KGP_CUDA(c, cudaMemcpyAsync(g->hypm.p, m.hyp_mean.data(), nmean * 8, cudaMemcpyHostToDevice, st));
  KGP_CUDA(c, cudaMemcpyAsync(g->eoff.p, g->eoff_h, GRID_MAXP * 4, cudaMemcpyHostToDevice, st));
  double prior = 1.0;
  EigBatch eb; memset(&eb, 0, sizeof eb); size_t eig_smem = 0;
  bool toep[GRID_MAXP] = {false}, any_toep = false;
  for (int d = 0; d < G.p; ++d) { toep[d] = d < (int)c->grid.toep.size() && c->grid.toep[d] != 0; any_toep |= toep[d]; }
  if (g->any_neq && (any_toep || m.opt.cg_precond || m.opt.pred_var != 0 || slq))
    return c->fail(KGP_BAD_ARG, "non-equispaced grid axes (neqinterp) and opt.deg != 3 run with the plain CG and the Fiedler / Chebyshev log-det only");
  if (any_toep && (m.opt.cg_precond || m.opt.pred_var != 0))
    return c->fail(KGP_BAD_ARG, "opt.cg_precond / opt.pred_var need the factor eigenvectors: not available for Toeplitz factors (nested xg)");
  for (int d = 0; d < G.p; ++d) {
    const double am = axis_mean(c->grid.axes[d]);
    const long nn = (long)G.ng[d] * G.ng[d];
    factor_build_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(kind, Q, g->hyp.d() + d * nh, G.ax[d], G.ng[d], am, g->Kf.d() + g->koff[d]);
    KGP_CUDA(c, cudaMemcpyAsync(g->Af.d() + g->koff[d], g->Kf.d() + g->koff[d], nn * 8, cudaMemcpyDeviceToDevice, st));
    if (d == 0) transpose_small_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(g->Kf.d(), G.ng[0], g->KfT.d());
    const size_t need = 2 * (size_t)G.ng[d] * (G.ng[d] | 1) * 8;
    eb.A[d] = g->Af.d() + g->koff[d]; eb.V[d] = g->Vf.d() + g->koff[d]; eb.ev[d] = g->ee.d() + g->eoff_h[d];
    eb.ng[d] = toep[d] ? 0 : G.ng[d]; eb.use_smem[d] = need <= 190 * 1024; if (eb.use_smem[d] && !toep[d]) eig_smem = std::max(eig_smem, need);
    {
      // mirror-symmetric axis (every equispaced one) + stationary factor kernel => centrosymmetric factor: two half-order
      // eigenproblems (jacobi_eig_kernel).  The test is on the spacings, to rounding: an axis that merely passes equi()'s
      // 1e-9 does not qualify.
      static const bool split_ok = !(getenv("KGP_EIG_SPLIT") && atoi(getenv("KGP_EIG_SPLIT")) == 0);
      const std::vector<double>& ax = c->grid.axes[d];
      const int na = (int)ax.size(), nh = (na + 1) / 2;
      const size_t need3 = 3 * (size_t)nh * (nh | 1) * 8;
      bool sym = split_ok && !toep[d] && na == G.ng[d] && na >= 2 && need3 <= 190 * 1024;
      if (sym) {
        double amax = 0.0; for (double v : ax) amax = std::max(amax, fabs(v));
        for (int a = 0; a + 1 < na && sym; ++a)
          if (fabs((ax[a + 1] - ax[a]) - (ax[na - 1 - a] - ax[na - 2 - a])) > 8.0 * 2.220446049250313e-16 * amax) sym = false;
      }
      eb.split[d] = sym ? 1 : 0;
      if (sym) eig_smem = std::max(eig_smem, need3);
    }
    c->stats.kernel_launches += 1;
    if (kind == 0) prior *= exp(2.0 * m.hyp_cov[d * nh + 1]);
    else { double sws = 0; for (int q = 0; q < Q; ++q) sws += exp(m.hyp_cov[d * nh + q]); prior *= sws; }
  }
  g->prior_var = prior;
  // eigen-decompositions of all factors: one CTA each, on the side stream, overlapped with the CG solve
  cudaEventRecord(c->ev[10], st);
  cudaStreamWaitEvent(c->stream2, c->ev[10], 0);
  g->eig_ok = !cheby || m.opt.cg_precond || m.opt.pred_var != 0;
  if (g->eig_ok) {
    // Jacobi converges quadratically from a nearly diagonal start: within an optimiser run the hyper-parameters move
    // little between evaluations, so A = V_prev' K V_prev needs 2-4 sweeps instead of ~14.  Cold start on a new grid
    // and every 32nd evaluation (keeps the accumulated V orthogonal to rounding).
    static const bool warm_ok = !(getenv("KGP_EIG_WARM") && atoi(getenv("KGP_EIG_WARM")) == 0);
    const bool warm = warm_ok && g->eig_have && g->eig_grid_ver == c->grid.version && (g->eig_evals % 32) != 0;
    if (warm) {
      long ksz = 0; for (int d = 0; d < G.p; ++d) ksz += (long)G.ng[d] * G.ng[d];
      KGP_CUDA(c, g->eigT.ensure(ksz * 8));
      for (int d = 0; d < G.p; ++d) {
        if (toep[d]) continue;
        if (eb.split[d]) { eb.warm[d] = 1; continue; }           // the split kernel transforms its half problems itself
        const int ngd = G.ng[d]; const unsigned nbk = (unsigned)(((long)ngd * ngd + 255) / 256);
        matmul_small_kernel<<<nbk, 256, 0, c->stream2>>>(g->Kf.d() + g->koff[d], g->Vf.d() + g->koff[d], g->eigT.d() + g->koff[d], ngd, 0);
        matmul_small_kernel<<<nbk, 256, 0, c->stream2>>>(g->Vf.d() + g->koff[d], g->eigT.d() + g->koff[d], g->Af.d() + g->koff[d], ngd, 1);
        eb.warm[d] = 1;
      }
    }
    jacobi_eig_kernel<<<2 * G.p, EIG_THREADS, eig_smem, c->stream2>>>(eb, getenv("KGP_EIG_SWEEPS") ? atoi(getenv("KGP_EIG_SWEEPS")) : 40);
    for (int d = 0; d < G.p; ++d)                               // Toeplitz factors: Whittle-circulant eigenvalues (apxGrid.m:576-580)
      if (toep[d]) {
        whittle_eig_kernel<<<1, 1024, (size_t)G.ng[d] * 8, c->stream2>>>(kind, Q, g->hyp.d() + d * nh, G.ng[d], G.range[d] / (double)(G.ng[d] - 1),
                                                                        g->ee.d() + g->eoff_h[d]);
        c->stats.kernel_launches++;
      }
    g->eig_have = true; g->eig_grid_ver = c->grid.version; g->eig_evals++;
  }
  cudaEventRecord(c->ev[11], c->stream2);
  c->stats.kernel_launches++;
  resid_mean_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(S.X.d(), S.has_y ? S.y.d() : nullptr, n, D, m.mean_mask, g->hypm.d(), g->ym.d(), n);
  c->stats.kernel_launches++;
  cudaEventRecord(c->ev[1], st);

  int maxit = m.opt.cg_maxit > 0 ? m.opt.cg_maxit : (int)std::min<long>(n, 20);
  static const bool legacy_cg = getenv("KGP_CG_LEGACY") && atoi(getenv("KGP_CG_LEGACY"));
  if (m.opt.cg_precond) {
    rc = pcg_solve(c, g, G, slot, sw, m.opt.cg_tol, maxit, g->ym.d(), g->x.d(), true);
    if (rc) return rc;
  } else if (G.p <= 2 && !legacy_cg && !g->any_neq && c->px_ok && dist_rows_sharded(c->nranks, n, G.N)) {
    // row-sharded solve (SURVEY 8(e) KISS-GP (ii)): this rank's contiguous share of the points, the two reductions
    // of every iteration cross the ranks over peer memory inside the persistent kernel
    const long lo = dist_row_begin(c->rank, c->nranks, n), hi = dist_row_begin(c->rank + 1, c->nranks, n);
    swap_sorted(g);
    if (g->sub_lo != lo || g->sub_n != hi - lo) { g->sub_lo = lo; g->sub_n = hi - lo; g->sort_slot = -1; }
    rc = cg_solve_sorted(c, g, G, slot, g->ym.d(), n, 1, sw, m.opt.cg_tol, maxit);
    if (!rc) {
      KGP_CUDA(c, cudaMemsetAsync(g->x.p, 0, n * 8, st));
      cg_unsort_kernel<<<(unsigned)((hi - lo + 255) / 256), 256, 0, st>>>(g->sx.d(), (const int*)g->perm.p, hi - lo, sw, g->x.d() + lo);
    }
    swap_sorted(g);
    if (rc) return rc;
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(g->x.p, g->x.p, (size_t)n, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
  } else if (G.p <= 2 && !legacy_cg && !g->any_neq) {          // (non-equispaced axes: general streaming kernels below)
    rc = cg_solve_sorted(c, g, G, slot, g->ym.d(), n, 1, sw, m.opt.cg_tol, maxit);
    if (rc) return rc;
    cg_unsort_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->sx.d(), (const int*)g->perm.p, n, sw, g->x.d());   // alpha = sW q
  } else {
    rc = cg_solve(c, g, G, S.X.d(), n, D, sw, m.opt.cg_tol, maxit);
    if (rc) return rc;
    axpy_scale_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, sw, g->x.d(), g->x.d());      // alpha = sW q
  }
  dot_kernel<<<1, 1024, 0, st>>>(g->ym.d(), g->x.d(), n, g->out.d() + 0);
  dot_kernel<<<1, 1024, 0, st>>>(g->x.d(), g->x.d(), n, g->out.d() + 1);
  cudaEventRecord(c->ev[2], st);

  // Fiedler bound (apx.m:236-240,312-318) or the Chebyshev-Hutchinson estimator (apx.m:336-391)
  cudaStreamWaitEvent(st, c->ev[11], 0);
  const unsigned char* keep = nullptr;
  if (cheby) {
    const int it0 = c->stats.cg_iters; const double rel0 = c->stats.cg_relres;      // keep the alpha solve's CG statistics
    rc = slq ? slq_logdet(c, g, G, slot, sw, want >= 1) : cheby_logdet(c, g, G, slot, sw, want >= 1, kind, Q, nh);
    if (rc) return rc;
    c->stats.cg_iters = it0; c->stats.cg_relres = rel0;
  } else {
  if (n < G.N) {
    long Npad = 1; while (Npad < G.N) Npad <<= 1;
    KGP_CUDA(c, g->sortk.ensure(Npad * 8)); KGP_CUDA(c, g->sorti.ensure(Npad * 4)); KGP_CUDA(c, g->keep.ensure(G.N));
    kron_eig_kernel<<<(unsigned)((Npad + 255) / 256), 256, 0, st>>>(G, g->ee.d(), (const int*)g->eoff.p, g->sortk.d(), (int*)g->sorti.p, Npad);
    for (long k = 2; k <= Npad; k <<= 1)
      for (long j = k >> 1; j > 0; j >>= 1) {
        bitonic_stage_kernel<<<(unsigned)((Npad + 255) / 256), 256, 0, st>>>(g->sortk.d(), (int*)g->sorti.p, Npad, j, k);
        c->stats.kernel_launches++;
      }
    KGP_CUDA(c, cudaMemsetAsync(g->keep.p, 0, G.N, st));
    mark_keep_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>((const int*)g->sorti.p, n, G.N, (unsigned char*)g->keep.p);
    keep = (const unsigned char*)g->keep.p;
  }
  const int fb = nblocks_for(G.N);
  fiedler_kernel<<<fb, 256, 0, st>>>(G, g->ee.d(), (const int*)g->eoff.p, (double)n / (double)G.N, w, keep, g->de.d(), g->partA.d(), g->partB.d());
  sum_small_kernel<<<1, 256, 0, st>>>(g->partA.d(), fb, g->out.d() + 2);
  sum_small_kernel<<<1, 256, 0, st>>>(g->partB.d(), fb, g->out.d() + 3);
  c->stats.kernel_launches += 6;
  }

  // beta = Kg Mx' alpha (needed by predict and deriv_x); T = Mx' alpha kept for dirder
  scatter(c, G, S.X.d(), n, D, g->x.d(), 1.0, g->tgrid.d());
  KGP_CUDA(c, cudaMemcpyAsync(g->g0.p, g->tgrid.p, G.N * 8, cudaMemcpyDeviceToDevice, st));
  {
    double* b = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1);
    KGP_CUDA(c, cudaMemcpyAsync(g->beta.p, b, G.N * 8, cudaMemcpyDeviceToDevice, st));
  }
  if (want >= 1) {
    for (int d = 0; d < G.p; ++d) {
      const double am = axis_mean(c->grid.axes[d]);
      const long nn = (long)G.ng[d] * G.ng[d];
      if (cheby) {      // log-det part: dK_d of the accumulated Krylov Gram matrix (cheby_logdet)
        factor_dirder_kernel<<<nh, 256, 0, st>>>(kind, Q, g->hyp.d() + d * nh, G.ax[d], G.ng[d], am, g->chP.d() + g->koff[d], 1.0, g->dhyp.d() + d * nh, 0);
      } else {
      // Fiedler part: dK_d( V diag(z) V' ), or for a Toeplitz factor the sum over the Whittle lags (apx.m:256-266)
      fiedler_z_kernel<<<G.ng[d], 256, 0, st>>>(G, d, g->ee.d(), (const int*)g->eoff.p, g->de.d(), g->zv.d());
      if (toep[d]) {
        toep_fiedler_dirder_kernel<<<nh, 256, (size_t)G.ng[d] * 8, st>>>(kind, Q, g->hyp.d() + d * nh, G.ng[d], G.range[d] / (double)(G.ng[d] - 1),
                                                                     g->zv.d(), g->dhyp.d() + d * nh);
      } else {
      vdv_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(g->Vf.d() + g->koff[d], g->zv.d(), G.ng[d], g->Qm.d());
      factor_dirder_kernel<<<nh, 256, 0, st>>>(kind, Q, g->hyp.d() + d * nh, G.ax[d], G.ng[d], am, g->Qm.d(), 1.0, g->dhyp.d() + d * nh, 0);
      }
      }
      // data-fit part: -1/2 dK_d( P ),  P = shp(v) shp(T)',  v = T with every other factor applied
      KGP_CUDA(c, cudaMemcpyAsync(g->g0.p, g->tgrid.p, G.N * 8, cudaMemcpyDeviceToDevice, st));
      double* v = kron_apply(c, g, G, g->g0.d(), g->g1.d(), d);
      const long inner = G.stride[d], outer = G.N / (inner * G.ng[d]);
      dim3 gg(G.ng[d], G.ng[d]);
      mode_gram_kernel<<<gg, 128, 0, st>>>(v, g->tgrid.d(), G.ng[d], inner, outer, g->Pm.d());
      factor_dirder_kernel<<<nh, 256, 0, st>>>(kind, Q, g->hyp.d() + d * nh, G.ax[d], G.ng[d], am, g->Pm.d(), -0.5, g->dhyp.d() + d * nh, 1);
      c->stats.kernel_launches += 5;
    }
    if (nmean) { mean_grad_kernel<<<D + 1, 1024, 0, st>>>(S.X.d(), g->x.d(), n, D, g->dhyp.d() + nhyp); c->stats.kernel_launches++; }
  }
  if (want >= 1) {      // dnlZ/dX is one gather pass: always kept, so a following get_dlik_dx reuses this inference
    const double* dmean = nullptr;
    if (m.mean_mask & 1) { KGP_CUDA(c, cudaMemcpyAsync(g->dmean.p, m.hyp_mean.data(), D * 8, cudaMemcpyHostToDevice, st)); dmean = g->dmean.d(); }
    for_p(G.p, [&](auto P) {
      interp_dx_kernel<decltype(P)::value><<<(unsigned)((n + 255) / 256), 256, 0, st>>>(G, S.X.d(), n, D, g->beta.d(), g->x.d(), dmean, g->dX.d());
    });
    c->stats.kernel_launches++;
  }
  cudaEventRecord(c->ev[3], st);
  double ho[8]; std::vector<double> hd(nhyp + D + 2);
  KGP_CUDA(c, cudaMemcpyAsync(ho, g->out.p, 4 * 8, cudaMemcpyDeviceToHost, st));
  if (want >= 1) KGP_CUDA(c, cudaMemcpyAsync(hd.data(), g->dhyp.p, (nhyp + D + 1) * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  float ms = 0;
  cudaEventElapsedTime(&ms, c->ev[0], c->ev[3]); c->stats.ms_total = ms;
  cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]); c->stats.ms_kbuild = ms;
  cudaEventElapsedTime(&ms, c->ev[1], c->ev[2]); c->stats.ms_solve = ms;
  cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]); c->stats.ms_reduce = ms;
  const double yta = ho[0], ata = ho[1], ldB2 = ho[2], sdW = ho[3];
  const double nlZ = yta / 2.0 + ldB2 + (double)n * log(2.0 * M_PI * sn2) / 2.0;
  if (nlZ_out) *nlZ_out = nlZ;
  if (want >= 1 && dnlZ_out) {
    int o = 0;
    for (int i = 0; i < nhyp; ++i) dnlZ_out[o++] = hd[i];
    dnlZ_out[o++] = -sn2 * ata - 2.0 * sdW / sn2 + (double)n;          // infGaussLik.m:32
    if (m.mean_mask & 1) for (int d = 0; d < D; ++d) dnlZ_out[o++] = hd[nhyp + d];
    if (m.mean_mask & 2) dnlZ_out[o++] = hd[nhyp + D];
  }
  g->st_slot = slot; g->st_slot_ver = S.version; g->st_model_ver = m.version; g->st_grid_ver = c->grid.version;
  g->st_level = want >= 1 ? 2 : 0;
  return KGP_OK;
}

int grid_dlik(kgp_ctx* c, int slot, double* out, bool out_is_dev) {
  GridState* g = gs(c);
  if (!gstate_valid(c, g, slot, 2)) {
    int rc = grid_eval(c, slot, 2, nullptr, nullptr);     // like dlikGrid.m:29-36: the inference is re-run
    if (rc) return rc;
  }
  Slot& S = c->slots[slot];
  KGP_CUDA(c, cudaMemcpyAsync(out, g->dX.p, (size_t)S.n * S.D * 8, out_is_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGP_OK;
}

int grid_alpha(kgp_ctx* c, double* out) {
  GridState* g = gs(c);
  if (g->st_level < 0) return c->fail(KGP_NO_DATA, "no posterior computed yet");
  KGP_CUDA(c, cudaMemcpy(out, g->x.p, (size_t)c->slots[g->st_slot].n * 8, cudaMemcpyDeviceToHost));
  return KGP_OK;
}

// post.L applied to m right-hand sides at once: out = sW .* CG(B, sW .* r)  (apx.m:219-222 with the batched conjgrad
// of apx.m:281-302: per-column convergence masks).  R and out are m x n, one right-hand side per row (host).
int grid_solve(kgp_ctx* c, int slot, const double* R, int m, double* out) {
  GridState* g = gs(c);
  if (!gstate_valid(c, g, slot, 0)) {
    int rc = grid_eval(c, slot, 0, nullptr, nullptr);
    if (rc) return rc;
  }
  GridDesc G; int rc = grid_setup(c, g, G); if (rc) return rc;
  if (G.p > 2) return c->fail(KGP_BAD_ARG, "batched solve is built for 1-D and 2-D grids");
  if (g->any_neq) return c->fail(KGP_BAD_ARG, "batched solve needs equispaced grid axes (cell-sorted CG)");
  Slot& S = c->slots[slot]; Model& mo = c->model; const long n = S.n;
  const double sn2 = exp(2.0 * mo.hyp_lik), sw = 1.0 / sqrt(sn2);
  const int maxit = mo.opt.cg_maxit > 0 ? mo.opt.cg_maxit : (int)std::min<long>(n, 20);
  cudaStream_t st = c->stream;
  DevBuf rhs, res;
  KGP_CUDA(c, rhs.ensure((size_t)CG_MAXCOLS * n * 8)); KGP_CUDA(c, res.ensure((size_t)CG_MAXCOLS * n * 8));
  for (int c0 = 0; c0 < m; c0 += CG_MAXCOLS) {
    const int mb = std::min(CG_MAXCOLS, m - c0);
    KGP_CUDA(c, cudaMemcpyAsync(rhs.p, R + (long)c0 * n, (size_t)mb * n * 8, cudaMemcpyHostToDevice, st));
    rc = cg_solve_sorted(c, g, G, slot, rhs.d(), n, mb, sw, mo.opt.cg_tol, maxit);
    if (rc) { rhs.release(); res.release(); return rc; }
    for (int col = 0; col < mb; ++col)
      cg_unsort_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->sx.d() + (long)col * g->npad, (const int*)g->perm.p, n, sw, res.d() + (long)col * n);
    KGP_CUDA(c, cudaMemcpyAsync(out + (long)c0 * n, res.p, (size_t)mb * n * 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
  }
  rhs.release(); res.release();
  g->st_level = -1;                 // the CG work vectors were reused: the cached posterior is gone
  return KGP_OK;
}

// ---- opt.pred_var: explained variance on the grid by perturb-and-MAP (infGrid.m:91-101, sample() :147-152) -------
// standard normals from a counter-based generator (splitmix64 pairs + Box-Muller) when no draws are injected
__global__ void normal_kernel(unsigned long long seed, long count, double* __restrict__ out) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= count) return;
  auto mix = [](unsigned long long z) { z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; return z ^ (z >> 31); };
  const unsigned long long a = mix(seed + 0x9E3779B97F4A7C15ULL * (unsigned long long)(2 * q + 1));
  const unsigned long long b = mix(seed + 0x9E3779B97F4A7C15ULL * (unsigned long long)(2 * q + 2));
  const double u1 = ((double)(a >> 11) + 1.0) * (1.0 / 9007199254740992.0), u2 = (double)(b >> 11) * (1.0 / 9007199254740992.0);
  out[q] = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
}
__global__ void sq_acc_kernel(long N, double coef, const double* __restrict__ z, double* __restrict__ vg) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < N) vg[q] = fma(coef * z[q], z[q], vg[q]);
}
__global__ void absmax_flag_kernel(long N, const double* __restrict__ v, double* __restrict__ out) {
  __shared__ double sh[32];
  double a = 0.0;
  for (long q = threadIdx.x; q < N; q += blockDim.x) a += fabs(v[q]);
  const double t = block_sum(a, sh);
  if (threadIdx.x == 0) out[0] = t;
}
// Rg: ns x N, Rn: ns x n (host, one draw per row; both NULL -> generated from seed).  vg_out (host, N) optional.
int grid_pred_var(kgp_ctx* c, int slot, const double* Rg, const double* Rn, int ns, unsigned long long seed, double* vg_out) {
  GridState* g = gs(c); Model& mo = c->model;
  if (ns <= 0) return c->fail(KGP_BAD_ARG, "pred_var needs ns > 0 samples");
  if ((Rg == nullptr) != (Rn == nullptr)) return c->fail(KGP_BAD_ARG, "pred_var: inject both draw matrices or neither");
  if (!gstate_valid(c, g, slot, 0) || !g->eig_ok) {
    if ((mo.opt.ldB2_cheby || mo.opt.ldB2_slq) && !mo.opt.cg_precond && !mo.opt.pred_var)
      return c->fail(KGP_BAD_ARG, "pred_var with opt.ldB2_cheby needs opt.pred_var set (factor eigen-decompositions)");
    int rc = grid_eval(c, slot, 1, nullptr, nullptr); if (rc) return rc;
  }
  GridDesc G; int rc = grid_setup(c, g, G); if (rc) return rc;
  if (G.p > 2) return c->fail(KGP_BAD_ARG, "pred_var is built for 1-D and 2-D grids");
  if (g->any_neq) return c->fail(KGP_BAD_ARG, "pred_var needs equispaced grid axes (cell-sorted CG)");
  Slot& S = c->slots[slot]; const long n = S.n, N = G.N; const int D = S.D;
  const double sn2 = exp(2.0 * mo.hyp_lik), sw = 1.0 / sqrt(sn2), sn = sqrt(sn2);
  const int maxit = mo.opt.cg_maxit > 0 ? mo.opt.cg_maxit : (int)std::min<long>(n, 20);
  cudaStream_t st = c->stream;
  const bool pre = mo.opt.cg_precond != 0;
  const int MB = pre ? 1 : CG_MAXCOLS;
  long ksz = 0; for (int d = 0; d < G.p; ++d) ksz += (long)G.ng[d] * G.ng[d];
  KGP_CUDA(c, g->VT.ensure(ksz * 8)); KGP_CUDA(c, g->se.ensure(N * 8));
  KGP_CUDA(c, g->vg.ensure(N * 8)); KGP_CUDA(c, g->pvG.ensure((size_t)2 * N * 8));
  KGP_CUDA(c, g->pvA.ensure((size_t)MB * n * 8)); KGP_CUDA(c, g->pvS.ensure((size_t)MB * n * 8));
  cudaStreamWaitEvent(st, c->ev[11], 0);
  if (!pre) {                                                        // (the preconditioner set-up already made these)
    for (int d = 0; d < G.p; ++d) {
      const long nn = (long)G.ng[d] * G.ng[d];
      transpose_small_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, st>>>(g->Vf.d() + g->koff[d], G.ng[d], g->VT.d() + g->koff[d]);
    }
    sqrt_kron_eig_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(G, g->ee.d(), (const int*)g->eoff.p, g->se.d());
  }
  KGP_CUDA(c, cudaMemsetAsync(g->vg.p, 0, N * 8, st));
  const int nb = nblocks_for(n);
  double* zg = g->pvG.d(); double* tg = g->pvG.d() + N;
  int worst_it = 0; double worst_rel = 0.0;
  for (int c0 = 0; c0 < ns; c0 += MB) {
    const int mb = std::min(MB, ns - c0);
    for (int col = 0; col < mb; ++col) {
      double* Acol = g->pvA.d() + (long)col * n;
      if (Rg) {
        KGP_CUDA(c, cudaMemcpyAsync(g->g0.p, Rg + (long)(c0 + col) * N, N * 8, cudaMemcpyHostToDevice, st));
        KGP_CUDA(c, cudaMemcpyAsync(Acol, Rn + (long)(c0 + col) * n, n * 8, cudaMemcpyHostToDevice, st));
      } else {
        normal_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(seed + 2ULL * (unsigned long long)(c0 + col) * 0x51ED27ULL, N, g->g0.d());
        normal_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(seed + (2ULL * (unsigned long long)(c0 + col) + 1ULL) * 0x51ED27ULL, n, Acol);
      }
      double* t = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1, g->VT.d());                     // V' r
      mul_pad_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(t, g->se.d(), N, N, g->g0.d());   // E^1/2 .
      double* a = kron_apply(c, g, G, g->g0.d(), g->g1.d(), -1, g->Vf.d());                      // a ~ N(0, Ku)
      vec_scale_copy_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, sn, Acol, Acol);          // randn / sW
      for_p(G.p, [&](auto P) {
        interp_gather_bz_kernel<decltype(P)::value><<<nb, 256, 0, st>>>(G, S.X.d(), n, D, a, 1.0, Acol, Acol, nullptr);   // A = M a + randn/sW
      });
      c->stats.kernel_launches += 5;
    }
    // A <- post.L(A) = -sW CG(B, sW A)  (the sign drops out of z.^2)
    if (pre) {
      rc = pcg_solve(c, g, G, slot, sw, mo.opt.cg_tol, maxit, g->pvA.d(), g->pvS.d(), false); if (rc) return rc;
    } else {
      rc = cg_solve_sorted(c, g, G, slot, g->pvA.d(), n, mb, sw, mo.opt.cg_tol, maxit); if (rc) return rc;
      for (int col = 0; col < mb; ++col)
        cg_unsort_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->sx.d() + (long)col * g->npad, (const int*)g->perm.p, n, sw, g->pvS.d() + (long)col * n);
    }
    worst_it = std::max(worst_it, c->stats.cg_iters); worst_rel = std::max(worst_rel, c->stats.cg_relres);
    for (int col = 0; col < mb; ++col) {                                   // z = Ku M' A ; vg += z.^2 / ns
      scatter(c, G, S.X.d(), n, D, g->pvS.d() + (long)col * n, 1.0, tg);
      kron_to(c, g, G, tg, zg, -1);
      sq_acc_kernel<<<(unsigned)((N + 255) / 256), 256, 0, st>>>(N, 1.0 / (double)ns, zg, g->vg.d());
      c->stats.kernel_launches += 2;
    }
  }
  c->stats.cg_iters = worst_it; c->stats.cg_relres = worst_rel;
  absmax_flag_kernel<<<1, 1024, 0, st>>>(N, g->vg.d(), g->out.d() + 10);
  double l1 = 0.0;
  KGP_CUDA(c, cudaMemcpyAsync(&l1, g->out.d() + 10, 8, cudaMemcpyDeviceToHost, st));
  if (vg_out) KGP_CUDA(c, cudaMemcpyAsync(vg_out, g->vg.p, N * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  g->vg_nonzero = l1 > 1e-10;                                               // infGrid.m:141
  g->vg_slot = slot; g->vg_slot_ver = S.version; g->vg_model_ver = mo.version; g->vg_grid_ver = c->grid.version;
  return KGP_OK;
}
static bool vg_valid(kgp_ctx* c, GridState* g, int slot) {
  return g->vg_slot == slot && g->vg_slot_ver == c->slots[slot].version && g->vg_model_ver == c->model.version &&
         g->vg_grid_ver == c->grid.version;
}

// post.predict (infGrid.m:135-145): fmu = ms + Ms (Kg Mx' alpha); fs2 = max(k** - Ms vg, 0), vg = 0 unless opt.pred_var
int grid_predict(kgp_ctx* c, const double* Xs, long ns, double* ymu, double* ys2, double* fmu, double* fs2) {
  GridState* g = gs(c);
  const int slot = KGP_SLOT_TR;
  if (!gstate_valid(c, g, slot, 0)) {
    int rc = grid_eval(c, slot, 1, nullptr, nullptr);     // gpml.py:22-25 always asks for dnlZ as well
    if (rc) return rc;
  }
  GridDesc G; int rc = grid_setup(c, g, G); if (rc) return rc;
  Model& m = c->model; const int D = c->slots[slot].D;
  const double sn2 = exp(2.0 * m.hyp_lik);
  cudaStream_t st = c->stream;
  KGP_CUDA(c, c->pred_X.ensure((size_t)ns * D * 8)); KGP_CUDA(c, c->pred_a.ensure((size_t)ns * 8));
  KGP_CUDA(c, c->pred_b.ensure((size_t)ns * 8));
  KGP_CUDA(c, cudaMemcpyAsync(c->pred_X.p, Xs, (size_t)ns * D * 8, cudaMemcpyDefault, st));    // host, or the projected inputs on the device (hyp.proj)
  resid_mean_kernel<<<(unsigned)((ns + 255) / 256), 256, 0, st>>>(c->pred_X.d(), nullptr, ns, D, m.mean_mask, g->hypm.d(), c->pred_b.d(), ns);   // = -ms
  for_p(G.p, [&](auto P) {
    interp_gather_bz_kernel<decltype(P)::value><<<nblocks_for(ns), 256, 0, st>>>(G, c->pred_X.d(), ns, D, g->beta.d(), 1.0, nullptr, c->pred_a.d(), nullptr);
  });
  std::vector<double> a(ns), b(ns);
  KGP_CUDA(c, cudaMemcpyAsync(a.data(), c->pred_a.p, ns * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaMemcpyAsync(b.data(), c->pred_b.p, ns * 8, cudaMemcpyDeviceToHost, st));
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  std::vector<double> ve;
  if ((fs2 || ys2) && m.opt.pred_var != 0) {                               // explained variance (infGrid.m:94,141-143)
    if (!vg_valid(c, g, slot)) {
      const int nsmp = std::max((int)ceil(fabs((double)m.opt.pred_var)), 20);
      rc = grid_pred_var(c, slot, nullptr, nullptr, nsmp, m.opt.ldB2_cheby_seed ^ 0xC0FFEEULL, nullptr); if (rc) return rc;
    }
    if (g->vg_nonzero) {
      for_p(G.p, [&](auto P) {
        interp_gather_bz_kernel<decltype(P)::value><<<nblocks_for(ns), 256, 0, st>>>(G, c->pred_X.d(), ns, D, g->vg.d(), 1.0, nullptr, c->pred_a.d(), nullptr);
      });
      ve.resize(ns);
      KGP_CUDA(c, cudaMemcpyAsync(ve.data(), c->pred_a.p, ns * 8, cudaMemcpyDeviceToHost, st));
      KGP_CUDA(c, cudaStreamSynchronize(st));
    }
  }
  for (long i = 0; i < ns; ++i) {
    const double f = a[i] - b[i];
    const double v = fmax(g->prior_var - (ve.empty() ? 0.0 : ve[i]), 0.0);
    if (fmu) fmu[i] = f; if (ymu) ymu[i] = f;
    if (fs2) fs2[i] = v; if (ys2) ys2[i] = v + sn2;
  }
  return KGP_OK;
}

// ---- covGrid('create', X, eq, k)   (apxGrid.m:389-416) ------------------------------------------
__global__ void col_minmax_kernel(const double* __restrict__ X, long n, int D, int col, double* __restrict__ out) {
  __shared__ double smin[32], smax[32];
  double lo = INFINITY, hi = -INFINITY;
  for (long i = threadIdx.x; i < n; i += blockDim.x) { double v = X[i * D + col]; lo = fmin(lo, v); hi = fmax(hi, v); }
  for (int o = 16; o > 0; o >>= 1) { lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o)); hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o)); }
  if ((threadIdx.x & 31) == 0) { smin[threadIdx.x >> 5] = lo; smax[threadIdx.x >> 5] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < (int)(blockDim.x >> 5); ++w) { lo = fmin(lo, smin[w]); hi = fmax(hi, smax[w]); }
    out[0] = lo; out[1] = hi;
  }
}
__global__ void col_gather_kernel(const double* __restrict__ X, long n, int D, int col, double* __restrict__ key,
                                  int* __restrict__ idx, long Npad) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Npad) return;
  key[i] = (i < n) ? X[i * D + col] : -INFINITY; idx[i] = (int)i;
}
__global__ void count_unique_kernel(const double* __restrict__ key, long n, unsigned long long* __restrict__ cnt) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (i == 0 || key[i] != key[i - 1]) atomicAdd(cnt, 1ULL);     // integer count: order-independent
}

// 1-D k-means of the sorted unique values u (ascending) into K clusters, for covGrid('create', x, 0, k) (apxGrid.m:411).  The
// reference calls kmeans() of Octave's statistics package, whose random initialisation cannot be reproduced; this is Lloyd's
// algorithm from an equal-count initial partition, deterministic, every step in a fixed order of FP64 operations so that a
// CPU restatement can agree to the last bit (the parity tests check that).  In one dimension the nearest-centroid rule is a set
// of thresholds at the midpoints of neighbouring centroids; segment means come from sequential prefix sums.
int kmeans1d_host(const double* u, long nu, long K, double* cent) {
  if (!u || !cent || K < 1 || K > nu) return -1;
  std::vector<double> S(nu + 1); S[0] = 0.0;
  for (long i = 0; i < nu; ++i) S[i + 1] = S[i] + u[i];
  std::vector<long> bnd(K + 1), nb(K + 1);
  for (long cc = 0; cc <= K; ++cc) bnd[cc] = (cc * nu) / K;
  for (long cc = 0; cc < K; ++cc) cent[cc] = (S[bnd[cc + 1]] - S[bnd[cc]]) / (double)(bnd[cc + 1] - bnd[cc]);
  for (int it = 0; it < 200; ++it) {
    nb[0] = 0; nb[K] = nu;
    for (long cc = 0; cc + 1 < K; ++cc) {
      const double mid = 0.5 * (cent[cc] + cent[cc + 1]);                  // a tie goes to the lower cluster
      nb[cc + 1] = (long)(std::upper_bound(u, u + nu, mid) - u);
      if (nb[cc + 1] < nb[cc]) nb[cc + 1] = nb[cc];
    }
    if (nb == bnd) break;
    bnd = nb;
    for (long cc = 0; cc < K; ++cc)
      if (bnd[cc + 1] > bnd[cc]) cent[cc] = (S[bnd[cc + 1]] - S[bnd[cc]]) / (double)(bnd[cc + 1] - bnd[cc]);   // an emptied cluster keeps its centre
  }
  return 0;
}

int grid_create(kgp_ctx* c, int slot, int eq, double k) {
  GridState* g = gs(c); Slot& S = c->slots[slot];
  const long n = S.n; const int D = S.D;
  if (D > GRID_MAXP) return c->fail(KGP_BAD_ARG, "at most %d grid dimensions", GRID_MAXP);
  cudaStream_t st = c->stream;
  KGP_CUDA(c, g->out.ensure(64 * 8));
  std::vector<std::vector<double>> axes;
  std::vector<double> uq;                             // sorted unique values of a column (eq = 0 only)
  for (int j = 0; j < D; ++j) {
    double mm[2];
    col_minmax_kernel<<<1, 1024, 0, st>>>(S.X.d(), n, D, j, g->out.d());
    KGP_CUDA(c, cudaMemcpyAsync(mm, g->out.p, 16, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    if (!(mm[1] > mm[0])) return c->fail(KGP_BAD_ARG, "Two few unique points.");
    double ngj;                                       // stays fractional in du (apxGrid.m:404-406); only linspace floors it
    const bool frac = (k >= 0.0 && k <= 1.0);
    if (frac || !eq) {                                // u = sort(unique(x(:,j))): descending bitonic sort, padding -inf at the end
      long Npad = 1; while (Npad < n) Npad <<= 1;
      KGP_CUDA(c, g->sortk.ensure(Npad * 8)); KGP_CUDA(c, g->sorti.ensure(Npad * 4));
      col_gather_kernel<<<(unsigned)((Npad + 255) / 256), 256, 0, st>>>(S.X.d(), n, D, j, g->sortk.d(), (int*)g->sorti.p, Npad);
      for (long kk = 2; kk <= Npad; kk <<= 1)
        for (long jj = kk >> 1; jj > 0; jj >>= 1)
          bitonic_stage_kernel<<<(unsigned)((Npad + 255) / 256), 256, 0, st>>>(g->sortk.d(), (int*)g->sorti.p, Npad, jj, kk);
      unsigned long long* cnt = (unsigned long long*)(g->out.d() + 8);
      KGP_CUDA(c, cudaMemsetAsync(cnt, 0, 8, st));
      count_unique_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(g->sortk.d(), n, cnt);
      unsigned long long hc = 0;
      KGP_CUDA(c, cudaMemcpyAsync(&hc, cnt, 8, cudaMemcpyDeviceToHost, st));
      if (!eq) { uq.resize(n); KGP_CUDA(c, cudaMemcpyAsync(uq.data(), g->sortk.p, (size_t)n * 8, cudaMemcpyDeviceToHost, st)); }
      KGP_CUDA(c, cudaStreamSynchronize(st));
      ngj = frac ? ceil(k * (double)hc) : k;          // fraction of the unique values (apxGrid.m:402-403)
    } else ngj = k;
    if (!(ngj >= 1.0)) return c->fail(KGP_BAD_ARG, "bad grid size k=%g", k);
    const double du = (mm[1] - mm[0]) / ngj, b0 = mm[0] - 5.0 * du, b1 = mm[1] + 5.0 * du;
    auto linspace = [](double lo, double hi, long np_) {            // Octave linspace: base + i*delta, last point exact
      std::vector<double> v(np_);
      const double delta = (np_ > 1) ? (hi - lo) / (double)(np_ - 1) : 0.0;
      for (long i = 0; i < np_; ++i) v[i] = lo + (double)i * delta;
      if (np_ > 0) v[np_ - 1] = hi;
      return v;
    };
    if (eq) {
      axes.push_back(linspace(b0, b1, (long)floor(std::max(ngj, 5.0))));   // Octave's linspace(a, b, n) uses floor(n) points
    } else {                                          // non-equispaced axis: cluster centres + boundary points (apxGrid.m:410-415)
      std::reverse(uq.begin(), uq.end());
      uq.erase(std::unique(uq.begin(), uq.end()), uq.end());
      const long nu = (long)uq.size(), ngi = (long)floor(ngj);
      const long K = std::min(nu, ngi - 2);
      if (K < 1) return c->fail(KGP_BAD_ARG, "covGrid('create', x, 0, k): k=%g leaves no cluster centres (need k >= 3)", k);
      std::vector<double> cent(K);
      if (kmeans1d_host(uq.data(), nu, K, cent.data())) return c->fail(KGP_BAD_ARG, "kmeans failed");
      const long nb = ngi - K, nb1 = nb / 2, nb2 = nb - nb1;
      std::vector<double> x1 = linspace(b0, cent[0], nb1 + 1), x2 = linspace(cent[K - 1], b1, nb2 + 1), ax;
      ax.insert(ax.end(), x1.begin(), x1.begin() + nb1);
      ax.insert(ax.end(), cent.begin(), cent.end());
      ax.insert(ax.end(), x2.begin() + 1, x2.end());
      if (ax.size() < 4) return c->fail(KGP_BAD_ARG, "grid axis %d needs >= 4 points", j);
      for (size_t i = 1; i < ax.size(); ++i)
        if (!(ax[i] - ax[i - 1] >= 1e-10)) return c->fail(KGP_BAD_ARG, "k-means grid axis %d is not strictly increasing (an emptied cluster)", j);
      axes.push_back(ax);
    }
  }
  KGP_CUDA(c, cudaGetLastError());
  c->grid.axes = axes; c->grid.toep.assign(axes.size(), 0); c->grid.neq.assign(axes.size(), 0);
  // interpgrid dispatches per axis on equi() (apxGrid.m:684-695): a k-means axis runs neqinterp unless it happens to be equispaced
  if (!eq) for (size_t a = 0; a < axes.size(); ++a) c->grid.neq[a] = kgp_grid_axis_equispaced(axes[a].data(), (int)axes[a].size()) ? 0 : 1;
  c->grid.set = true; c->grid.version++;
  return KGP_OK;
}

int grid_init(kgp_ctx* c) {
  KGP_CUDA(c, cudaFuncSetAttribute(jacobi_eig_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 190 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(kron2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(mode_product_rb_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(interp_scatter_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(interp_scatter_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(interp_scatter_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  KGP_CUDA(c, cudaFuncSetAttribute(interp_scatter_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  return KGP_OK;
}

}  // namespace kgp
