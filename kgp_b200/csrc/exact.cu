// Host orchestration of the exact-GP path on one B200.
//
// What the reference does per evaluation (GPML/inf/infGaussLik.m:19-33 with
// GPML/cov/apx.m:141-173): K, L = chol(I + K/sn2), Q = B^{-1} by two n-RHS
// triangular solves, alpha, nlZ, dnlZ; then dlikExact.m re-runs it all for dnlZ/dX.
// Here: one blocked FP64 Cholesky, one triangular inverse Z = L^{-T}, one product
// B^{-1} = Z Z^T (n^3 flops in total instead of 7n^3/3), all O(n^3) work in
// dgemm_nt_kernel (DMMA), and ONE fused O(n^2) pass for every gradient.
//
// Storage (row-major, order np = n rounded up to 128, identity padding):
//   M  : lower triangle = B, then L (kept: prediction solves with it)
//   W2 : upper triangle = Z = L^{-T} (so that every GEMM operand is K-contiguous),
//        then the full symmetric B^{-1}
//   Dws: the diagonal NB x NB blocks of B^{-1} while W2's diagonal blocks still hold Z
#include "ctx.h"
#include "exact_kernels2.cuh"
#include "exact_dist.cuh"
#include "exact_sm.cuh"
#include <math.h>
#include <string.h>

namespace kgp {

static inline double count_gemm(kgp_ctx* c, const GemmArgs& g) {
  c->stats.gemm_launches++; c->stats.kernel_launches++;
  double f = 0.0;
  if (g.nbatch > 0) {
    for (int p = 0; p < g.nbatch; ++p) f += 2.0 * 128.0 * 128.0 * (double)(g.kend - g.kbeg) * (g.ntm - g.bat_roff[p] / 128) * g.ntn;
    c->stats.gemm_flops += f;
    return f;
  }
  for (int tm = 0; tm < g.ntm; ++tm) {
    int ntn = (g.shape == GEMM_LOWER) ? (tm + 1 < g.ntn ? tm + 1 : g.ntn) : g.ntn;
    int kb = g.kbeg + (g.ktri == KTRI_FROM_M0 ? tm * 128 : 0);
    f += 2.0 * 128.0 * 128.0 * (double)(g.kend - kb) * ntn;
  }
  if (g.own_mod > 1) f /= g.own_mod;
  c->stats.gemm_flops += f;
  return f;
}
// launches of the reference-sized problems (np <= KGP_SMALL_NP, default 2048) that cannot fill the GPU with 128-tiles go
// to the small-tile kernel; the decision depends only on the problem, never on the number of ranks
static const GemmSmallWs* small_ws(kgp_ctx* c, const GemmArgs& g, GemmSmallWs& sw) {
  static const int small_np = getenv("KGP_SMALL_NP") ? atoi(getenv("KGP_SMALL_NP")) : 2048;
  if (c->np > small_np || c->nranks > 1) return nullptr;
  int nchmax = 0; size_t need = 0;
  if (!gemm_small_eligible(g, gemm_num_tiles(g), nchmax, need)) return nullptr;
  const int w = (c->stream_p && c->stream == c->stream_p) ? 1 : 0;
  constexpr long NCNT = 96 * 16;
  if (!c->gcnt[w].p) {
    if (c->gcnt[w].ensure(NCNT * sizeof(unsigned)) != cudaSuccess) return nullptr;
    cudaMemsetAsync(c->gcnt[w].p, 0, NCNT * sizeof(unsigned), c->stream);
  }
  if (need > c->gws[w].bytes) {
    cudaStreamSynchronize(c->stream);                      // a launch in flight may still use the old buffer
    if (c->gws[w].ensure(need) != cudaSuccess) return nullptr;
  }
  sw.ws = c->gws[w].d(); sw.ws_bytes = c->gws[w].bytes; sw.cnt = (unsigned*)c->gcnt[w].p; sw.cnt_n = NCNT;
  return &sw;
}
static inline void gemm(kgp_ctx* c, const GemmArgs& g, bool tile_ctas = false) {
  if (g.ntm <= 0 || g.ntn <= 0) return;
  const double fl = count_gemm(c, g);
  GemmSmallWs sws; const GemmSmallWs* sw = small_ws(c, g, sws);
  if (c->profiling) {
    if (c->gev_used + 2 > c->gev.size()) {
      size_t old = c->gev.size(); c->gev.resize(old + 256);
      for (size_t i = old; i < c->gev.size(); ++i) cudaEventCreate(&c->gev[i]);
    }
    cudaEventRecord(c->gev[c->gev_used], c->stream);
    gemm_launch(g, c->stream, tile_ctas, sw);
    cudaEventRecord(c->gev[c->gev_used + 1], c->stream);
    c->gev_class.push_back(c->gemm_class); c->gev_flops.push_back(fl);
    c->gev_used += 2;
  } else gemm_launch(g, c->stream, tile_ctas, sw);
}
static void gemm_prof_collect(kgp_ctx* c) {
  double ms = 0.0;
  for (size_t i = 0; i + 1 < c->gev_used; i += 2) {
    float t = 0; cudaEventElapsedTime(&t, c->gev[i], c->gev[i + 1]); ms += t;
    int cl = c->gev_class[i / 2] & 7; c->stats.gemm_class_ms[cl] += t; c->stats.gemm_class_flops[cl] += c->gev_flops[i / 2];
  }
  c->stats.ms_gemm = ms; c->gev_used = 0; c->gev_class.clear(); c->gev_flops.clear();
}
static inline GemmArgs mkgemm(const double* A, long lda, const double* B, long ldb, double* C, long ldc,
                              int ntm, int ntn, int kbeg, int kend, int shape, int ktri, double alpha, double beta) {
  GemmArgs g; memset(&g, 0, sizeof g);
  g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.C = C; g.ldc = ldc; g.ntm = ntm; g.ntn = ntn;
  g.kbeg = kbeg; g.kend = kend; g.shape = shape; g.ktri = ktri; g.alpha = alpha; g.beta = beta;
  g.Cdiag = nullptr; g.nb = NB; g.cdiag_only = 0;
  return g;
}

// 32-row CTAs while a launch has too few 128-row blocks to occupy the GPU (the reference-sized problems)
static inline bool trsm_small(kgp_ctx* c, long mrows) {
  static const int on = getenv("KGP_SMALL_TRSM") ? atoi(getenv("KGP_SMALL_TRSM")) : 1;
  return on && mrows <= 4096;
}
// X (mrows x w, leading dim ldx) <- X * Lblk^{-T}, Lblk lower w x w (w multiple of 128, <= NB)
static void trsm_block(kgp_ctx* c, double* X, long ldx, long mrows, const double* Lblk, long ldl, int w) {
  const int nc = w / 128;
  for (int cb = 0; cb < nc; ++cb) {
    trsm128_rows_launch(X + cb * 128, ldx, mrows, Lblk + (long)cb * 128 * ldl + cb * 128, ldl, c->stream, trsm_small(c, mrows));
    c->stats.kernel_launches++;
    if (cb + 1 < nc) c->gemm_class = 3;
    if (cb + 1 < nc)
      gemm(c, mkgemm(X + cb * 128, ldx, Lblk + (long)(cb + 1) * 128 * ldl + cb * 128, ldl,
                     X + (cb + 1) * 128, ldx, (int)(mrows / 128), nc - cb - 1, 0, 128, GEMM_FULL, KTRI_NONE, -1.0, 1.0));
  }
}

// one NB-wide panel of the blocked Cholesky: factor the diagonal block and the block column below it (on c->stream)
static void factor_panel(kgp_ctx* c, long k0, int w) {
  double* M = c->M.d(); const long ld = c->np, np = c->np;
  for (int cb = 0; cb < w / 128; ++cb) {
    const long j = k0 + cb * 128, r0 = j + 128, mrows = np - r0;
    potrf_diag128_kernel<<<1, PD_THREADS, PD_SMEM, c->stream>>>(M + j * ld + j, ld, (int)j, c->d_info);
    c->stats.kernel_launches++;
    if (mrows > 0) {
      trsm128_rows_launch(M + r0 * ld + j, ld, mrows, M + j * ld + j, ld, c->stream, trsm_small(c, mrows));
      c->stats.kernel_launches++;
      const long ncols = k0 + w - r0;
      c->gemm_class = 0;
      if (ncols > 0)
        gemm(c, mkgemm(M + r0 * ld + j, ld, M + r0 * ld + j, ld, M + r0 * ld + r0, ld,
                       (int)(mrows / 128), (int)(ncols / 128), 0, 128, GEMM_LOWER, KTRI_NONE, -1.0, 1.0));
    }
  }
}
// blocked right-looking Cholesky of the lower triangle of M (two levels: NB / 128)
static void potrf(kgp_ctx* c) {
  double* M = c->M.d(); const long ld = c->np, np = c->np;
  for (long k0 = 0; k0 < np; k0 += NB) {
    const int w = (int)((np - k0 < NB) ? np - k0 : NB);
    factor_panel(c, k0, w);
    const long r0 = k0 + w, mt = (np - r0) / 128;
    c->gemm_class = 1;
    if (mt > 0)
      gemm(c, mkgemm(M + r0 * ld + k0, ld, M + r0 * ld + k0, ld, M + r0 * ld + r0, ld,
                     (int)mt, (int)mt, 0, w, GEMM_LOWER, KTRI_NONE, -1.0, 1.0));
  }
}

// high-priority panel stream + its events (also used by the multi-GPU factorisation)
static int ensure_panel_stream(kgp_ctx* c) {
  if (c->stream_p) return KGP_OK;
  int lo = 0, hi = 0; cudaDeviceGetStreamPriorityRange(&lo, &hi);
  KGP_CUDA(c, cudaStreamCreateWithPriority(&c->stream_p, cudaStreamNonBlocking, hi));
  for (int i = 0; i < 8; ++i) KGP_CUDA(c, cudaEventCreateWithFlags(&c->dev_ev[i], cudaEventDisableTiming));
  for (int i = 0; i < 8; ++i) KGP_CUDA(c, cudaEventCreateWithFlags(&c->zev[i], cudaEventDisableTiming));
  return KGP_OK;
}
// Single-GPU Cholesky with look-ahead 1: panel k+1 is factored on the high-priority panel stream while the rest of
// trailing update k runs; the updates use one CTA per tile so that the panel kernels get SMs at tile granularity.
static int potrf_lookahead(kgp_ctx* c) {
  int rc = ensure_panel_stream(c); if (rc) return rc;
  double* M = c->M.d(); const long ld = c->np, np = c->np;
  cudaStream_t sm = c->stream, sp = c->stream_p;
  cudaEvent_t* ev_panel = c->dev_ev; cudaEvent_t ev_la = c->dev_ev[4];
  KGP_CUDA(c, cudaEventRecord(ev_la, sm));
  int k = 0;
  for (long k0 = 0; k0 < np; k0 += NB, ++k) {
    const int w = (int)((np - k0 < NB) ? np - k0 : NB);
    KGP_CUDA(c, cudaStreamWaitEvent(sp, ev_la, 0));
    c->stream = sp; factor_panel(c, k0, w); c->stream = sm;
    KGP_CUDA(c, cudaEventRecord(ev_panel[k & 1], sp));
    KGP_CUDA(c, cudaStreamWaitEvent(sm, ev_panel[k & 1], 0));
    const long r0 = k0 + w, mt = (np - r0) / 128;
    if (mt <= 0) continue;
    const int w1 = (int)((np - r0 < NB) ? np - r0 : NB), nt1 = w1 / 128;
    c->gemm_class = 1;
    gemm(c, mkgemm(M + r0 * ld + k0, ld, M + r0 * ld + k0, ld, M + r0 * ld + r0, ld, (int)mt, nt1, 0, w, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), true);
    KGP_CUDA(c, cudaEventRecord(ev_la, sm));                       // block column k+1 is up to date: its panel may be factored
    const long r1 = r0 + w1;
    if (mt > nt1)
      gemm(c, mkgemm(M + r1 * ld + k0, ld, M + r1 * ld + k0, ld, M + r1 * ld + r1, ld, (int)(mt - nt1), (int)(mt - nt1), 0, w,
                     GEMM_LOWER, KTRI_NONE, -1.0, 1.0), true);
  }
  return KGP_OK;
}

// Z = L^{-T} into the upper triangle of W2, block column by block column
static void zsweep(kgp_ctx* c) {
  const double* M = c->M.d(); double* W = c->W2.d(); const long ld = c->np, np = c->np;
  for (long k0 = 0; k0 < np; k0 += NB) {
    const int w = (int)((np - k0 < NB) ? np - k0 : NB);
    c->gemm_class = 2;
    if (k0 > 0)
      gemm(c, mkgemm(W, ld, M + k0 * ld, ld, W + k0, ld, (int)(k0 / 128), w / 128, 0, (int)k0,
                     GEMM_FULL, KTRI_FROM_M0, -1.0, 0.0));
    set_identity_kernel<<<(unsigned)(((long)w * w + 255) / 256), 256, 0, c->stream>>>(W + k0 * ld + k0, ld, w);
    c->stats.kernel_launches++;
    trsm_block(c, W + k0, ld, k0 + w, M + k0 * ld + k0, ld, w);
  }
}

// B^{-1} = Z Z^T : off-diagonal NB-blocks into W2's lower triangle, diagonal blocks into Dws
static void zzt(kgp_ctx* c) {
  double* W = c->W2.d(); const long ld = c->np; const int nt = (int)(c->np / 128);
  GemmArgs g = mkgemm(W, ld, W, ld, W, ld, nt, nt, 0, (int)c->np, GEMM_LOWER, KTRI_FROM_M0, 1.0, 0.0);
  g.Cdiag = c->Dws.d(); g.nb = NB; g.cdiag_only = 0;
  c->gemm_class = 4;
  gemm(c, g);
  dim3 grid((unsigned)(c->np / 32), (unsigned)(c->np / 32)), blk(32, 8);
  mirror_finalize_kernel<<<grid, blk, 0, c->stream>>>(W, ld, c->np, c->Dws.d(), NB);
  c->stats.kernel_launches++;
}

// ===== multi-GPU variants (see exact_dist.cuh) ============================================================
int dist_unique_id(char* out128, const char* libpath, std::string& err) {
  NcclApi& api = nccl_api(libpath);
  if (!api.ok) { err = api.err; return KGP_CUDA_ERR; }
  kgp_ncclUniqueId id; int r = api.GetUniqueId(&id);
  if (r != 0) { err = api.GetErrorString(r); return KGP_CUDA_ERR; }
  memcpy(out128, id.internal, 128);
  return KGP_OK;
}
int dist_init(kgp_ctx* c, int nranks, int rank, const char* id128, const char* libpath) {
  if (nranks < 1 || rank < 0 || rank >= nranks) return c->fail(KGP_BAD_ARG, "bad rank %d of %d", rank, nranks);
  if (nranks == 1) { c->nranks = 1; c->rank = 0; return KGP_OK; }
  NcclApi& api = nccl_api(libpath);
  if (!api.ok) return c->fail(KGP_CUDA_ERR, "%s", api.err.c_str());
  kgp_ncclUniqueId id; memcpy(id.internal, id128, 128);
  KGP_CUDA(c, cudaSetDevice(c->device));
  KGP_NCCL(c, api, api.CommInitRank(&c->nccl_comm, nranks, id, rank));
  c->nranks = nranks; c->rank = rank;
  // peer window for the row-sharded KISS-GP CG: allocate, exchange the IPC handles over the new communicator, map
  c->px_ok = false;
  if (nranks <= 8 && !(getenv("KGP_DIST_CG") && atoi(getenv("KGP_DIST_CG")) == 0)) {
    cudaIpcMemHandle_t mine; char* hdev = nullptr; std::vector<cudaIpcMemHandle_t> all(nranks);
    bool ok = cudaMalloc(&c->px_local, PX_BYTES) == cudaSuccess && cudaMemset(c->px_local, 0, PX_BYTES) == cudaSuccess &&
              cudaMalloc((void**)&c->px_abort, 64) == cudaSuccess && cudaMemset(c->px_abort, 0, 64) == cudaSuccess &&
              cudaIpcGetMemHandle(&mine, c->px_local) == cudaSuccess &&
              cudaMalloc((void**)&hdev, (size_t)nranks * sizeof(mine)) == cudaSuccess;
    if (ok) {
      cudaMemcpy(hdev + (size_t)rank * sizeof(mine), &mine, sizeof(mine), cudaMemcpyHostToDevice);
      for (int r = 0; r < nranks && ok; ++r)
        ok = api.Broadcast(hdev + (size_t)r * sizeof(mine), hdev + (size_t)r * sizeof(mine), sizeof(mine), KGP_NCCL_INT8, r, c->nccl_comm, c->stream) == 0;
      ok = ok && cudaStreamSynchronize(c->stream) == cudaSuccess &&
           cudaMemcpy(all.data(), hdev, (size_t)nranks * sizeof(mine), cudaMemcpyDeviceToHost) == cudaSuccess;
      for (int r = 0; r < nranks && ok; ++r) {
        if (r == rank) { c->px_peer[r] = c->px_local; continue; }
        ok = cudaIpcOpenMemHandle(&c->px_peer[r], all[r], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      }
    }
    if (hdev) cudaFree(hdev);
    if (!ok) { cudaGetLastError(); if (getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] rank %d: peer window unavailable, KISS-GP CG stays replicated\n", rank); }
    // every rank must agree: all-reduce the flag
    double* flag = nullptr; double hv = ok ? 1.0 : 0.0;
    if (cudaMalloc((void**)&flag, 8) == cudaSuccess) {
      cudaMemcpy(flag, &hv, 8, cudaMemcpyHostToDevice);
      api.AllReduce(flag, flag, 1, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, c->stream);
      cudaStreamSynchronize(c->stream);
      cudaMemcpy(&hv, flag, 8, cudaMemcpyDeviceToHost); cudaFree(flag);
      c->px_ok = hv > (double)nranks - 0.5;
    }
    c->px_epoch = 0;
  }
  return ensure_panel_stream(c);
}
void dist_free(kgp_ctx* c) {
  if (c->nccl_comm) { NcclApi& api = nccl_api(nullptr); if (api.ok) api.CommDestroy(c->nccl_comm); c->nccl_comm = nullptr; }
  if (c->stream_p) { cudaStreamDestroy(c->stream_p); for (int i = 0; i < 8; ++i) { cudaEventDestroy(c->dev_ev[i]); cudaEventDestroy(c->zev[i]); } c->stream_p = nullptr; }
  c->PB[0].release(); c->PB[1].release();
  for (int r = 0; r < 8; ++r) {
    if (c->w2_peer[r] && r != c->rank) cudaIpcCloseMemHandle(c->w2_peer[r]);
    c->w2_peer[r] = nullptr;
    if (c->zstream[r]) { cudaStreamDestroy(c->zstream[r]); c->zstream[r] = nullptr; }
    if (c->zpush_ev[r]) { cudaEventDestroy(c->zpush_ev[r]); c->zpush_ev[r] = nullptr; }
  }
  if (c->zcol_ev) { cudaEventDestroy(c->zcol_ev); c->zcol_ev = nullptr; }
  if (c->zbar) { cudaFree(c->zbar); c->zbar = nullptr; }
  c->w2_mapped = nullptr; c->w2_push_ok = false;
  for (void* q : c->w2_retired) cudaFree(q);
  c->w2_retired.clear();
  for (int r = 0; r < 8; ++r) { if (c->px_peer[r] && c->px_peer[r] != c->px_local) cudaIpcCloseMemHandle(c->px_peer[r]); c->px_peer[r] = nullptr; }
  if (c->px_local) { cudaFree(c->px_local); c->px_local = nullptr; }
  if (c->px_abort) { cudaFree(c->px_abort); c->px_abort = nullptr; }
  c->px_ok = false;
}

// host-synchronous sum of one double over the ranks (set-up decisions every rank must take alike)
static int dist_allreduce_host(kgp_ctx* c, double* v) {
  NcclApi& api = nccl_api(nullptr);
  if (!c->zbar) KGP_CUDA(c, cudaMalloc((void**)&c->zbar, 64));
  KGP_CUDA(c, cudaMemcpyAsync(c->zbar, v, 8, cudaMemcpyHostToDevice, c->stream));
  KGP_NCCL(c, api, api.AllReduce(c->zbar, c->zbar, 1, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, c->stream));
  KGP_CUDA(c, cudaMemcpyAsync(v, c->zbar, 8, cudaMemcpyDeviceToHost, c->stream));
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGP_OK;
}
// Z's all-gather runs as peer copies straight into the other ranks' W2 (zsweep_dist), so every rank needs a mapping of every
// peer's W2.  Called at the start of an evaluation, collectively, BEFORE W2 may be (re)allocated: if any rank is about to
// reallocate, or holds an allocation its peers have not mapped, every rank closes its mappings first (an exporter must not free
// memory a peer still maps), then allocates, then the IPC handles are exchanged over the communicator.  Steady state: one 8-byte
// all-reduce per evaluation.  On failure (no peer access) the broadcast all-gather stays in use -- decided alike on all ranks.
static int w2_peer_sync(kgp_ctx* c, size_t need_bytes) {
  NcclApi& api = nccl_api(nullptr);
  const int P = c->nranks, me = c->rank;
  double flag = (need_bytes > c->W2.bytes || !c->W2.p || c->W2.p != c->w2_mapped) ? 1.0 : 0.0;
  int rc = dist_allreduce_host(c, &flag); if (rc) return rc;
  if (flag == 0.0) return KGP_OK;
  for (int r = 0; r < P; ++r) { if (c->w2_peer[r] && r != me) cudaIpcCloseMemHandle(c->w2_peer[r]); c->w2_peer[r] = nullptr; }
  c->w2_mapped = nullptr; c->w2_push_ok = false;
  double zero = 0.0;
  rc = dist_allreduce_host(c, &zero); if (rc) return rc;          // every rank has closed its mappings: owners may free now
  for (void* q : c->w2_retired) cudaFree(q);
  c->w2_retired.clear();
  KGP_CUDA(c, c->W2.ensure(need_bytes));
  cudaIpcMemHandle_t mine; char* hdev = nullptr; std::vector<cudaIpcMemHandle_t> all(P);
  bool ok = cudaIpcGetMemHandle(&mine, c->W2.p) == cudaSuccess && cudaMalloc((void**)&hdev, (size_t)P * sizeof(mine)) == cudaSuccess;
  double okall = ok ? 1.0 : 0.0;
  rc = dist_allreduce_host(c, &okall); if (rc) { if (hdev) cudaFree(hdev); return rc; }
  if (okall > (double)P - 0.5) {
    cudaMemcpy(hdev + (size_t)me * sizeof(mine), &mine, sizeof(mine), cudaMemcpyHostToDevice);
    for (int r = 0; r < P && ok; ++r)
      ok = api.Broadcast(hdev + (size_t)r * sizeof(mine), hdev + (size_t)r * sizeof(mine), sizeof(mine), KGP_NCCL_INT8, r, c->nccl_comm, c->stream) == 0;
    ok = ok && cudaStreamSynchronize(c->stream) == cudaSuccess &&
         cudaMemcpy(all.data(), hdev, (size_t)P * sizeof(mine), cudaMemcpyDeviceToHost) == cudaSuccess;
    for (int r = 0; r < P && ok; ++r) {
      if (r == me) { c->w2_peer[r] = c->W2.p; continue; }
      ok = cudaIpcOpenMemHandle(&c->w2_peer[r], all[r], cudaIpcMemLazyEnablePeerAccess) == cudaSuccess;
      if (!ok) c->w2_peer[r] = nullptr;
    }
    for (int r = 0; r < P && ok; ++r) {
      if (r == me) continue;
      if (!c->zstream[r]) ok = cudaStreamCreateWithFlags(&c->zstream[r], cudaStreamNonBlocking) == cudaSuccess;
      if (ok && !c->zpush_ev[r]) ok = cudaEventCreateWithFlags(&c->zpush_ev[r], cudaEventDisableTiming) == cudaSuccess;
    }
    if (ok && !c->zcol_ev) ok = cudaEventCreateWithFlags(&c->zcol_ev, cudaEventDisableTiming) == cudaSuccess;
    if (!ok) cudaGetLastError();
    okall = ok ? 1.0 : 0.0;
    rc = dist_allreduce_host(c, &okall);
    if (rc) { if (hdev) cudaFree(hdev); return rc; }
  }
  if (hdev) cudaFree(hdev);
  c->w2_mapped = c->W2.p;                                          // (also when mapping failed: do not retry every evaluation)
  c->w2_push_ok = okall > (double)P - 0.5;
  if (!c->w2_push_ok && getenv("KGP_DEBUG")) fprintf(stderr, "[kgp] rank %d: peer mapping of W2 unavailable, Z is gathered by broadcasts\n", me);
  return KGP_OK;
}

// Cholesky with block columns dealt cyclically over the ranks, panel broadcast, look-ahead 1
static int potrf_dist(kgp_ctx* c) {
  NcclApi& api = nccl_api(nullptr);
  double* M = c->M.d(); const long ld = c->np, np = c->np;
  const int P = c->nranks, me = c->rank, nblk = (int)((np + NB - 1) / NB);
  cudaStream_t sm = c->stream, sp = c->stream_p;
  cudaEvent_t *ev_panel = c->dev_ev, *ev_free = c->dev_ev + 2, ev_la = c->dev_ev[4], ev_second = c->dev_ev[5];
  for (int b = 0; b < 2; ++b) KGP_CUDA(c, c->PB[b].ensure((size_t)np * NB * 8));
  // trailing updates run one CTA per tile so that the high-priority panel stream (next panel's factorisation, the NCCL
  // broadcast kernel) gets SMs at tile granularity; persistent CTAs would hold every SM until the whole update ends
  static const bool tile_mode = !(getenv("KGP_DIST_TILE_CTAS") && atoi(getenv("KGP_DIST_TILE_CTAS")) == 0);
  // look-ahead 2 (opt-in, KGP_DIST_LA2=1; measured neutral at 4 GPUs: potrf 229.6 vs 231.0 ms): at step k the main
  // stream updates column k+2 first ("second"), then the rest; the update
  // of column k+1 with panel k ("first") runs on the PANEL stream right in front of panel k+1's factorisation and only
  // waits for second(k-1) -- the serial chain is broadcast -> first -> factor, independent of the bulk update.
  static const bool la2 = getenv("KGP_DIST_LA2") && atoi(getenv("KGP_DIST_LA2")) != 0;
  KGP_CUDA(c, cudaEventRecord(ev_la, sm));                          // block column 0 is ready once the kernel matrix is built
  KGP_CUDA(c, cudaEventRecord(ev_second, sm));
  for (int k = 0; k < nblk; ++k) {
    const long k0 = (long)k * NB; const int w = (int)((np - k0 < NB) ? np - k0 : NB);
    const long rows = np - k0; const int owner = dist_col_owner(k, P), b = k & 1;
    double* PB = c->PB[b].d();
    // ---- panel stream: factor (owner), broadcast, unpack
    if (k >= 2) KGP_CUDA(c, cudaStreamWaitEvent(sp, ev_free[b], 0));
    if (me == owner) {
      if (!la2 || k == 0) KGP_CUDA(c, cudaStreamWaitEvent(sp, ev_la, 0));
      c->stream = sp; factor_panel(c, k0, w); c->stream = sm;
      KGP_CUDA(c, cudaMemcpy2DAsync(PB, (size_t)w * 8, M + k0 * ld + k0, (size_t)ld * 8, (size_t)w * 8, (size_t)rows, cudaMemcpyDeviceToDevice, sp));
    }
    KGP_NCCL(c, api, api.Broadcast(PB, PB, (size_t)rows * w, KGP_NCCL_DOUBLE, owner, c->nccl_comm, sp));
    KGP_CUDA(c, cudaEventRecord(ev_panel[b], sp));                  // the updates read the PACKED panel: they need not wait for the unpack
    if (me != owner)
      KGP_CUDA(c, cudaMemcpy2DAsync(M + k0 * ld + k0, (size_t)ld * 8, PB, (size_t)w * 8, (size_t)w * 8, (size_t)rows, cudaMemcpyDeviceToDevice, sp));
    const long t0 = k0 + w;                                         // first trailing row
    const double* Pt = PB + (long)w * w;                            // panel rows from t0 on (row pitch w)
    const int ntm = (int)((np - t0) / 128);
    auto launch = [&](const std::vector<int>& blocks) {
      if (blocks.empty()) return;
      // blocks of full width NB go into one batched launch; a narrower last block gets its own
      GemmArgs g = mkgemm(Pt, w, Pt, w, M + t0 * ld + t0, ld, ntm, NB / 128, 0, w, GEMM_FULL, KTRI_NONE, -1.0, 1.0);
      for (int J : blocks) {
        const long j0 = (long)J * NB; const int wj = (int)((np - j0 < NB) ? np - j0 : NB);
        if (wj == NB && g.nbatch < GEMM_MAX_BATCH) g.bat_roff[g.nbatch++] = (int)(j0 - t0);
        else {
          GemmArgs h = mkgemm(Pt + (j0 - t0) * w, w, Pt + (j0 - t0) * w, w, M + j0 * ld + j0, ld, (int)((np - j0) / 128), wj / 128,
                              0, w, GEMM_FULL, KTRI_NONE, -1.0, 1.0);
          c->gemm_class = 1; gemm(c, h, tile_mode);
        }
      }
      if (g.nbatch > 0) { c->gemm_class = 1; gemm(c, g, tile_mode); }
    };
    std::vector<int> first, second, rest;
    for (int J = k + 1; J < nblk; ++J)
      if (dist_col_owner(J, P) == me) (J == k + 1 ? first : (la2 && J == k + 2) ? second : rest).push_back(J);
    if (t0 < np && la2) {                                           // first(k) on the panel stream, after second(k-1)
      KGP_CUDA(c, cudaStreamWaitEvent(sp, ev_second, 0));
      c->stream = sp; launch(first); c->stream = sm;
    }
    // ---- main stream: update the block columns this rank owns with the packed panel
    KGP_CUDA(c, cudaStreamWaitEvent(sm, ev_panel[b], 0));
    if (t0 < np) {
      if (la2) {
        launch(second);
        KGP_CUDA(c, cudaEventRecord(ev_second, sm));
      } else {
        launch(first);
        if (!first.empty()) KGP_CUDA(c, cudaEventRecord(ev_la, sm));   // look-ahead: panel k+1 may be factored now
      }
      launch(rest);
    }
    KGP_CUDA(c, cudaEventRecord(ev_free[b], sm));
  }
  KGP_CUDA(c, cudaEventRecord(c->dev_ev[6], sp));                   // every panel is unpacked into M behind this point
  KGP_CUDA(c, cudaStreamWaitEvent(sm, c->dev_ev[6], 0));
  return KGP_OK;
}

// Z = L^-T on this rank's rows [a, b) (right-looking), then every rank's rows are broadcast to all
static int zsweep_dist(kgp_ctx* c) {
  NcclApi& api = nccl_api(nullptr);
  const double* M = c->M.d(); double* W = c->W2.d(); const long ld = c->np, np = c->np;
  const int P = c->nranks, me = c->rank;
  const long a = dist_zrow_begin(me, P, np, NB), b = dist_zrow_begin(me + 1, P, np, NB);
  cudaStream_t st = c->stream;
  // Pipelined all-gather by peer copies (default for P > 2; KGP_DIST_ZPUSH=0/1 forces): the sweep is right-looking, so after the
  // triangular solve of block column k0 the columns < k0 + w of this rank's rows are final.  Finished column ranges are copied
  // straight into every peer's W2 (cudaMemcpy2DAsync on one copy stream per peer: copy engines over NVLink, no SMs, no packing)
  // while the sweep goes on -- eight block columns at a time, the last ones in smaller pieces so that the exposed tail is short.
  // Rows >= c1 of a range [c0, c1) are zero there and never read by Z Z^T (its K range starts at the tile's own row).
  const bool zpush = c->w2_push_ok && (getenv("KGP_DIST_ZPUSH") ? atoi(getenv("KGP_DIST_ZPUSH")) != 0 : (P > 2));
  if (b > a) {
    KGP_CUDA(c, cudaMemsetAsync(W + a * ld, 0, (size_t)(b - a) * ld * 8, st));
    long c0 = (a / NB) * NB; int since = 0;
    for (long k0 = (a / NB) * NB; k0 < np; k0 += NB) {
      const int w = (int)((np - k0 < NB) ? np - k0 : NB);
      const long hi = (b < k0 + w) ? b : k0 + w, nact = hi - a;
      if (k0 >= a && k0 < b) {
        set_identity_kernel<<<(unsigned)(((long)w * w + 255) / 256), 256, 0, st>>>(W + k0 * ld + k0, ld, w);
        c->stats.kernel_launches++;
      }
      trsm_block(c, W + a * ld + k0, ld, nact, M + k0 * ld + k0, ld, w);
      if (zpush) {
        const long rem = (np - k0 - w + NB - 1) / NB;             // block columns still to come
        ++since;
        if (rem < 2 || (rem < 8 && since >= 2) || since >= 8) {
          const long c1 = k0 + w;
          KGP_CUDA(c, cudaEventRecord(c->zcol_ev, st));
          for (int p = 0; p < P; ++p) {
            if (p == me) continue;
            KGP_CUDA(c, cudaStreamWaitEvent(c->zstream[p], c->zcol_ev, 0));
            KGP_CUDA(c, cudaMemcpy2DAsync((double*)c->w2_peer[p] + a * ld + c0, (size_t)ld * 8, W + a * ld + c0, (size_t)ld * 8,
                                          (size_t)(c1 - c0) * 8, (size_t)nact, cudaMemcpyDefault, c->zstream[p]));
          }
          c0 = c1; since = 0;
        }
      }
      const long rest = np - k0 - w;
      c->gemm_class = 2;
      if (rest > 0)
        gemm(c, mkgemm(W + a * ld + k0, ld, M + (k0 + w) * ld + k0, ld, W + a * ld + k0 + w, ld, (int)(nact / 128), (int)(rest / 128),
                       0, w, GEMM_FULL, KTRI_NONE, -1.0, 1.0));
    }
  }
  cudaEventRecord(c->ev[12], st);
  if (zpush) {
    // this rank's copies are complete -> tiny all-reduce: when it returns on a rank, every peer's copies into that rank's W2 have
    // completed as well (a rank joins only after its own copy streams have drained)
    for (int p = 0; p < P; ++p) {
      if (p == me) continue;
      KGP_CUDA(c, cudaEventRecord(c->zpush_ev[p], c->zstream[p]));
      KGP_CUDA(c, cudaStreamWaitEvent(st, c->zpush_ev[p], 0));
    }
    KGP_NCCL(c, api, api.AllReduce(c->zbar, c->zbar, 1, KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, st));
    c->zpipe_active = false;
    cudaEventRecord(c->ev[13], st);
    return KGP_OK;
  }
  // all-gather of Z as one broadcast per rank, in rank order, on the (high-priority) panel stream: rank r's rows are the
  // top rows with the long K ranges, i.e. the tiles of Z Z^T that become computable first.  One event per rank lets
  // zzt_reduce_dist start its tiles while the later (larger) broadcasts are still in flight.
  // Measured (n = 40 000): 2 GPUs 1035 -> 1024 ms per evaluation; 8 GPUs 344.7 -> 348.2 ms -- P sequential broadcasts lose the
  // concurrency of the grouped form (eight roots sending at once) and the one-CTA-per-tile stages run ~5 % below the
  // persistent launch.  Default: pipelined at P = 2 only (KGP_DIST_ZPIPE=0/1 forces it).
  const bool zpipe = getenv("KGP_DIST_ZPIPE") ? atoi(getenv("KGP_DIST_ZPIPE")) != 0 : (P == 2);
  cudaStream_t sc = zpipe ? c->stream_p : st;
  if (zpipe) KGP_CUDA(c, cudaStreamWaitEvent(sc, c->ev[12], 0));
  if (!zpipe) KGP_NCCL(c, api, api.GroupStart());
  for (int r = 0; r < P; ++r) {
    const long ra = dist_zrow_begin(r, P, np, NB), rb = dist_zrow_begin(r + 1, P, np, NB);
    if (rb > ra) KGP_NCCL(c, api, api.Broadcast(W + ra * ld, W + ra * ld, (size_t)(rb - ra) * ld, KGP_NCCL_DOUBLE, r, c->nccl_comm, sc));
    if (zpipe) KGP_CUDA(c, cudaEventRecord(c->zev[r], sc));
  }
  if (!zpipe) KGP_NCCL(c, api, api.GroupEnd());
  c->zpipe_active = zpipe;
  cudaEventRecord(c->ev[13], st);
  return KGP_OK;
}

template <int DP>
static void launch_reduce_tiles(kgp_ctx* c, const GemmArgs& g, long ntiles, long n, double sf2, double inv_sn2) {
  static bool attr = false;
  if (!attr) { cudaFuncSetAttribute(reduce_grad_tile_kernel<DP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)reduce_tile_smem<DP>()); attr = true; }
  const long grid = std::min<long>(ntiles, 148L * 8);
  reduce_grad_tile_kernel<DP><<<(unsigned)grid, 128, reduce_tile_smem<DP>(), c->stream>>>(g, ntiles, c->W2.d(), c->np, n, c->xs.d(),
                                                                                          c->alpha.d(), sf2, inv_sn2, c->rows.d());
}
// B^-1 = Z Z^T on this rank's share of the lower tiles + the gradient pass over the same tiles + all-reduce of the row sums
static int zzt_reduce_dist(kgp_ctx* c, long n, double sf2, double inv_sn2) {
  double* W = c->W2.d(); const long ld = c->np; const int nt = (int)(c->np / 128); const int Dp = c->Dp;
  GemmArgs g = mkgemm(W, ld, W, ld, W, ld, nt, nt, 0, (int)c->np, GEMM_LOWER, KTRI_FROM_M0, 1.0, 0.0);
  g.Cdiag = c->Dws.d(); g.nb = NB; g.own_mod = c->nranks; g.own_rem = c->rank;
  c->gemm_class = 4;
  if (c->nranks > 1 && c->zpipe_active) {
    // staged: the global tile enumeration runs band by band (8 tile rows) from the top; a prefix of it needs only the Z rows of
    // the first ranks.  Stage r covers the bands that lie completely above the end of rank r's rows and starts when that
    // rank's broadcast has arrived; one CTA per tile, so that the NCCL kernels of the later broadcasts (high-priority
    // stream) get SMs as tiles retire.  Ownership stays the global idx = t P + rank rule, so the gradient pass below still
    // enumerates exactly the tiles this rank produced.
    const int P = c->nranks, me = c->rank;
    const double fl = count_gemm(c, g);                           // flops of this rank's share, counted once for all stages
    const bool prof = c->profiling != 0;
    if (prof) {
      if (c->gev_used + 2 > c->gev.size()) { size_t old = c->gev.size(); c->gev.resize(old + 256); for (size_t i = old; i < c->gev.size(); ++i) cudaEventCreate(&c->gev[i]); }
      cudaEventRecord(c->gev[c->gev_used], c->stream);
    }
    long tile_lo = 0;                                              // global tile index where the next stage begins
    for (int r = 0; r < P; ++r) {
      const long rb = dist_zrow_begin(r + 1, P, c->np, NB);
      long tile_hi;
      if (r == P - 1) tile_hi = (long)nt * (nt + 1) / 2;
      else { const long rows_t = rb / 128, bands = rows_t / GEMM_BAND, tr = bands * GEMM_BAND; tile_hi = tr * (tr + 1) / 2; }
      KGP_CUDA(c, cudaStreamWaitEvent(c->stream, c->zev[r], 0));
      if (tile_hi > tile_lo) {
        // owned global indices in [tile_lo, tile_hi): first one G0 = smallest >= tile_lo with G0 % P == me
        long G0 = tile_lo + ((me - tile_lo % P) + P) % P;
        if (G0 < tile_hi) {
          GemmArgs h = g; h.tile_off = G0 - me; h.tile_cnt = (tile_hi - G0 + P - 1) / P;
          gemm_launch(h, c->stream, true);
        }
        tile_lo = tile_hi;
      }
    }
    if (prof) {                                                    // (one event pair for the staged launch: it includes waiting for rows)
      cudaEventRecord(c->gev[c->gev_used + 1], c->stream);
      c->gev_class.push_back(c->gemm_class); c->gev_flops.push_back(fl); c->gev_used += 2;
    }
  } else gemm(c, g);
  cudaEventRecord(c->ev[5], c->stream);
  // the same enumeration for the reduction (band table as gemm_launch builds it)
  long base = 0; int nb_ = 0;
  for (int r0 = 0; r0 < g.ntm && nb_ < GEMM_MAX_BANDS; r0 += 8) {
    g.band_base[nb_++] = (int)base;
    for (int r = r0; r < r0 + 8 && r < g.ntm; ++r) base += (r + 1 < g.ntn) ? r + 1 : g.ntn;
  }
  g.nbands = nb_;
  long ntiles = gemm_num_tiles(g);
  ntiles = (ntiles > g.own_rem) ? (ntiles - g.own_rem + g.own_mod - 1) / g.own_mod : 0;
  const size_t rbytes = (size_t)c->np * (Dp + 3) * 8;
  KGP_CUDA(c, cudaMemsetAsync(c->rows.p, 0, rbytes, c->stream));
  if (ntiles > 0) {
    switch (Dp) {
      case 4: launch_reduce_tiles<4>(c, g, ntiles, n, sf2, inv_sn2); break;
      case 8: launch_reduce_tiles<8>(c, g, ntiles, n, sf2, inv_sn2); break;
      case 16: launch_reduce_tiles<16>(c, g, ntiles, n, sf2, inv_sn2); break;
      default: launch_reduce_tiles<32>(c, g, ntiles, n, sf2, inv_sn2); break;
    }
    c->stats.kernel_launches++;
  }
  if (c->nranks > 1) {                                   // (NCCL is bound only here: single-GPU use never loads libnccl)
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(c->rows.p, c->rows.p, (size_t)c->np * (Dp + 3), KGP_NCCL_DOUBLE, KGP_NCCL_SUM, c->nccl_comm, c->stream));
  }
  return KGP_OK;
}

static int pick_dp(int D) { return D <= 4 ? 4 : D <= 8 ? 8 : D <= 16 ? 16 : 32; }

template <int DP>
static void launch_reduce(kgp_ctx* c, long n, double sf2, double inv_sn2, int nsplit, int per) {
  dim3 grid((unsigned)(c->np / 128), (unsigned)nsplit);
  reduce_grad_kernel<DP><<<grid, 128, 0, c->stream>>>(c->W2.d(), c->np, n, c->np, c->xs.d(), c->alpha.d(),
                                                     sf2, inv_sn2, per, c->parts.d());
}

static int ensure_exact_buffers(kgp_ctx* c, long n, int D) {
  const long np = ((n + 127) / 128) * 128;
  const int Dp = pick_dp(D);
  c->np = np; c->Dp = Dp;
  KGP_CUDA(c, c->M.ensure((size_t)np * np * 8));
  if ((size_t)np * np * 8 > c->W2.bytes && c->W2.p && c->W2.p == c->w2_mapped) {
    // peers hold CUDA-IPC mappings of this allocation (w2_peer_sync): an exporter must not free it under them.  Outside the
    // collective re-mapping (e.g. the KISS-GP preconditioner growing the matrix on a sharded context) it is set aside instead.
    c->w2_retired.push_back(c->W2.p); c->W2.p = nullptr; c->W2.bytes = 0;
  }
  KGP_CUDA(c, c->W2.ensure((size_t)np * np * 8));
  KGP_CUDA(c, c->Dws.ensure((size_t)np * NB * 8));
  KGP_CUDA(c, c->xs.ensure((size_t)np * Dp * 8));
  KGP_CUDA(c, c->s.ensure((size_t)np * 8));
  KGP_CUDA(c, c->mu.ensure(64 * 8));
  KGP_CUDA(c, c->inv_ell.ensure(64 * 8));
  KGP_CUDA(c, c->fac.ensure(64 * 8));
  KGP_CUDA(c, c->hypm.ensure(128 * 8));
  KGP_CUDA(c, c->ym.ensure((size_t)np * 8));
  KGP_CUDA(c, c->alpha.ensure((size_t)np * 8));
  KGP_CUDA(c, c->tvec.ensure((size_t)np * 8));
  KGP_CUDA(c, c->scal.ensure(256 * 8));
  KGP_CUDA(c, c->rows.ensure((size_t)np * (Dp + 3) * 8));
  KGP_CUDA(c, c->dX.ensure((size_t)n * D * 8));
  if (!c->d_info) KGP_CUDA(c, cudaMalloc(&c->d_info, sizeof(int)));
  return KGP_OK;
}

// covSM (exact path): host-side unpacking of hyp = [log w; log m (D x Q); log sqrt v (D x Q)] (covSM.m:127-129) into the
// kernel tables Vq = 4 pi^2 v, Mq = 2 pi m ([q][Dp], zero padded)
static int sm_prepare(kgp_ctx* c, int D, int Dp, SmHyp& h, std::vector<double>& Vq, std::vector<double>& Mq) {
  Model& m = c->model; const int Q = m.sm_Q;
  if (Q < 1 || Q > SM_MAXQ) return c->fail(KGP_BAD_ARG, "covSM needs 1 <= Q <= %d (got %d)", SM_MAXQ, Q);
  if (D > 16) return c->fail(KGP_BAD_ARG, "exact covSM supports input_dim <= 16 (got %d)", D);
  if ((int)m.hyp_cov.size() != (1 + 2 * D) * Q)
    return c->fail(KGP_BAD_ARG, "Wrong number of hyperparameters (cov): got %d, need %d", (int)m.hyp_cov.size(), (1 + 2 * D) * Q);
  memset(&h, 0, sizeof h); h.Q = Q;
  Vq.assign((size_t)Q * Dp, 0.0); Mq.assign((size_t)Q * Dp, 0.0);
  for (int q = 0; q < Q; ++q) {
    h.w[q] = exp(m.hyp_cov[q]);
    for (int d = 0; d < D; ++d) {
      Mq[(size_t)q * Dp + d] = 2.0 * M_PI * exp(m.hyp_cov[Q + q * D + d]);
      Vq[(size_t)q * Dp + d] = 4.0 * M_PI * M_PI * exp(2.0 * m.hyp_cov[Q + D * Q + q * D + d]);
    }
  }
  KGP_CUDA(c, c->smV.ensure(Vq.size() * 8)); KGP_CUDA(c, c->smM.ensure(Mq.size() * 8));
  KGP_CUDA(c, cudaMemcpyAsync(c->smV.p, Vq.data(), Vq.size() * 8, cudaMemcpyHostToDevice, c->stream));
  KGP_CUDA(c, cudaMemcpyAsync(c->smM.p, Mq.data(), Mq.size() * 8, cudaMemcpyHostToDevice, c->stream));
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));               // Vq / Mq are locals of the caller
  return KGP_OK;
}
template <int DP>
static void launch_sm_build(kgp_ctx* c, const double* xr, long nR, long rowsp, const double* xc, long nC, long colsp, const SmHyp& h,
                            double scale, int diag_add, int lower_only, double* out, long ld) {
  dim3 grid((unsigned)(colsp / 32), (unsigned)(rowsp / 128));
  sm_build_kernel<DP><<<grid, 128, 0, c->stream>>>(xr, nR, xc, nC, h, c->smV.d(), c->smM.d(), scale, diag_add, lower_only, out, ld);
}
static void sm_build(kgp_ctx* c, int Dp, const double* xr, long nR, long rowsp, const double* xc, long nC, long colsp, const SmHyp& h,
                     double scale, int diag_add, int lower_only, double* out, long ld) {
  switch (Dp) {
    case 4: launch_sm_build<4>(c, xr, nR, rowsp, xc, nC, colsp, h, scale, diag_add, lower_only, out, ld); break;
    case 8: launch_sm_build<8>(c, xr, nR, rowsp, xc, nC, colsp, h, scale, diag_add, lower_only, out, ld); break;
    default: launch_sm_build<16>(c, xr, nR, rowsp, xc, nC, colsp, h, scale, diag_add, lower_only, out, ld); break;
  }
  c->stats.kernel_launches++;
}
template <int DP>
static void launch_sm_reduce(kgp_ctx* c, long n, int D, const SmHyp& h, double inv_sn2, int nsplit, int per, double* acc) {
  dim3 grid((unsigned)(c->np / 128), (unsigned)nsplit, (unsigned)h.Q);
  sm_reduce_kernel<DP><<<grid, 128, 0, c->stream>>>(c->W2.d(), c->np, n, c->xs.d(), c->alpha.d(), h, c->smV.d(), c->smM.d(), inv_sn2, per, D, acc, c->dX.d());
}

static float ev_ms(cudaEvent_t a, cudaEvent_t b) { float ms = 0; cudaEventElapsedTime(&ms, a, b); return ms; }

// ---------------------------------------------------------------------------
int exact_eval(kgp_ctx* c, int slot, int want, double* nlZ_out, double* dnlZ_out) {
  Slot& S = c->slots[slot]; Model& m = c->model;
  const long n = S.n; const int D = S.D;
  if (D > 32) return c->fail(KGP_BAD_ARG, "exact path supports input_dim <= 32 (got %d)", D);
  const bool ard = m.cov_kind == KGP_COV_SEARD, sm = m.cov_kind == KGP_COV_SM;
  const int nl = ard ? D : 1;
  if (!sm && (int)m.hyp_cov.size() != nl + 1) return c->fail(KGP_BAD_ARG, "Wrong number of hyperparameters (cov): got %d, need %d", (int)m.hyp_cov.size(), nl + 1);
  if (sm && c->nranks > 1) return c->fail(KGP_BAD_ARG, "exact covSM runs on one GPU (the sharded exact path covers covSEiso / covSEard)");
  const int nmean = ((m.mean_mask & 1) ? D : 0) + ((m.mean_mask & 2) ? 1 : 0);
  if ((int)m.hyp_mean.size() != nmean) return c->fail(KGP_BAD_ARG, "Number of mean function hyperparameters disagree with mean function");
  int rc;
  // (the peer mappings are set up only where the peer-copy gather is in use: more than two ranks, or forced by KGP_DIST_ZPUSH=1)
  if (c->nranks > 1 && c->nranks <= 8 && !sm &&
      (getenv("KGP_DIST_ZPUSH") ? atoi(getenv("KGP_DIST_ZPUSH")) != 0 : c->nranks > 2)) {
    const long npp = ((n + 127) / 128) * 128;
    rc = w2_peer_sync(c, (size_t)npp * npp * 8); if (rc) return rc;
  }
  rc = ensure_exact_buffers(c, n, D); if (rc) return rc;
  const long np = c->np; const int Dp = c->Dp;
  cudaStream_t st = c->stream;
  memset(&c->stats, 0, sizeof c->stats); c->stats.n = n; c->stats.np = np;
  c->st_level = -1;
  SmHyp smh; std::vector<double> smVq, smMq;
  if (sm) { rc = sm_prepare(c, D, Dp, smh, smVq, smMq); if (rc) return rc; }

  const double sn2 = exp(2.0 * m.hyp_lik), sf2 = sm ? 1.0 : exp(2.0 * m.hyp_cov[nl]);
  double h_inv_ell[64], h_fac[64];
  for (int d = 0; d < D; ++d) {
    double ell = sm ? 1.0 : exp(m.hyp_cov[ard ? d : 0]);
    h_inv_ell[d] = 1.0 / ell;
    h_fac[d] = (ard && m.opt.ard_reference_form) ? ell / exp(2.0 * m.hyp_cov[0]) : 1.0 / ell;
  }
  KGP_CUDA(c, cudaMemcpyAsync(c->inv_ell.p, h_inv_ell, D * 8, cudaMemcpyHostToDevice, st));
  KGP_CUDA(c, cudaMemcpyAsync(c->fac.p, h_fac, D * 8, cudaMemcpyHostToDevice, st));
  if (nmean) KGP_CUDA(c, cudaMemcpyAsync(c->hypm.p, m.hyp_mean.data(), nmean * 8, cudaMemcpyHostToDevice, st));
  int big = 1 << 30;
  KGP_CUDA(c, cudaMemcpyAsync(c->d_info, &big, sizeof(int), cudaMemcpyHostToDevice, st));

  cudaEventRecord(c->ev[0], st);
  col_mean_kernel<<<D, 1024, 0, st>>>(S.X.d(), n, D, c->mu.d());
  scale_x_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(S.X.d(), n, D, c->mu.d(), c->inv_ell.d(), c->xs.d(), c->s.d(), np, Dp);
  resid_mean_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(S.X.d(), S.has_y ? S.y.d() : nullptr, n, D, m.mean_mask, c->hypm.d(), c->ym.d(), np);
  if (sm) sm_build(c, Dp, c->xs.d(), n, np, c->xs.d(), n, np, smh, 1.0 / sn2, 1, 1, c->M.d(), np);
  else {
    dim3 grid((unsigned)(np / 128), (unsigned)(np / 128));
    kbuild_kernel<<<grid, KB_THREADS, kbuild_smem(Dp), st>>>(c->xs.d(), c->s.d(), n, c->xs.d(), c->s.d(), n, Dp,
                                                            sf2 / sn2, 1, 1, c->M.d(), np);
  }
  c->stats.kernel_launches += 4;
  cudaEventRecord(c->ev[1], st);
  const bool dist = c->nranks > 1;
  // look-ahead (panel k+1 factored on the high-priority stream under trailing update k) is the default; it is switched
  // off while per-launch GEMM profiling is on, because overlapped launches blur the event pairs around each GEMM
  static const bool lookahead = !(getenv("KGP_LOOKAHEAD") && atoi(getenv("KGP_LOOKAHEAD")) == 0);
  if (dist) {
    rc = potrf_dist(c); if (rc) return rc;
    // the failing-pivot flag is written by the owner of the failing panel only: make it global (min over the ranks) so
    // that every rank returns the same status -- otherwise the ranks' optimisers diverge and the next collective hangs
    NcclApi& api = nccl_api(nullptr);
    KGP_NCCL(c, api, api.AllReduce(c->d_info, c->d_info, 1, KGP_NCCL_INT32, KGP_NCCL_MIN, c->nccl_comm, st));
  }
  else if (lookahead && !c->profiling && c->np >= 4 * NB) { rc = potrf_lookahead(c); if (rc) return rc; }
  else potrf(c);
  logdet_kernel<<<1, 1024, 0, st>>>(c->M.d(), np, n, c->scal.d() + 100);
  c->stats.kernel_launches++;
  cudaEventRecord(c->ev[2], st);
  if (dist) { rc = zsweep_dist(c); if (rc) return rc; } else zsweep(c);
  cudaEventRecord(c->ev[3], st);
  {  // alpha = B^{-1} ym / sn2 = Z (Z^T ym) / sn2
    const int chunk = np <= 8192 ? 64 : 512;          // small problems: shorter serial loops, more CTAs (latency-bound)
    const int nch = (int)((np + chunk - 1) / chunk);
    KGP_CUDA(c, c->parts.ensure((size_t)nch * np * 8));
    dim3 grid((unsigned)(np / 128), (unsigned)nch);
    gemv_t_upper_kernel<<<grid, 128, 0, st>>>(c->W2.d(), np, np, c->ym.d(), c->parts.d(), chunk);
    sum_parts_kernel<<<(unsigned)((np + 255) / 256), 256, 0, st>>>(c->parts.d(), np, nch, 1.0, c->tvec.d());
    gemv_n_upper_kernel<<<(unsigned)((np + 7) / 8), 256, 0, st>>>(c->W2.d(), np, np, c->tvec.d(), 1.0 / sn2, c->alpha.d());
    dot_kernel<<<1, 1024, 0, st>>>(c->ym.d(), c->alpha.d(), n, c->scal.d() + 101);
    c->stats.kernel_launches += 4;
  }
  cudaEventRecord(c->ev[4], st);
  // symmetric gradient pass over the lower tiles of B^-1 (reads 4n^2 bytes instead of 8n^2, no mirror pass) is the default
  static const bool tile_reduce = !(getenv("KGP_TILE_REDUCE") && atoi(getenv("KGP_TILE_REDUCE")) == 0);
  // small problems (a handful of 128-tiles) do not fill the GPU with one CTA per tile: the row-split full-square pass is faster there
  const bool use_tiles = dist || (tile_reduce && np > 2048);
  if (want >= 1 && sm) {
    // covSM: full symmetric B^-1 (Z Z^T + mirror), then one pass per mixture component (exact_sm.cuh)
    zzt(c);
    cudaEventRecord(c->ev[5], st);
    const int ntr = (int)(np / 128), nt32 = (int)(np / 32);
    int nsplit = (1184 + ntr * smh.Q - 1) / (ntr * smh.Q); if (nsplit < 1) nsplit = 1; if (nsplit > nt32) nsplit = nt32;
    const int per = (nt32 + nsplit - 1) / nsplit; nsplit = (nt32 + per - 1) / per;
    const size_t accn = 1 + (size_t)smh.Q * (1 + 2 * Dp);
    KGP_CUDA(c, c->smAcc.ensure(accn * 8));
    KGP_CUDA(c, cudaMemsetAsync(c->smAcc.p, 0, accn * 8, st));
    KGP_CUDA(c, cudaMemsetAsync(c->dX.p, 0, (size_t)n * D * 8, st));
    switch (Dp) {
      case 4: launch_sm_reduce<4>(c, n, D, smh, 1.0 / sn2, nsplit, per, c->smAcc.d()); break;
      case 8: launch_sm_reduce<8>(c, n, D, smh, 1.0 / sn2, nsplit, per, c->smAcc.d()); break;
      default: launch_sm_reduce<16>(c, n, D, smh, 1.0 / sn2, nsplit, per, c->smAcc.d()); break;
    }
    dot_kernel<<<1, 1024, 0, st>>>(c->alpha.d(), c->alpha.d(), n, c->scal.d() + 3);
    if (nmean) mean_grad_kernel<<<D + 1, 1024, 0, st>>>(S.X.d(), c->alpha.d(), n, D, c->scal.d() + 110);
    c->stats.kernel_launches += 2 + (nmean ? 1 : 0);
    cudaEventRecord(c->ev[6], st);
  } else if (want >= 1 && use_tiles) {
    rc = zzt_reduce_dist(c, n, sf2, 1.0 / sn2); if (rc) return rc;
    grad_scalars_kernel<<<4 + Dp, 1024, 0, st>>>(c->rows.d(), n, Dp, c->xs.d(), c->alpha.d(), c->scal.d());
    dx_from_rows_kernel<<<(unsigned)((n * D + 255) / 256), 256, 0, st>>>(c->rows.d(), n, D, Dp, c->fac.d(), c->dX.d());
    if (nmean) mean_grad_kernel<<<D + 1, 1024, 0, st>>>(S.X.d(), c->alpha.d(), n, D, c->scal.d() + 110);
    c->stats.kernel_launches += 2 + (nmean ? 1 : 0);
    cudaEventRecord(c->ev[6], st);
  } else if (want >= 1) {
    zzt(c);
    cudaEventRecord(c->ev[5], st);
    const int ntr = (int)(np / 128), nt32 = (int)(np / 32);
    int nsplit = (1184 + ntr - 1) / ntr; if (nsplit < 1) nsplit = 1; if (nsplit > nt32) nsplit = nt32;
    const int per = (nt32 + nsplit - 1) / nsplit; nsplit = (nt32 + per - 1) / per;
    KGP_CUDA(c, c->parts.ensure((size_t)nsplit * np * (Dp + 3) * 8));
    switch (Dp) {
      case 4: launch_reduce<4>(c, n, sf2, 1.0 / sn2, nsplit, per); break;
      case 8: launch_reduce<8>(c, n, sf2, 1.0 / sn2, nsplit, per); break;
      case 16: launch_reduce<16>(c, n, sf2, 1.0 / sn2, nsplit, per); break;
      default: launch_reduce<32>(c, n, sf2, 1.0 / sn2, nsplit, per); break;
    }
    const long tot = np * (Dp + 3);
    reduce_rows_kernel<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(c->parts.d(), np, nsplit, Dp + 3, c->rows.d());
    grad_scalars_kernel<<<4 + Dp, 1024, 0, st>>>(c->rows.d(), n, Dp, c->xs.d(), c->alpha.d(), c->scal.d());
    dx_from_rows_kernel<<<(unsigned)((n * D + 255) / 256), 256, 0, st>>>(c->rows.d(), n, D, Dp, c->fac.d(), c->dX.d());
    if (nmean) mean_grad_kernel<<<D + 1, 1024, 0, st>>>(S.X.d(), c->alpha.d(), n, D, c->scal.d() + 110);
    c->stats.kernel_launches += 4 + (nmean ? 1 : 0);
    cudaEventRecord(c->ev[6], st);
  }
  double h[256];
  KGP_CUDA(c, cudaMemcpyAsync(h, c->scal.p, 256 * 8, cudaMemcpyDeviceToHost, st));
  int info = 0;
  KGP_CUDA(c, cudaMemcpyAsync(&info, c->d_info, sizeof(int), cudaMemcpyDeviceToHost, st));
  cudaEventRecord(c->ev[7], st);
  KGP_CUDA(c, cudaStreamSynchronize(st));
  KGP_CUDA(c, cudaGetLastError());
  if (c->profiling) gemm_prof_collect(c);
  c->stats.ms_kbuild = ev_ms(c->ev[0], c->ev[1]);
  c->stats.ms_potrf = ev_ms(c->ev[1], c->ev[2]);
  c->stats.ms_inverse = ev_ms(c->ev[2], c->ev[3]) + (want >= 1 ? ev_ms(c->ev[4], c->ev[5]) : 0.f);
  c->stats.ms_solve = ev_ms(c->ev[3], c->ev[4]);
  c->stats.ms_reduce = want >= 1 ? ev_ms(c->ev[5], c->ev[6]) : 0.f;
  c->stats.ms_total = ev_ms(c->ev[0], c->ev[7]);
  if (info != (1 << 30)) {
    c->stats.potrf_info = info;
    return c->fail(KGP_NOT_PD, "chol: input matrix must be positive definite (pivot %d)", info);
  }
  const double ldB2 = h[100], yta = h[101];
  const double nlZ = yta / 2.0 + ldB2 + (double)n * log(2.0 * M_PI * sn2) / 2.0;     // infGaussLik.m:29
  if (nlZ_out) *nlZ_out = nlZ;
  c->st_nlZ = nlZ; c->st_dnlZ.clear();
  if (want >= 1 && sm) {
    const int Q = smh.Q; const size_t accn = 1 + (size_t)Q * (1 + 2 * Dp);
    std::vector<double> a(accn);
    KGP_CUDA(c, cudaMemcpy(a.data(), c->smAcc.p, accn * 8, cudaMemcpyDeviceToHost));
    std::vector<double> g((size_t)(1 + 2 * D) * Q);
    for (int q = 0; q < Q; ++q) {                                          // dnlZ.cov = dKdhyp(R) / 2 (covSM.m:150-153, apx.m:173)
      const double* aq = a.data() + 1 + (size_t)q * (1 + 2 * Dp);
      g[q] = 0.5 * smh.w[q] * aq[0];
      for (int d = 0; d < D; ++d) {
        g[Q + q * D + d] = -0.5 * smh.w[q] * smMq[(size_t)q * Dp + d] * aq[1 + d];
        g[Q + D * Q + q * D + d] = -0.5 * smh.w[q] * smVq[(size_t)q * Dp + d] * aq[1 + Dp + d];
      }
    }
    g.push_back(-sn2 * h[3] - a[0] / sn2 + (double)n);                     // infGaussLik.m:32
    if (m.mean_mask & 1) for (int d = 0; d < D; ++d) g.push_back(h[110 + d]);
    if (m.mean_mask & 2) g.push_back(h[110 + D]);
    c->st_dnlZ = g;
    if (dnlZ_out) memcpy(dnlZ_out, g.data(), g.size() * 8);
  } else if (want >= 1) {
    std::vector<double> g;
    if (ard) for (int d = 0; d < D; ++d) g.push_back(h[4 + d]);
    else g.push_back(0.5 * h[1]);
    g.push_back(h[0]);                                                     // d/d log sf
    g.push_back(-sn2 * h[3] - h[2] / sn2 + (double)n);                     // infGaussLik.m:32
    if (m.mean_mask & 1) for (int d = 0; d < D; ++d) g.push_back(h[110 + d]);
    if (m.mean_mask & 2) g.push_back(h[110 + D]);
    c->st_dnlZ = g;
    if (dnlZ_out) memcpy(dnlZ_out, g.data(), g.size() * 8);
  }
  c->st_slot = slot; c->st_slot_ver = S.version; c->st_model_ver = m.version;
  c->st_level = want >= 1 ? 2 : 0;
  return KGP_OK;
}

static bool state_valid(kgp_ctx* c, int slot, int level) {
  return c->st_level >= level && c->st_slot == slot && c->st_slot_ver == c->slots[slot].version &&
         c->st_model_ver == c->model.version;
}

int exact_dlik(kgp_ctx* c, int slot, double* out, bool out_is_dev) {
  if (!state_valid(c, slot, 2)) {
    int rc = exact_eval(c, slot, 2, nullptr, nullptr);
    if (rc) return rc;
  }
  Slot& S = c->slots[slot];
  KGP_CUDA(c, cudaMemcpyAsync(out, c->dX.p, (size_t)S.n * S.D * 8,
                              out_is_dev ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, c->stream));
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));
  return KGP_OK;
}

int exact_alpha(kgp_ctx* c, double* out) {
  if (c->st_level < 0) return c->fail(KGP_NO_DATA, "no posterior computed yet");
  KGP_CUDA(c, cudaMemcpy(out, c->alpha.p, (size_t)c->slots[c->st_slot].n * 8, cudaMemcpyDeviceToHost));
  return KGP_OK;
}

// gp.m:128-181 for infExact: batches of <=1024 test points
int exact_predict(kgp_ctx* c, const double* Xs, long ns, double* ymu, double* ys2, double* fmu, double* fs2) {
  const int slot = KGP_SLOT_TR;
  if (!state_valid(c, slot, 0)) {
    int rc = exact_eval(c, slot, 0, nullptr, nullptr);
    if (rc) return rc;
  }
  Slot& S = c->slots[slot]; Model& m = c->model;
  const long n = S.n, np = c->np; const int D = S.D, Dp = c->Dp;
  const bool ard = m.cov_kind == KGP_COV_SEARD, sm = m.cov_kind == KGP_COV_SM; const int nl = ard ? D : 1;
  SmHyp smh; std::vector<double> smVq, smMq;
  if (sm) { int rc = sm_prepare(c, D, Dp, smh, smVq, smMq); if (rc) return rc; }
  double sf2 = 0.0;                                               // prior variance k(x*, x*): sf^2, or sum_q w_q for covSM
  if (sm) for (int q = 0; q < smh.Q; ++q) sf2 += smh.w[q]; else sf2 = exp(2.0 * m.hyp_cov[nl]);
  const double sn2 = exp(2.0 * m.hyp_lik);
  const bool want_var = ys2 || fs2;
  cudaStream_t st = c->stream;
  const long BATCH = 1024;
  KGP_CUDA(c, c->pred_K.ensure((size_t)BATCH * np * 8));
  KGP_CUDA(c, c->pred_X.ensure((size_t)BATCH * D * 8));
  KGP_CUDA(c, c->pred_x.ensure((size_t)BATCH * Dp * 8));
  KGP_CUDA(c, c->pred_s.ensure((size_t)BATCH * 8));
  KGP_CUDA(c, c->pred_a.ensure((size_t)BATCH * 8));
  KGP_CUDA(c, c->pred_b.ensure((size_t)BATCH * 8));
  KGP_CUDA(c, c->pred_out.ensure((size_t)BATCH * 4 * 8));
  std::vector<double> h(BATCH * 4);
  for (long s0 = 0; s0 < ns; s0 += BATCH) {
    const long nsb = (ns - s0 < BATCH) ? ns - s0 : BATCH, nsp = ((nsb + 127) / 128) * 128;
    KGP_CUDA(c, cudaMemcpyAsync(c->pred_X.p, Xs + s0 * D, (size_t)nsb * D * 8, cudaMemcpyHostToDevice, st));
    scale_x_kernel<<<(unsigned)((nsp + 255) / 256), 256, 0, st>>>(c->pred_X.d(), nsb, D, c->mu.d(), c->inv_ell.d(),
                                                                  c->pred_x.d(), c->pred_s.d(), nsp, Dp);
    if (sm) sm_build(c, Dp, c->pred_x.d(), nsb, nsp, c->xs.d(), n, np, smh, 1.0, 0, 0, c->pred_K.d(), np);
    else {
      dim3 grid((unsigned)(np / 128), (unsigned)(nsp / 128));
      kbuild_kernel<<<grid, KB_THREADS, kbuild_smem(Dp), st>>>(c->pred_x.d(), c->pred_s.d(), nsb, c->xs.d(), c->s.d(), n,
                                                              Dp, sf2, 0, 0, c->pred_K.d(), np);
    }
    gemv_rows_kernel<<<(unsigned)((nsp + 7) / 8), 256, 0, st>>>(c->pred_K.d(), np, nsp, np, c->alpha.d(), c->pred_a.d());
    if (want_var) {
      double* X = c->pred_K.d(); const double* M = c->M.d();
      for (long k0 = 0; k0 < np; k0 += NB) {
        const int w = (int)((np - k0 < NB) ? np - k0 : NB);
        trsm_block(c, X + k0, np, nsp, M + k0 * np + k0, np, w);
        const long rest = np - k0 - w;
        c->gemm_class = 5;
        if (rest > 0)
          gemm(c, mkgemm(X + k0, np, M + (k0 + w) * np + k0, np, X + k0 + w, np, (int)(nsp / 128), (int)(rest / 128),
                         0, w, GEMM_FULL, KTRI_NONE, -1.0, 1.0));
      }
      rowsumsq_kernel<<<(unsigned)((nsp + 7) / 8), 256, 0, st>>>(X, np, nsp, np, c->pred_b.d());
      scale_matrix_kernel<<<(unsigned)((nsp + 255) / 256), 256, 0, st>>>(c->pred_b.d(), 1, nsp, 1, 1.0 / sn2);
    }
    double* o = c->pred_out.d();
    predict_finish_kernel<<<(unsigned)((nsb + 255) / 256), 256, 0, st>>>(c->pred_X.d(), nsb, D, m.mean_mask, c->hypm.d(),
        c->pred_a.d(), c->pred_b.d(), sf2, sn2, want_var ? 1 : 0, o, o + BATCH, o + 2 * BATCH, o + 3 * BATCH);
    KGP_CUDA(c, cudaMemcpyAsync(h.data(), o, (size_t)BATCH * 4 * 8, cudaMemcpyDeviceToHost, st));
    KGP_CUDA(c, cudaStreamSynchronize(st));
    KGP_CUDA(c, cudaGetLastError());
    if (ymu) memcpy(ymu + s0, h.data(), nsb * 8);
    if (ys2) memcpy(ys2 + s0, h.data() + BATCH, nsb * 8);
    if (fmu) memcpy(fmu + s0, h.data() + 2 * BATCH, nsb * 8);
    if (fs2) memcpy(fs2 + s0, h.data() + 3 * BATCH, nsb * 8);
  }
  return KGP_OK;
}

// ===== allocation-free predictive mean (kgp_predict_mean_noalloc_dev) ==================================================
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the predictive-mean step of
// exact_predict -- there the cross-covariances and partial results go through cudaMalloc'd scratch (pred_K, pred_x, pred_s,
// pred_a ...) -- with its scratch in ONE file-scope device array of fixed size:
//   parts[chunk][i] = sf2 sum_{j in chunk} exp(-|xs*_i - xs_j|^2 / 2) alpha_j   (chunks of 512 training points)
//   ymu[i] = m(x*_i) + sum_chunk parts[chunk][i]                                 (fixed chunk order: deterministic)
// Every GPU that loads the library holds the array (2 MiB) until exit.  All calls share it and it keeps the last call's
// contents; every element a call reads it has written first, so nothing needs clearing -- but two calls that overlap on
// one GPU (two contexts / two host threads) race on it.
constexpr long NOALLOC_SCRATCH_DOUBLES = 262144;                 // 2 MiB: ns * ceil(n / 512) doubles at most
static __device__ __align__(256) double g_noalloc_scratch[NOALLOC_SCRATCH_DOUBLES];
constexpr int NOALLOC_MAX_DEV = 64;
static double* g_noalloc_ptr[NOALLOC_MAX_DEV] = {nullptr};       // DEVICE address of the array per GPU (the symbol itself is a host-side stand-in)
constexpr int PM_CHUNK = 512, PM_SUB = 64, PM_THREADS = 128;

template <int DP>
__global__ void __launch_bounds__(PM_THREADS) predict_mean_parts_kernel(const double* __restrict__ Xs, long ns, int D,
    const double* __restrict__ mu, const double* __restrict__ inv_ell, const double* __restrict__ xs, const double* __restrict__ alpha,
    long n, double sf2, double* __restrict__ parts) {
  __shared__ double sx[PM_SUB][DP + 1];
  __shared__ double sa[PM_SUB];
  const int tid = threadIdx.x;
  const long i = (long)blockIdx.x * PM_THREADS + tid;
  double xi[DP];
#pragma unroll
  for (int d = 0; d < DP; ++d) xi[d] = (i < ns && d < D) ? (Xs[i * D + d] - mu[d]) * inv_ell[d] : 0.0;   // as scale_x_kernel
  const long j0 = (long)blockIdx.y * PM_CHUNK, j1 = (j0 + PM_CHUNK < n) ? j0 + PM_CHUNK : n;
  double acc = 0.0;
  for (long jb = j0; jb < j1; jb += PM_SUB) {
    __syncthreads();
    for (int q = tid; q < PM_SUB * DP; q += PM_THREADS) {
      const int r = q / DP, d = q - r * DP;
      sx[r][d] = (jb + r < j1) ? xs[(jb + r) * DP + d] : 0.0;
    }
    if (tid < PM_SUB) sa[tid] = (jb + tid < j1) ? alpha[jb + tid] : 0.0;
    __syncthreads();
    const int cnt = (j1 - jb < PM_SUB) ? (int)(j1 - jb) : PM_SUB;
    for (int r = 0; r < cnt; ++r) {
      double d2 = 0.0;
#pragma unroll
      for (int d = 0; d < DP; ++d) { const double t = xi[d] - sx[r][d]; d2 = fma(t, t, d2); }
      acc = fma(exp(-0.5 * d2), sa[r], acc);
    }
  }
  if (i < ns) parts[(long)blockIdx.y * ns + i] = sf2 * acc;
}
// ymu[i] = m(x*_i) + sum of the chunk partials in chunk order  (gp.m:155-164; the mean as in predict_finish_kernel)
__global__ void predict_mean_finish_kernel(const double* __restrict__ Xs, long ns, int D, int mean_mask, const double* __restrict__ hyp_mean,
                                           const double* __restrict__ parts, int nchunk, double* __restrict__ ymu) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns) return;
  double m = 0.0; int o = 0;
  if (mean_mask & 1) { for (int d = 0; d < D; ++d) m += Xs[i * D + d] * hyp_mean[d]; o = D; }
  if (mean_mask & 2) m += hyp_mean[o];
  double k = 0.0;
  for (int ch = 0; ch < nchunk; ++ch) k += parts[(long)ch * ns + i];
  ymu[i] = m + k;
}
// set-up, once per device, with that device current (kgp_ctx_create -> exact_init)
static int noalloc_setup(kgp_ctx* c) {
  if (c->device < 0 || c->device >= NOALLOC_MAX_DEV || g_noalloc_ptr[c->device]) return KGP_OK;
  double* p = nullptr;
  KGP_CUDA(c, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
  g_noalloc_ptr[c->device] = p;
  return KGP_OK;
}
template <int DP>
static void launch_predict_mean_parts(kgp_ctx* c, const double* Xs, long ns, int D, long n, double sf2, int nchunk, double* parts) {
  dim3 grid((unsigned)((ns + PM_THREADS - 1) / PM_THREADS), (unsigned)nchunk);
  predict_mean_parts_kernel<DP><<<grid, PM_THREADS, 0, c->stream>>>(Xs, ns, D, c->mu.d(), c->inv_ell.d(), c->xs.d(), c->alpha.d(), n, sf2, parts);
}
}  // namespace kgp

int kgp_predict_mean_noalloc_dev(kgp_ctx* c, const double* Xs_dev, long ns, double* ymu_dev) {
  using namespace kgp;
  if (!c) return KGP_BAD_ARG;
  if (!Xs_dev || !ymu_dev || ns <= 0) return c->fail(KGP_BAD_ARG, "bad test inputs / output");
  if (!c->model.set || c->model.inf_kind != KGP_INF_EXACT || c->model.cov_kind == KGP_COV_SM)
    return c->fail(KGP_BAD_ARG, "kgp_predict_mean_noalloc_dev serves infExact with covSEiso / covSEard");
  const int slot = KGP_SLOT_TR;
  if (!state_valid(c, slot, 0))
    return c->fail(KGP_NO_DATA, "no posterior for the current data / hyper-parameters: call kgp_eval first (this entry point "
                   "cannot factorise, that would allocate)");
  const Slot& S = c->slots[slot]; const Model& m = c->model;
  const long n = S.n; const int D = S.D, Dp = c->Dp;
  const long nchunk = (n + PM_CHUNK - 1) / PM_CHUNK;
  if (ns * nchunk > NOALLOC_SCRATCH_DOUBLES)
    return c->fail(KGP_BAD_ARG, "ns = %ld test points x %ld chunks of %d training points = %ld doubles of scratch; the static array holds %ld "
                   "(2 MiB): at most %ld test points per call at n = %ld", ns, nchunk, PM_CHUNK, ns * nchunk, NOALLOC_SCRATCH_DOUBLES,
                   NOALLOC_SCRATCH_DOUBLES / nchunk, n);
  if (c->device < 0 || c->device >= NOALLOC_MAX_DEV || !g_noalloc_ptr[c->device])
    return c->fail(KGP_CUDA_ERR, "static scratch not resolved on device %d", c->device);
  cudaSetDevice(c->device);
  double* parts = g_noalloc_ptr[c->device];
  const int nl = m.cov_kind == KGP_COV_SEARD ? D : 1;
  const double sf2 = exp(2.0 * m.hyp_cov[nl]);
  switch (Dp) {
    case 4: launch_predict_mean_parts<4>(c, Xs_dev, ns, D, n, sf2, (int)nchunk, parts); break;
    case 8: launch_predict_mean_parts<8>(c, Xs_dev, ns, D, n, sf2, (int)nchunk, parts); break;
    case 16: launch_predict_mean_parts<16>(c, Xs_dev, ns, D, n, sf2, (int)nchunk, parts); break;
    case 32: launch_predict_mean_parts<32>(c, Xs_dev, ns, D, n, sf2, (int)nchunk, parts); break;
    default: return c->fail(KGP_BAD_ARG, "unexpected padded input dimension %d", Dp);
  }
  predict_mean_finish_kernel<<<(unsigned)((ns + 255) / 256), 256, 0, c->stream>>>(Xs_dev, ns, D, m.mean_mask, c->hypm.d(), parts, (int)nchunk, ymu_dev);
  KGP_CUDA(c, cudaStreamSynchronize(c->stream));
  KGP_CUDA(c, cudaGetLastError());
  return KGP_OK;
}

namespace kgp {
int exact_init(kgp_ctx* c) {
  { int rc = noalloc_setup(c); if (rc) return rc; }
  KGP_CUDA(c, gemm_init());
  KGP_CUDA(c, gemm_tma_init());
  KGP_CUDA(c, cudaFuncSetAttribute(potrf_diag128_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PD_SMEM));
  KGP_CUDA(c, cudaFuncSetAttribute(trsm128_rows_kernel_t<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, TR_SMEM));
  KGP_CUDA(c, cudaFuncSetAttribute(trsm128_rows_kernel_t<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, TR_SMEM));
  KGP_CUDA(c, cudaFuncSetAttribute(kbuild_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kbuild_smem(32)));
  return KGP_OK;
}

}  // namespace kgp
