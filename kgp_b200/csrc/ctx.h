// Engine context: device buffers + model state.  One per kgp_ctx (C ABI).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <stdarg.h>
#include <stdio.h>
#include "../../include/kgp_b200.h"

namespace kgp {

struct DevBuf {
  void* p = nullptr; size_t bytes = 0;
  cudaError_t ensure(size_t need) {
    if (need <= bytes) return cudaSuccess;
    if (p) cudaFree(p);
    p = nullptr; bytes = 0;
    cudaError_t e = cudaMalloc(&p, need);
    if (e == cudaSuccess) bytes = need;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; bytes = 0; }
  double* d() const { return (double*)p; }
};

struct Slot {
  DevBuf X, y; long n = 0; int D = 0; bool has_y = false;
  DevBuf Xraw; int Draw = 0;               // hyp.proj set: the un-projected inputs (n x Draw); X then holds Xraw hP' (n x d)
  unsigned long version = 0;
};

struct Model {
  int inf_kind = KGP_INF_EXACT, cov_kind = KGP_COV_SEISO, sm_Q = 0, mean_mask = 0;
  std::vector<double> hyp_cov, hyp_mean; double hyp_lik = 0.0;
  kgp_opt opt; bool set = false;
  unsigned long version = 0;
};

struct Grid {
  std::vector<std::vector<double>> axes;   // xg{d} = equispaced axis values
  std::vector<int> toep;                   // xg{d} was a nested cell {axis}: Toeplitz factor (apxGrid.m:258-281)
  std::vector<int> neq;                    // axis fails equi() (apxGrid.m:684-695): linear neqinterp instead of cubic eqinterp
  bool set = false; unsigned long version = 0;
};

constexpr int NB = 512;     // outer block width of the blocked factorisations

}  // namespace kgp

struct kgp_ctx {
  int device = 0;
  cudaStream_t stream = nullptr, stream2 = nullptr;   // main stream; side stream for work overlapped with it
  cudaEvent_t ev[16];
  std::string err;
  kgp::Slot slots[KGP_NUM_SLOTS];
  kgp::Model model;
  kgp::Grid grid;
  kgp_stats stats;
  int profiling = 0;
  std::vector<cudaEvent_t> gev; size_t gev_used = 0; std::vector<int> gev_class; std::vector<double> gev_flops; int gemm_class = 0;   // event pairs around dgemm_nt launches (profiling)

  // exact-path device state
  kgp::DevBuf M, W2, Dws;             // L (lower) | Z upper -> full B^{-1} | diagonal nb-blocks of B^{-1}
  kgp::DevBuf gws[2], gcnt[2];        // small-tile GEMM (dgemm_nt_small.cuh): split-K partials + arrival counters, [main | panel stream]
  kgp::DevBuf xs, s, mu, inv_ell, ym, alpha, tvec, parts, rows, scal, hypm, dX, fac;
  kgp::DevBuf pred_K, pred_x, pred_s, pred_a, pred_b, pred_out, pred_X;
  kgp::DevBuf smV, smM, smAcc;        // covSM exact path: 4 pi^2 v, 2 pi m tables and the reduction accumulators
  kgp::DevBuf projP, projXs, projdX;  // hyp.proj (infGrid.m:56-65): effective projector hP (d x D), projected test inputs, dnlZ/d(x hP')
  int proj_d = 0, proj_D = 0;
  int* d_info = nullptr;
  long np = 0; int Dp = 0;
  // what the factorisation state corresponds to
  unsigned long st_slot_ver = 0, st_model_ver = 0; int st_slot = -1; int st_level = -1;  // -1 none, 0 L+alpha, 2 +dX
  double st_nlZ = 0.0; std::vector<double> st_dnlZ;
  // multi-GPU (exact_dist.cu): one context per rank, NCCL communicator over all ranks of the job
  int nranks = 1, rank = 0; void* nccl_comm = nullptr;
  cudaStream_t stream_p = nullptr;                    // panel stream: factorisation of the next panel + its broadcast
  cudaEvent_t dev_ev[8];                              // panel ready [2], panel buffer free [2], look-ahead column updated
  cudaEvent_t zev[8]; bool zpipe_active = false;      // "rank r's rows of Z = L^-T have arrived" (pipelined all-gather under Z Z^T)
  kgp::DevBuf PB[2];                                  // packed panels (rows x NB)
  // pipelined all-gather of Z = L^-T by peer copies (exact.cu, zsweep_dist): every rank's W2 mapped into every peer with CUDA IPC;
  // w2_mapped = the local allocation the peers hold a mapping of; one copy stream + event per peer; zbar = 8-byte all-reduce word
  void* w2_peer[8] = {nullptr}; void* w2_mapped = nullptr; bool w2_push_ok = false;
  std::vector<void*> w2_retired;                      // allocations outgrown while peers still map them: freed once the peers have closed
  cudaStream_t zstream[8] = {nullptr}; cudaEvent_t zpush_ev[8] = {nullptr}; cudaEvent_t zcol_ev = nullptr; double* zbar = nullptr;
  // peer exchange window of the row-sharded KISS-GP CG (grid_cg.cuh, CgDist): one cudaMalloc per rank, mapped into
  // every peer with CUDA IPC; px_epoch = tag of the last exchange (monotone over all solves, equal on all ranks)
  void* px_local = nullptr; void* px_peer[8] = {nullptr}; bool px_ok = false; unsigned long long px_epoch = 0;
  int* px_abort = nullptr;
  double dist_ms_comm = 0.0;
  // grid-path state lives in grid.cu (opaque)
  void* gridstate = nullptr;

  int fail(int code, const char* fmt, ...) {
    char buf[512]; va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    err = buf; return code;
  }
};

#define KGP_CUDA(ctx, call)                                                              \
  do { cudaError_t e__ = (call);                                                         \
       if (e__ != cudaSuccess)                                                           \
         return (ctx)->fail(KGP_CUDA_ERR, "%s:%d %s: %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
  } while (0)
