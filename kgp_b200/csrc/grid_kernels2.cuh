// KISS-GP path, part 2: CG vector kernels, symmetric eigen-solver for the Kronecker factors,
// Fiedler bound and its derivatives.  See grid_kernels.cuh for citations.
#pragma once
#include "grid_kernels.cuh"

namespace kgp {

// device scalars of the CG solve (one double array):
enum { CG_R2A = 0, CG_R2B = 1, CG_NB = 2, CG_RELRES = 3, CG_DONE = 4, CG_ITER = 5, CG_YTA = 6, CG_ATA = 7, CG_NSCAL = 16 };

// sum of `cnt` partials, computed identically by every thread block (deterministic order)
__device__ __forceinline__ double sum_partials(const double* __restrict__ part, int cnt, double* sh) {
  double a = 0.0;
  for (int q = threadIdx.x; q < cnt; q += blockDim.x) a += part[q];
  return block_sum(a, sh);
}

// b = sw*(y-m) -> r, d ; x = 0 ; partial r.r
__global__ void cg_init_kernel(const double* __restrict__ ym, double sw, long n, double* __restrict__ x,
                               double* __restrict__ r, double* __restrict__ d, double* __restrict__ part) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    double b = sw * ym[i];
    x[i] = 0.0; r[i] = b; d[i] = b; acc = fma(b, b, acc);
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}
// scalars after init: r2, nb, relres, done (apx.m:284-287)
__global__ void cg_init_scalars_kernel(const double* __restrict__ part, int cnt, double tol, double* __restrict__ S) {
  __shared__ double sh[32];
  double r2 = sum_partials(part, cnt, sh);
  if (threadIdx.x == 0) {
    S[CG_R2A] = r2; S[CG_R2B] = r2; S[CG_NB] = sqrt(r2);
    double rel = sqrt(r2) / sqrt(r2);
    S[CG_RELRES] = rel; S[CG_ITER] = 1.0;
    S[CG_DONE] = (rel >= tol) ? 0.0 : 1.0;            // NaN (b == 0) -> not "todo" -> flag 1, like apx.m:286
  }
}
// a = r2/(d.z); x += a d; r -= a z; partial r.r     (apx.m:291-294)
__global__ void cg_update1_kernel(long n, const double* __restrict__ partdz, int cnt, const double* __restrict__ S,
                                  int parity, double* __restrict__ x, double* __restrict__ r,
                                  const double* __restrict__ d, const double* __restrict__ z, double* __restrict__ part) {
  __shared__ double sh[32];
  if (S[CG_DONE] != 0.0) return;
  const double dz = sum_partials(partdz, cnt, sh);
  const double a = S[parity ? CG_R2B : CG_R2A] / dz;
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    x[i] = fma(a, d[i], x[i]);
    double ri = r[i] - a * z[i];
    r[i] = ri; acc = fma(ri, ri, acc);
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}
// relres, convergence, Fletcher-Reeves beta, d = r + beta d   (apx.m:295-301)
__global__ void cg_update2_kernel(long n, const double* __restrict__ part, int cnt, double* __restrict__ S,
                                  int parity, double tol, int iter, const double* __restrict__ r, double* __restrict__ d) {
  __shared__ double sh[32];
  if (S[CG_DONE] != 0.0) return;
  const double r2new = sum_partials(part, cnt, sh);
  const double r2old = S[parity ? CG_R2B : CG_R2A];
  const double rel = sqrt(r2new) / S[CG_NB];
  const bool conv = rel < tol;
  if (!conv) {
    const double beta = r2new / r2old;
    for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
      d[i] = fma(beta, d[i], r[i]);
  }
  // Scalars for LATER launches: the other parity slot and the flags.  Blocks of this launch that
  // start after block 0 finished may already see CG_DONE = 1 and return early, which is harmless
  // (a converged iteration updates nothing but these scalars).
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    S[parity ? CG_R2A : CG_R2B] = r2new;
    S[CG_RELRES] = rel; S[CG_ITER] = (double)iter;
    if (conv) S[CG_DONE] = 1.0;
  }
}

}  // namespace kgp
