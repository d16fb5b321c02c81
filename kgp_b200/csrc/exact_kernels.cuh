// Device kernels of the exact-GP path (infExact).  Reference arithmetic being
// reproduced: GPML/cov/covMaha.m:125-154 (centred squared distances, clamp),
// covSE.m:23 / covScale.m:74 (K = sf^2 exp(-D2/2)), GPML/cov/apx.m:160-166
// (chol, log-det, inverse), GPML/inf/infGaussLik.m:25-32 (alpha, nlZ, dnlZ),
// kgp/backend/utils/dlikExact.m:19-25 (input gradient), GPML/gp.m:151-164 (predict).
#pragma once
#include "dgemm_nt_tma.cuh"
#include <math.h>

namespace kgp {

// ---------------------------------------------------------------------------
// small deterministic reductions
// ---------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// block-wide sum, result valid in thread 0 (and broadcast through smem to all)
__device__ __forceinline__ double block_sum(double v, double* sh /* >= 32 doubles */) {
  int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) sh[w] = v;
  __syncthreads();
  double r = 0.0;
  if (w == 0) {
    r = lane < nw ? sh[lane] : 0.0;
    r = warp_sum(r);
    if (lane == 0) sh[0] = r;
  }
  __syncthreads();
  return sh[0];
}

// column means of X (n x D): one CTA per column   [covMaha.m:138 mu = mean(x,1)]
__global__ void col_mean_kernel(const double* __restrict__ X, long n, int D, double* __restrict__ mu) {
  __shared__ double sh[32];
  int d = blockIdx.x;
  double a = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) a += X[i * D + d];
  double t = block_sum(a, sh);
  if (threadIdx.x == 0) mu[d] = t / (double)n;
}

// xs[i][d] = (x[i][d]-mu[d]) * inv_ell[d]  (zero padded to np x Dp);  s[i] = sum_d xs^2
__global__ void scale_x_kernel(const double* __restrict__ X, long n, int D, const double* __restrict__ mu,
                               const double* __restrict__ inv_ell, double* __restrict__ xs,
                               double* __restrict__ s, long np, int Dp) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  double acc = 0.0;
  for (int d = 0; d < Dp; ++d) {
    double v = 0.0;
    if (i < n && d < D) v = (X[i * D + d] - mu[d]) * inv_ell[d];
    xs[i * Dp + d] = v;
    acc += v * v;
  }
  s[i] = acc;
}

// ym = y - m(x), m = meanSum{meanLinear, meanConst} selected by mask  (meanSum.m:13-37)
__global__ void resid_mean_kernel(const double* __restrict__ X, const double* __restrict__ y, long n, int D,
                                  int mean_mask, const double* __restrict__ hyp_mean, double* __restrict__ ym,
                                  long np) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  if (i >= n) { ym[i] = 0.0; return; }
  double m = 0.0; int o = 0;
  if (mean_mask & 1) { for (int d = 0; d < D; ++d) m += X[i * D + d] * hyp_mean[d]; o = D; }
  if (mean_mask & 2) m += hyp_mean[o];
  ym[i] = (y ? y[i] : 0.0) - m;
}

// ---------------------------------------------------------------------------
// kernel-matrix tiles:  out[i][j] = scale*exp(-max(s_i+s_j-2 xs_i.xs_j,0)/2) (+1 on the diagonal)
// rows/cols beyond nR/nC (padding) give the identity pattern.  The Gram part runs
// on the FP64 tensor pipe (DMMA.8x8x4 over Dp/4 k-steps).
// ---------------------------------------------------------------------------
constexpr int KB_THREADS = 256;
__global__ void __launch_bounds__(KB_THREADS) kbuild_kernel(
    const double* __restrict__ xr, const double* __restrict__ sr, long nR,
    const double* __restrict__ xc, const double* __restrict__ sc, long nC,
    int Dp, double scale, int diag_add, int lower_only, double* __restrict__ out, long ld) {
  const int tn = blockIdx.x, tm = blockIdx.y;
  if (lower_only && tn > tm) return;
  extern __shared__ __align__(16) double sm[];
  const int ldx = ((Dp + 11) / 16) * 16 + 4;          // ldx % 16 == 4 -> conflict-free LDS.64 fragments
  double* sxr = sm; double* sxc = sm + 128 * ldx;
  double* ssr = sxc + 128 * ldx; double* ssc = ssr + 128;
  const long i0 = (long)tm * 128, j0 = (long)tn * 128;
  for (int q = threadIdx.x; q < 128 * Dp; q += KB_THREADS) {
    int r = q / Dp, d = q % Dp;
    sxr[r * ldx + d] = xr[(i0 + r) * Dp + d];
    sxc[r * ldx + d] = xc[(j0 + r) * Dp + d];
  }
  if (threadIdx.x < 128) ssr[threadIdx.x] = sr[i0 + threadIdx.x];
  else ssc[threadIdx.x - 128] = sc[j0 + threadIdx.x - 128];
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gq = lane >> 2, tq = lane & 3;
  double acc[2][16][2];
#pragma unroll
  for (int a = 0; a < 2; ++a)
#pragma unroll
    for (int b = 0; b < 16; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;
  for (int k = 0; k < Dp; k += 4) {
    double af[2];
#pragma unroll
    for (int a = 0; a < 2; ++a) af[a] = sxr[(warp * 16 + a * 8 + gq) * ldx + k + tq];
#pragma unroll
    for (int b = 0; b < 16; ++b) {
      double bf = sxc[(b * 8 + gq) * ldx + k + tq];
#pragma unroll
      for (int a = 0; a < 2; ++a) dmma884(acc[a][b][0], acc[a][b][1], af[a], bf);
    }
  }
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    int rl = warp * 16 + a * 8 + gq;
    long gi = i0 + rl;
    double si = ssr[rl];
#pragma unroll
    for (int b = 0; b < 16; ++b) {
      double2 v;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        int cl = b * 8 + tq * 2 + h;
        long gj = j0 + cl;
        double val;
        if (gi < nR && gj < nC) {
          double d2 = fmax(si + ssc[cl] - 2.0 * acc[a][b][h], 0.0);     // covMaha.m:150
          val = scale * exp(-0.5 * d2);
        } else val = 0.0;
        if (diag_add && gi == gj) val += 1.0;
        if (h == 0) v.x = val; else v.y = val;
      }
      *reinterpret_cast<double2*>(out + gi * ld + j0 + b * 8 + tq * 2) = v;
    }
  }
}
inline size_t kbuild_smem(int Dp) { int ldx = ((Dp + 11) / 16) * 16 + 4; return (size_t)(2 * 128 * ldx + 256) * 8; }

// ---------------------------------------------------------------------------
// Cholesky of one 128x128 diagonal block in shared memory (lower, row-major).
// info: 1-based global index of the first non-positive pivot (atomicMin), else unchanged.
// ---------------------------------------------------------------------------
constexpr int PD_OWNERS = 512, PD_THREADS = PD_OWNERS + 32;
constexpr int PD_SMEM = 0;
// Cholesky of one 128 x 128 diagonal block (lower triangle, in place), one CTA: 16 owner warps + ONE PANEL WARP.
// The block lives in the owners' registers: thread (ti, tj) holds rows 4 ti..4 ti+3 x columns 8 tj..8 tj+7.  Columns are processed
// EIGHT at a time: the owners publish the 128 x 8 panel in shared memory, the panel warp (which holds no part of the block, so its
// registers carry only the panel: lane l = rows 4 l..4 l+3) factors it with shuffles alone -- eight dependent
// rsqrt / scale / update steps without a block barrier --, then every owner applies the rank-8 update to its 32 elements from
// shared memory.  Three block barriers per eight columns (48) instead of one per column (128): the kernel is a chain of dependent
// FP64 operations and sits on the critical path of every panel (small problems; the multi-GPU panel chain, where it was
// 4 x ~100 us of every ~1 ms step).  The two roles run in separate branches (register allocation = the larger of the two, not
// their sum) and meet at a NAMED barrier with an explicit thread count (bar.sync 1, 544).
// Shared layout: element (row, k) at pan[k][(row & 7) * 16 + (row >> 3)] -- the rank-8 update reads rows 8 tj + c across a
// half-warp, which this makes contiguous (conflict-free).
__device__ __forceinline__ int pd_slot(int row) { return (row & 7) * 16 + (row >> 3); }
__device__ __forceinline__ void pd_bar() { asm volatile("bar.sync 1, %0;" ::"n"(PD_THREADS) : "memory"); }
__global__ void __launch_bounds__(PD_THREADS) potrf_diag128_kernel(double* __restrict__ A, long ld, int j0, int* info) {
  __shared__ double pan[8][128];
  const int tid = threadIdx.x;
  if (tid >= PD_OWNERS) {                              // ---- panel warp
    const int lane = tid & 31;
    for (int kb = 0; kb < 16; ++kb) {
      pd_bar();                                        // panel published
      double p[4][8];
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) p[r][c] = pan[c][pd_slot(4 * lane + r)];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int krow = 8 * kb + k;                   // pivot row: lane 2 kb + (k >> 2), local row k & 3
        double akk = __shfl_sync(0xffffffffu, p[k & 3][k], 2 * kb + (k >> 2));
        if (!(akk > 0.0)) { if (lane == 0) atomicMin(info, j0 + krow + 1); akk = nan(""); }
        const double inv = rsqrt(akk), dk = akk * inv;
        double lc[8];
#pragma unroll
        for (int c = k + 1; c < 8; ++c) lc[c] = __shfl_sync(0xffffffffu, p[c & 3][k], 2 * kb + (c >> 2)) * inv;   // L[8 kb + c][k]
#pragma unroll
        for (int r = 0; r < 4; ++r) {
          const int row = 4 * lane + r;
          if (row > krow) {
            const double lr = p[r][k] * inv;
#pragma unroll
            for (int c = k + 1; c < 8; ++c) p[r][c] = fma(-lr, lc[c], p[r][c]);
            p[r][k] = lr;
          } else if (row == krow) p[r][k] = dk;
        }
      }
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) pan[c][pd_slot(4 * lane + r)] = p[r][c];
      pd_bar();                                        // panel factored
      pd_bar();                                        // owners are done reading it
    }
    return;
  }
  // ---- owners
  const int ti = tid >> 4, tj = tid & 15;
  const int r0 = ti * 4, c0 = tj * 8;
  const bool active = c0 <= r0 + 3;                    // the 4 x 8 block touches the lower triangle
  double a[4][8];
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 8; c += 2) {
      double2 v = make_double2(0.0, 0.0);
      if (active) v = *reinterpret_cast<const double2*>(A + (long)(r0 + r) * ld + c0 + c);
      a[r][c] = v.x; a[r][c + 1] = v.y;
    }
  for (int kb = 0; kb < 16; ++kb) {
    if (tj == kb) {                                    // publish the panel (rows above it are never read)
#pragma unroll
      for (int r = 0; r < 4; ++r)
#pragma unroll
        for (int c = 0; c < 8; ++c) pan[c][pd_slot(r0 + r)] = a[r][c];
    }
    pd_bar();
    pd_bar();                                          // (the panel warp factors)
    if (active) {
      if (tj == kb) {                                  // the factored panel back into the owners' registers
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 8; ++c) if (r0 + r >= 8 * kb) a[r][c] = pan[c][pd_slot(r0 + r)];
      } else if (tj > kb) {                            // rank-8 update of the columns to the right of the panel
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          double lr[4], lcv[8];
#pragma unroll
          for (int r = 0; r < 4; ++r) lr[r] = pan[k][pd_slot(r0 + r)];
#pragma unroll
          for (int c = 0; c < 8; ++c) lcv[c] = pan[k][pd_slot(c0 + c)];
#pragma unroll
          for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int c = 0; c < 8; ++c) a[r][c] = fma(-lr[r], lcv[c], a[r][c]);
        }
      }
    }
    pd_bar();                                          // the panel buffer is rewritten in the next round
  }
  if (active) {
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int c = 0; c < 8; ++c)
        if (c0 + c <= r0 + r) A[(long)(r0 + r) * ld + c0 + c] = a[r][c];
  }
}

// ---------------------------------------------------------------------------
// X <- X * L^{-T} for a 128x128 lower-triangular L.  One warp owns 8 rows; the row values live in the DMMA
// C-fragment layout (row gq, columns 2tq, 2tq+1 of each 8-column block), so everything is DMMA.8x8x4:
//   * right-looking over the 16 column blocks: as soon as block p is solved, the 15-p later blocks are updated
//     with independent DMMA pairs (k-permutation: step h uses columns 2t+h), B fragments straight from shared L;
//   * the 8x8 diagonal solve is a product with the inverse of the 8x8 diagonal block (formed once per CTA by
//     forward substitution, one thread per column): two more DMMAs instead of eight shuffle + divide steps.
// Only the 8x8 diagonal blocks are ever inverted (conditioning of an 8x8 block of a Cholesky factor); the
// off-diagonal part is plain substitution.  ~1.1 MFMA per CTA: 9 us at the FP64 tensor rate (was ~60 us of
// dependent shuffles and divisions).
// ---------------------------------------------------------------------------
constexpr int TR_THREADS = 512;                 // 16 warps x 8 rows = 128 rows per CTA
constexpr int TR_LDL = 136;                     // % 16 == 8 -> conflict-free LDS.128 B fragments
constexpr int TR_SMEM = (128 * TR_LDL + 16 * 64) * 8;
__global__ void __launch_bounds__(TR_THREADS, 1) trsm128_rows_kernel(
    double* __restrict__ X, long ldx, long nrows, const double* __restrict__ L, long ldl) {
  extern __shared__ __align__(16) double Ls[];
  double* Di = Ls + 128 * TR_LDL;               // [16][8][8] inverses of the diagonal 8x8 blocks (row-major)
  for (int q = threadIdx.x; q < 128 * 64; q += TR_THREADS) {
    int r = q >> 6, c2 = (q & 63) * 2;
    double2 v = make_double2(0.0, 0.0);
    if (c2 <= r) v = *reinterpret_cast<const double2*>(L + (long)r * ldl + c2);
    *reinterpret_cast<double2*>(Ls + r * TR_LDL + c2) = v;
  }
  __syncthreads();
  if (threadIdx.x < 128) {                      // column j of inv(D_blk): D y = e_j
    const int blk = threadIdx.x >> 3, j = threadIdx.x & 7;
    const double* D = Ls + (blk * 8) * TR_LDL + blk * 8;
    double y[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      double acc = (i == j) ? 1.0 : 0.0;
#pragma unroll
      for (int k = 0; k < i; ++k) acc = fma(-D[i * TR_LDL + k], y[k], acc);
      y[i] = (i < j) ? 0.0 : acc / D[i * TR_LDL + i];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) Di[blk * 64 + i * 8 + j] = y[i];
  }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, gq = lane >> 2, tq = lane & 3;
  const long row = (long)blockIdx.x * 128 + warp * 8 + gq;
  if ((long)blockIdx.x * 128 + warp * 8 >= nrows) return;
  double* xp = X + row * ldx + tq * 2;
  double x[16][2];
#pragma unroll
  for (int cb = 0; cb < 16; ++cb) {
    double2 v = *reinterpret_cast<const double2*>(xp + cb * 8);
    x[cb][0] = v.x; x[cb][1] = v.y;
  }
#pragma unroll
  for (int p = 0; p < 16; ++p) {
    // solve block p: x_p <- c_p * inv(D_p)^T
    {
      const double2 di = *reinterpret_cast<const double2*>(Di + p * 64 + gq * 8 + tq * 2);
      double n0 = 0.0, n1 = 0.0;
      dmma884(n0, n1, x[p][0], di.x);
      dmma884(n0, n1, x[p][1], di.y);
      x[p][0] = n0; x[p][1] = n1;
    }
    const double a0 = -x[p][0], a1 = -x[p][1];
#pragma unroll
    for (int cb = p + 1; cb < 16; ++cb) {       // c_cb -= x_p * L[cb, p]^T
      const double2 l = *reinterpret_cast<const double2*>(Ls + (cb * 8 + gq) * TR_LDL + p * 8 + tq * 2);
      dmma884(x[cb][0], x[cb][1], a0, l.x);
      dmma884(x[cb][0], x[cb][1], a1, l.y);
    }
  }
#pragma unroll
  for (int cb = 0; cb < 16; ++cb)
    *reinterpret_cast<double2*>(xp + cb * 8) = make_double2(x[cb][0], x[cb][1]);
}

// w x w block <- identity
__global__ void set_identity_kernel(double* __restrict__ A, long ld, int w) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)w * w) return;
  int r = (int)(q / w), c = (int)(q % w);
  A[(long)r * ld + c] = (r == c) ? 1.0 : 0.0;
}

// sum_i log(L_ii), i < n  (apx.m:161); single CTA, deterministic
__global__ void logdet_kernel(const double* __restrict__ L, long ld, long n, double* __restrict__ out) {
  __shared__ double sh[32];
  double a = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) a += log(L[i * ld + i]);
  double t = block_sum(a, sh);
  if (threadIdx.x == 0) out[0] = t;
}

}  // namespace kgp
