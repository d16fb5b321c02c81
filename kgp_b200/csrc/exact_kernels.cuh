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
// Cholesky of one 128 x 128 diagonal block (lower triangle, in place), one CTA: 16 owner warps + ONE PANEL WARP.
// info: 1-based global index of the first non-positive pivot (atomicMin), else unchanged.
//
// The kernel is a chain of dependent FP64 operations and sits on the critical path of every panel (the reference-sized
// problems, where four of these are most of the factorisation; the multi-GPU panel chain).  Round-2 form:
//   * the block lives in the owners' registers as 8 x 8 tiles in the DMMA C-fragment layout: warp w holds tile row w
//     (tiles (w, 0..w)), so the rank-8 update of a tile with a factored 128 x 8 panel is TWO DMMA.8x8x4 (k = 0..3, 4..7)
//     instead of 64 scalar DFMAs per thread -- the update stops being the larger half of a step;
//   * columns are processed eight at a time by the panel warp (lane l = rows l, l+32, l+64, l+96 of the panel), with
//     shuffles alone: eight dependent rsqrt / scale / update steps without a block barrier;
//   * look-ahead: the owners first update tile column kb+1 with panel kb and publish it, THEN update the rest while the
//     panel warp already factors panel kb+1 (two barriers per eight columns, the panel warp never waits for the bulk).
// Shared memory: pin[c][row] (the published panel, column-major: the panel warp reads it conflict-free),
// pout[2][row][12] (the factored panel, row stride 12 doubles: gq*12 + tq hits 16 distinct banks per half-warp for the
// owners' fragment loads; double-buffered because the owners still read panel kb while panel kb+1 is written).
// The two roles run in separate branches and meet at a NAMED barrier with an explicit thread count (bar.sync 1, 544).
// ---------------------------------------------------------------------------
constexpr int PD_OWNERS = 512, PD_THREADS = PD_OWNERS + 32;
constexpr int PD_SMEM = 0;
constexpr int PD_PIN_LD = 132, PD_POUT_LD = 12;
__device__ __forceinline__ void pd_bar() { asm volatile("bar.sync 1, %0;" ::"n"(PD_THREADS) : "memory"); }
__global__ void __launch_bounds__(PD_THREADS) potrf_diag128_kernel(double* __restrict__ A, long ld, int j0, int* info) {
  __shared__ __align__(16) double pin[8 * PD_PIN_LD];
  __shared__ __align__(16) double pout[2][128 * PD_POUT_LD];
  const int tid = threadIdx.x;
  if (tid >= PD_OWNERS) {                              // ---- panel warp
    const int lane = tid & 31;
#pragma unroll
    for (int q = 0; q < 4; ++q) {                      // panels 4q .. 4q+3 have their pivot rows in register row q
      for (int kk = 0; kk < 4; ++kk) {
        const int kb = 4 * q + kk;
        pd_bar();                                      // panel kb published
        double p[4][8];
#pragma unroll
        for (int r = q; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 8; ++c) p[r][c] = pin[c * PD_PIN_LD + lane + 32 * r];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int krow = 8 * kb + k, src = 8 * kk + k;          // pivot row krow = 32 q + src
          double akk = __shfl_sync(0xffffffffu, p[q][k], src);
          if (!(akk > 0.0)) { if (lane == 0) atomicMin(info, j0 + krow + 1); akk = nan(""); }
          const double inv = rsqrt(akk), dk = akk * inv;
          double lc[8];
#pragma unroll
          for (int c = k + 1; c < 8; ++c) lc[c] = __shfl_sync(0xffffffffu, p[q][k], 8 * kk + c) * inv;   // L[8 kb + c][k]
#pragma unroll
          for (int r = q; r < 4; ++r) {
            const int row = lane + 32 * r;
            if (row > krow) {
              const double lr = p[r][k] * inv;
#pragma unroll
              for (int c = k + 1; c < 8; ++c) p[r][c] = fma(-lr, lc[c], p[r][c]);
              p[r][k] = lr;
            } else if (row == krow) p[r][k] = dk;
          }
        }
        double* po = pout[kb & 1];
#pragma unroll
        for (int r = q; r < 4; ++r)
#pragma unroll
          for (int c = 0; c < 8; c += 2)
            *reinterpret_cast<double2*>(po + (lane + 32 * r) * PD_POUT_LD + c) = make_double2(p[r][c], p[r][c + 1]);
        pd_bar();                                      // panel kb factored
      }
    }
    return;
  }
  // ---- owners: warp w = tile row w
  const int w = tid >> 5, lane = tid & 31, gq = lane >> 2, tq = lane & 3;
  const int row = 8 * w + gq;
  double c[16][2];
#pragma unroll
  for (int J = 0; J < 16; ++J) {
    c[J][0] = c[J][1] = 0.0;
    if (J <= w) {
      const double2 v = *reinterpret_cast<const double2*>(A + (long)row * ld + 8 * J + 2 * tq);
      c[J][0] = v.x; c[J][1] = v.y;
    }
  }
  pin[(2 * tq) * PD_PIN_LD + row] = c[0][0];           // publish panel 0
  pin[(2 * tq + 1) * PD_PIN_LD + row] = c[0][1];
  pd_bar();
  pd_bar();
#pragma unroll
  for (int kb = 0; kb < 15; ++kb) {
    const double* po = pout[kb & 1];
    if (w >= kb) {                                     // final values of tile (w, kb)
      const double2 v = *reinterpret_cast<const double2*>(po + row * PD_POUT_LD + 2 * tq);
      c[kb][0] = v.x; c[kb][1] = v.y;
    }
    double a0 = 0.0, a1 = 0.0;
    if (w > kb) {
      a0 = -po[row * PD_POUT_LD + tq]; a1 = -po[row * PD_POUT_LD + tq + 4];
      const double b0 = po[(8 * (kb + 1) + gq) * PD_POUT_LD + tq], b1 = po[(8 * (kb + 1) + gq) * PD_POUT_LD + tq + 4];
      dmma884(c[kb + 1][0], c[kb + 1][1], a0, b0);
      dmma884(c[kb + 1][0], c[kb + 1][1], a1, b1);
      pin[(2 * tq) * PD_PIN_LD + row] = c[kb + 1][0];  // publish panel kb+1
      pin[(2 * tq + 1) * PD_PIN_LD + row] = c[kb + 1][1];
    }
    pd_bar();                                          // the panel warp factors kb+1 while the rest is updated
#pragma unroll
    for (int J = kb + 2; J < 16; ++J) {
      if (J <= w) {
        const double b0 = po[(8 * J + gq) * PD_POUT_LD + tq], b1 = po[(8 * J + gq) * PD_POUT_LD + tq + 4];
        dmma884(c[J][0], c[J][1], a0, b0);
        dmma884(c[J][0], c[J][1], a1, b1);
      }
    }
    pd_bar();
  }
  if (w == 15) {
    const double2 v = *reinterpret_cast<const double2*>(pout[1] + row * PD_POUT_LD + 2 * tq);
    c[15][0] = v.x; c[15][1] = v.y;
  }
#pragma unroll
  for (int J = 0; J < 16; ++J) {
    if (J < w) *reinterpret_cast<double2*>(A + (long)row * ld + 8 * J + 2 * tq) = make_double2(c[J][0], c[J][1]);
    else if (J == w) {
      if (2 * tq <= gq) A[(long)row * ld + 8 * J + 2 * tq] = c[J][0];
      if (2 * tq + 1 <= gq) A[(long)row * ld + 8 * J + 2 * tq + 1] = c[J][1];
    }
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
// WARPS = 16: 128 rows per CTA (large problems: the L block is loaded once per 128 rows).  WARPS = 4: 32 rows per
// CTA, for the reference-sized problems where a launch has only a few hundred rows: the work of a row block is
// bound by ONE SM's DMMA rate (9 us per 128 rows), so spreading the rows over 4x as many SMs cuts the latency.
constexpr int TR_THREADS = 512;                 // 16 warps x 8 rows = 128 rows per CTA
constexpr int TR_LDL = 136;                     // % 16 == 8 -> conflict-free LDS.128 B fragments
constexpr int TR_SMEM = (128 * TR_LDL + 16 * 64) * 8;
template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) trsm128_rows_kernel_t(
    double* __restrict__ X, long ldx, long nrows, const double* __restrict__ L, long ldl) {
  constexpr int TR_THREADS = WARPS * 32;
  extern __shared__ __align__(16) double Ls[];
  double* Di = Ls + 128 * TR_LDL;               // [16][8][8] inverses of the diagonal 8x8 blocks (row-major)
  // lower triangle only (nothing above the diagonal is ever read), straight into shared memory: all of a thread's
  // 16-byte copies are in flight together instead of a register-staged loop of dependent L2 round trips
  for (int q = threadIdx.x; q < 128 * 64; q += TR_THREADS) {
    const int r = q >> 6, c2 = (q & 63) * 2;
    if (c2 <= r) cp_async16(Ls + r * TR_LDL + c2, L + (long)r * ldl + c2);
  }
  cp_async_commit();
  cp_async_wait<0>();
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
  const long row = (long)blockIdx.x * (WARPS * 8) + warp * 8 + gq;
  if ((long)blockIdx.x * (WARPS * 8) + warp * 8 >= nrows) return;
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

// rows: multiple of 128.  small: 32 rows per CTA
inline void trsm128_rows_launch(double* X, long ldx, long nrows, const double* L, long ldl, cudaStream_t st, bool small) {
  if (small) trsm128_rows_kernel_t<4><<<(unsigned)(nrows / 32), 128, TR_SMEM, st>>>(X, ldx, nrows, L, ldl);
  else trsm128_rows_kernel_t<16><<<(unsigned)(nrows / 128), 512, TR_SMEM, st>>>(X, ldx, nrows, L, ldl);
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
