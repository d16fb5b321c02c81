// Exact-GP path, part 2: solves with Z = L^{-T}, inverse finalisation, the fused
// gradient reduction and prediction helpers.  See exact_kernels.cuh for citations.
#pragma once
#include "exact_kernels.cuh"

namespace kgp {

// ---------------------------------------------------------------------------
// t = Z^T b  with Z upper triangular (row-major, zeros below the diagonal inside
// diagonal blocks).  Thread owns output column i; rows are split in chunks of 512
// and reduced in a second pass (deterministic).
// ---------------------------------------------------------------------------
__global__ void gemv_t_upper_kernel(const double* __restrict__ Z, long ld, long np,
                                    const double* __restrict__ b, double* __restrict__ part, int chunk = 512) {
  const long i = (long)blockIdx.x * 128 + threadIdx.x;
  const long c0 = (long)blockIdx.y * chunk;
  double acc = 0.0;
  if (c0 <= (long)blockIdx.x * 128 + 127) {
    long c1 = c0 + chunk; if (c1 > np) c1 = np;
    long cmax = (long)blockIdx.x * 128 + 128; if (c1 > cmax) c1 = cmax;
    for (long c = c0; c < c1; ++c) acc += Z[c * ld + i] * b[c];     // Z[c][i]==0 for c>i inside the diag block
  }
  part[(long)blockIdx.y * np + i] = acc;
}
__global__ void sum_parts_kernel(const double* __restrict__ part, long np, int nparts, double scale,
                                 double* __restrict__ out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= np) return;
  double a = 0.0;
  for (int p = 0; p < nparts; ++p) a += part[(long)p * np + i];
  out[i] = a * scale;
}
// out[c] = scale * sum_{k>=c} Z[c][k] t[k]   (warp per row)
__global__ void gemv_n_upper_kernel(const double* __restrict__ Z, long ld, long np,
                                    const double* __restrict__ t, double scale, double* __restrict__ out) {
  const long c = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (c >= np) return;
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  for (long k = (c & ~31L) + lane; k < np; k += 32)
    if (k >= c) acc += Z[c * ld + k] * t[k];
  acc = warp_sum(acc);
  if (lane == 0) out[c] = acc * scale;
}
// out[r] = sum_k A[r][k] v[k]  (dense rows, warp per row) ; used by predict (Ks^T alpha)
__global__ void gemv_rows_kernel(const double* __restrict__ A, long ld, long nrows, long ncols,
                                 const double* __restrict__ v, double* __restrict__ out) {
  const long r = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= nrows) return;
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  for (long k = lane; k < ncols; k += 32) acc += A[r * ld + k] * v[k];
  acc = warp_sum(acc);
  if (lane == 0) out[r] = acc;
}
// out[r] = sum_k A[r][k]^2
__global__ void rowsumsq_kernel(const double* __restrict__ A, long ld, long nrows, long ncols,
                                double* __restrict__ out) {
  const long r = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= nrows) return;
  const int lane = threadIdx.x & 31;
  double acc = 0.0;
  for (long k = lane; k < ncols; k += 32) { double a = A[r * ld + k]; acc += a * a; }
  acc = warp_sum(acc);
  if (lane == 0) out[r] = acc;
}
// dot products  out[0] = a.b (single CTA, deterministic)
__global__ void dot_kernel(const double* __restrict__ a, const double* __restrict__ b, long n,
                           double* __restrict__ out) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) acc += a[i] * b[i];
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[0] = t;
}

// ---------------------------------------------------------------------------
// After Z Z^T: W holds B^{-1} in the lower tiles of off-diagonal nb-blocks, Dws the
// diagonal nb-blocks.  Make W the full symmetric B^{-1} (upper := mirror of lower).
// grid (np/32, np/32), block (32, 8).
// ---------------------------------------------------------------------------
__global__ void mirror_finalize_kernel(double* __restrict__ W, long ld, long np,
                                       const double* __restrict__ Dws, int nb) {
  __shared__ double t[32][33];
  const int bj = blockIdx.x, bi = blockIdx.y;
  if (bj > bi) return;
  const long i0 = (long)bi * 32, j0 = (long)bj * 32;
  const bool diagblk = (i0 / nb) == (j0 / nb);
  for (int r = threadIdx.y; r < 32; r += 8) {
    long i = i0 + r, j = j0 + threadIdx.x;
    double v;
    if (diagblk) {
      long blk = i0 / nb;
      v = Dws[blk * (long)nb * nb + (i - blk * nb) * nb + (j - blk * nb)];
    } else v = W[i * ld + j];
    t[r][threadIdx.x] = v;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += 8) {
    long i = i0 + r, j = j0 + threadIdx.x;
    if (bi != bj) {
      if (diagblk) W[i * ld + j] = t[r][threadIdx.x];
      W[(j0 + r) * ld + i0 + threadIdx.x] = t[threadIdx.x][r];            // transposed tile
    } else {
      // diagonal 32x32 tile: take the lower part, mirror it
      double v = (threadIdx.x <= r) ? t[r][threadIdx.x] : t[threadIdx.x][r];
      W[i * ld + j] = v;
    }
  }
}

// ---------------------------------------------------------------------------
// Fused gradient pass over Binv = B^{-1} (full symmetric, np x np):
//   k_ij = sf2 exp(-|xs_i-xs_j|^2/2),  G_ij = (Binv_ij/sn2 - alpha_i alpha_j) k_ij
//   per row i:  r_i = sum_j G_ij,  rd2_i = sum_j G_ij D2_ij,  rk_i = sum_j Binv_ij k_ij,
//               gd_id = sum_j G_ij (xs_id - xs_jd)
// i.e. everything infGaussLik.m:31-32 / apx.m:166,170-173 / covMaha.m:104,169-175 /
// dlikExact.m:19-25 need, without the n x n x D tensor.  Thread owns one row;
// grid (np/128, nsplit); partial[split][row][DP+3].
// ---------------------------------------------------------------------------
template <int DP>
__global__ void __launch_bounds__(128) reduce_grad_kernel(
    const double* __restrict__ Binv, long ld, long n, long np, const double* __restrict__ xs,
    const double* __restrict__ alpha, double sf2, double inv_sn2, int ntile32_per_split,
    double* __restrict__ partial) {
  __shared__ double sB[128][33];
  __shared__ double sxj[32][DP];
  __shared__ double saj[32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long i0 = (long)blockIdx.x * 128, i = i0 + tid;
  double xi[DP], gd[DP];
#pragma unroll
  for (int d = 0; d < DP; ++d) { xi[d] = xs[i * DP + d]; gd[d] = 0.0; }
  const double ai = alpha[i];
  double r = 0.0, rd2 = 0.0, rk = 0.0;
  const long jt0 = (long)blockIdx.y * ntile32_per_split;
  for (long jt = jt0; jt < jt0 + ntile32_per_split; ++jt) {
    const long j0 = jt * 32;
    if (j0 >= n) break;
    __syncthreads();
    for (int rr = 0; rr < 32; ++rr) sB[warp * 32 + rr][lane] = Binv[(i0 + warp * 32 + rr) * ld + j0 + lane];
    for (int q = tid; q < 32 * DP; q += 128) sxj[q / DP][q % DP] = xs[(j0 + q / DP) * DP + q % DP];
    if (tid < 32) saj[tid] = alpha[j0 + tid];
    __syncthreads();
    if (i < n) {
      int jn = (n - j0 < 32) ? (int)(n - j0) : 32;
      for (int j = 0; j < jn; ++j) {
        double df[DP], d2 = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) { df[d] = xi[d] - sxj[j][d]; d2 = fma(df[d], df[d], d2); }
        double k = sf2 * exp(-0.5 * d2);
        double bij = sB[tid][j];
        double g = (bij * inv_sn2 - ai * saj[j]) * k;
        r += g; rd2 = fma(g, d2, rd2); rk = fma(bij, k, rk);
#pragma unroll
        for (int d = 0; d < DP; ++d) gd[d] = fma(g, df[d], gd[d]);
      }
    }
  }
  double* p = partial + ((long)blockIdx.y * np + i) * (DP + 3);
  p[0] = r; p[1] = rd2; p[2] = rk;
#pragma unroll
  for (int d = 0; d < DP; ++d) p[3 + d] = gd[d];
}

// sum the splits; rows[i][0..DP+2] ; also diagB[i] = Binv_ii
__global__ void reduce_rows_kernel(const double* __restrict__ partial, long np, int nsplit, int W,
                                   double* __restrict__ rows) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= np * W) return;
  double a = 0.0;
  for (int s = 0; s < nsplit; ++s) a += partial[(long)s * np * W + q];
  rows[q] = a;
}

// Scalars from the row sums (one CTA per scalar, blockIdx.x = c):  out[0]=sum r, out[1]=sum rd2, out[2]=sum rk,
// out[3]=alpha.alpha, out[4+d] = sum_i xs_id gd_id  (d < DP)
__global__ void grad_scalars_kernel(const double* __restrict__ rows, long n, int DP,
                                    const double* __restrict__ xs, const double* __restrict__ alpha,
                                    double* __restrict__ out) {
  __shared__ double sh[32];
  const int W = DP + 3;
  {
    const int c = blockIdx.x;
    double a = 0.0;
    for (long i = threadIdx.x; i < n; i += blockDim.x) {
      double v;
      if (c < 3) v = rows[i * W + c];
      else if (c == 3) v = alpha[i] * alpha[i];
      else v = xs[i * DP + (c - 4)] * rows[i * W + 3 + (c - 4)];
      a += v;
    }
    double t = block_sum(a, sh);
    if (threadIdx.x == 0) out[c] = t;
  }
}

// dX[i][d] = -gd_id * f_d   (f_d = 1/ell_d, or ell_d/ell_1^2 in dcovSEard.m's form)
__global__ void dx_from_rows_kernel(const double* __restrict__ rows, long n, int D, int DP,
                                    const double* __restrict__ fac, double* __restrict__ dX) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * D) return;
  long i = q / D; int d = (int)(q % D);
  dX[q] = -rows[i * (DP + 3) + 3 + d] * fac[d];
}

// dnlZ.mean pieces: out[d] = -sum_i X[i][d] alpha_i (d<D, meanLinear), out[D] = -sum alpha (meanConst)
__global__ void mean_grad_kernel(const double* __restrict__ X, const double* __restrict__ alpha, long n, int D,
                                 double* __restrict__ out) {
  __shared__ double sh[32];
  int c = blockIdx.x;
  double a = 0.0;
  for (long i = threadIdx.x; i < n; i += blockDim.x) a += (c < D ? X[i * D + c] : 1.0) * alpha[i];
  double t = block_sum(a, sh);
  if (threadIdx.x == 0) out[c] = -t;
}

// predict epilogue  (gp.m:155-164, likGauss.m:34-39)
__global__ void predict_finish_kernel(const double* __restrict__ Xs, long ns, int D, int mean_mask,
                                      const double* __restrict__ hyp_mean, const double* __restrict__ kalpha,
                                      const double* __restrict__ vsq, double kss, double sn2, int have_var,
                                      double* __restrict__ ymu, double* __restrict__ ys2,
                                      double* __restrict__ fmu, double* __restrict__ fs2) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= ns) return;
  double m = 0.0; int o = 0;
  if (mean_mask & 1) { for (int d = 0; d < D; ++d) m += Xs[i * D + d] * hyp_mean[d]; o = D; }
  if (mean_mask & 2) m += hyp_mean[o];
  double f = m + kalpha[i];
  double v = have_var ? fmax(kss - vsq[i], 0.0) : 0.0;
  fmu[i] = f; ymu[i] = f; fs2[i] = v; ys2[i] = v + sn2;
}

// scale rows: A[r][c] *= s  over nrows x ncols
__global__ void scale_matrix_kernel(double* __restrict__ A, long ld, long nrows, long ncols, double s) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= nrows * ncols) return;
  long r = q / ncols, c = q % ncols;
  A[r * ld + c] *= s;
}

}  // namespace kgp
