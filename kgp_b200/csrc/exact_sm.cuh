// Exact-path kernels for the multivariate spectral-mixture covariance (GPML/cov/covSM.m:124-156, covSMfast, Q > 0):
//   K_ij = sum_q w_q exp(-1/2 sum_d T_d^2 v_dq) cos(sum_d T_d m_dq),   T = 2 pi (x_i - x_j)
// With delta = x_i - x_j the kernels work on  Vq = 4 pi^2 v  and  Mq = 2 pi m  (host-prepared, [q][DP], zero padded).
// Unlike covSE there is no single Gram matrix to put on the tensor pipe (every component q has its own metric and its
// own phase), and the cost is the exp / sincos per (i, j, q): these kernels are FP64-transcendental bound.
//   sm_build_kernel  : tiles of B = I + K / sn2 (or cross-covariance tiles for prediction)
//   sm_reduce_kernel : one pass over B^-1 per component q (blockIdx.z): the sums dKdhyp needs (covSM.m:150-153),
//                      sum B^-1 o K (apx.m:166) and -- an extension, the reference ships no dcovSM.m -- dnlZ/dX rows
//                      from covSMfast's x-derivative (covSM.m:158-168).
#pragma once
#include "exact_kernels.cuh"

namespace kgp {

constexpr int SM_MAXQ = 16;
struct SmHyp { double w[SM_MAXQ]; int Q; };

// out[i][j] = scale * K(xr_i, xc_j) (+1 where i == j and diag_add); rows/cols beyond nR/nC: identity pattern.
// CTA = 128 rows x 32 columns, thread = row; the strip goes through shared memory so that the stores are coalesced.
template <int DP>
__global__ void __launch_bounds__(128) sm_build_kernel(const double* __restrict__ xr, long nR, const double* __restrict__ xc, long nC,
    const SmHyp h, const double* __restrict__ Vq, const double* __restrict__ Mq, double scale, int diag_add, int lower_only,
    double* __restrict__ out, long ld) {
  const long i0 = (long)blockIdx.y * 128, j0 = (long)blockIdx.x * 32;
  if (lower_only && j0 > i0 + 127) return;
  __shared__ double sB[128][33];
  __shared__ double sxj[32][DP];
  __shared__ double sV[SM_MAXQ][DP], sM[SM_MAXQ][DP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int q = tid; q < h.Q * DP; q += 128) { sV[q / DP][q % DP] = Vq[q]; sM[q / DP][q % DP] = Mq[q]; }
  for (int q = tid; q < 32 * DP; q += 128) sxj[q / DP][q % DP] = xc[(j0 + q / DP) * DP + q % DP];
  const long i = i0 + tid;
  double xi[DP];
#pragma unroll
  for (int d = 0; d < DP; ++d) xi[d] = xr[i * DP + d];
  __syncthreads();
  for (int j = 0; j < 32; ++j) {
    const long jj = j0 + j;
    double val;
    if (i < nR && jj < nC) {
      double df[DP], d2[DP];
#pragma unroll
      for (int d = 0; d < DP; ++d) { df[d] = xi[d] - sxj[j][d]; d2[d] = df[d] * df[d]; }
      double k = 0.0;
      for (int q = 0; q < h.Q; ++q) {
        double t2 = 0.0, tm = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) { t2 = fma(d2[d], sV[q][d], t2); tm = fma(df[d], sM[q][d], tm); }
        k = fma(h.w[q] * exp(-0.5 * t2), cos(tm), k);
      }
      val = scale * k + ((diag_add && i == jj) ? 1.0 : 0.0);
    } else val = (diag_add && i == jj) ? 1.0 : 0.0;
    sB[tid][j] = val;
  }
  __syncthreads();
  for (int rr = 0; rr < 32; ++rr) out[(i0 + warp * 32 + rr) * ld + j0 + lane] = sB[warp * 32 + rr][lane];
}

// One component q = blockIdx.z over rows [i0, i0+128) x a range of 32-column strips of the full symmetric B^-1.
// acc (global, zeroed): [0] += sum B^-1_ij K_ij ; per q at 1 + q (1 + 2 DP): A_q = sum R H, B_dq = sum R S delta_d,
// C_dq = sum R H delta_d^2 with R = B^-1 / sn2 - alpha alpha', H = E cos, S = E sin.
// dX (n x D, zeroed) += sum_j R_ij w_q (-H Vq_d delta_d - S Mq_d).
template <int DP>
__global__ void __launch_bounds__(128) sm_reduce_kernel(const double* __restrict__ Binv, long ld, long n, const double* __restrict__ xs,
    const double* __restrict__ alpha, const SmHyp h, const double* __restrict__ Vq, const double* __restrict__ Mq, double inv_sn2,
    int ntile32_per_split, int D, double* __restrict__ acc, double* __restrict__ dX) {
  __shared__ double sB[128][33];
  __shared__ double sxj[32][DP];
  __shared__ double saj[32];
  __shared__ double sV[DP], sM[DP];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, q = blockIdx.z;
  if (tid < DP) { sV[tid] = Vq[q * DP + tid]; sM[tid] = Mq[q * DP + tid]; }
  const long i0 = (long)blockIdx.x * 128, i = i0 + tid;
  double xi[DP], gx[DP], bq[DP], cq[DP];
#pragma unroll
  for (int d = 0; d < DP; ++d) { xi[d] = xs[i * DP + d]; gx[d] = 0.0; bq[d] = 0.0; cq[d] = 0.0; }
  const double ai = alpha[i], wq = h.w[q];
  double aq = 0.0, rk = 0.0;
  const long jt0 = (long)blockIdx.y * ntile32_per_split;
  for (long jt = jt0; jt < jt0 + ntile32_per_split; ++jt) {
    const long j0 = jt * 32;
    if (j0 >= n) break;
    __syncthreads();
    for (int rr = 0; rr < 32; ++rr) sB[warp * 32 + rr][lane] = Binv[(i0 + warp * 32 + rr) * ld + j0 + lane];
    for (int t = tid; t < 32 * DP; t += 128) sxj[t / DP][t % DP] = xs[(j0 + t / DP) * DP + t % DP];
    if (tid < 32) saj[tid] = alpha[j0 + tid];
    __syncthreads();
    if (i < n) {
      const int jn = (n - j0 < 32) ? (int)(n - j0) : 32;
      for (int j = 0; j < jn; ++j) {
        double df[DP], t2 = 0.0, tm = 0.0;
#pragma unroll
        for (int d = 0; d < DP; ++d) { df[d] = xi[d] - sxj[j][d]; t2 = fma(df[d] * df[d], sV[d], t2); tm = fma(df[d], sM[d], tm); }
        double sn_, cs_; sincos(tm, &sn_, &cs_);
        const double E = exp(-0.5 * t2), H = E * cs_, S = E * sn_;
        const double bij = sB[tid][j], R = bij * inv_sn2 - ai * saj[j];
        rk = fma(bij * wq, H, rk);
        aq = fma(R, H, aq);
        const double RS = R * S, RH = R * H;
#pragma unroll
        for (int d = 0; d < DP; ++d) {
          bq[d] = fma(RS, df[d], bq[d]);
          cq[d] = fma(RH * df[d], df[d], cq[d]);
          gx[d] -= wq * (RH * sV[d] * df[d] + RS * sM[d]);
        }
      }
    }
  }
  if (i < n)
    for (int d = 0; d < D; ++d) atomicAdd(&dX[i * D + d], gx[d]);
  // block totals of the hyper-parameter sums
  double* out = acc + 1 + (long)q * (1 + 2 * DP);
  rk = warp_sum(rk); aq = warp_sum(aq);
  if (lane == 0) { atomicAdd(acc, rk); atomicAdd(out, aq); }
#pragma unroll
  for (int d = 0; d < DP; ++d) {
    const double b = warp_sum(bq[d]), c = warp_sum(cq[d]);
    if (lane == 0 && d < D) { atomicAdd(out + 1 + d, b); atomicAdd(out + 1 + DP + d, c); }
  }
}

}  // namespace kgp
