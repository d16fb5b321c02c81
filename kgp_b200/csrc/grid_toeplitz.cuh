// Toeplitz grid factors (equispaced axis handed over as a nested cell xg{i} = {axis}; GPML/cov/apxGrid.m:258-281).
// For such a factor GPML keeps the covariance as a function of the lag only, kii(l) = k((l-1) h, 0), and
//   * multiplies by FFT through a Strang circulant embedding (bttbmvmsymfft, apxGrid.m:335-346) -- exactly the Toeplitz
//     product, so here the MVM simply runs on the same factor matrix as the dense mode (K_ab = k((a-b) h));
//   * takes the Fiedler eigenvalues from a WHITTLE circulant embedding with s = 3 (circ, apxGrid.m:617-627; eigkron
//     :576-580; apx.m:236): c_w[j] = sum_{r=-s}^{s-1} k((j + r n) h), e = max(real(fft(c_w)), 0).  These are not the
//     eigenvalues of the Toeplitz matrix: ldB2 and its derivative differ from dense-factor mode;
//   * differentiates the bound as a sum over the 2 s n Whittle lags (apx.m:256-266):
//     dhyp = sum_l real(fft(z))[l mod n] dk(l h)/dhyp,  l = -s n .. s n - 1.
// The DFTs are evaluated directly (n^2 sincospi per factor, n <= a few thousand): exact to rounding for any n, no
// power-of-two restriction, and far off the critical path (the eigen-decomposition it replaces is O(n^3)).
#pragma once
#include "grid_kernels.cuh"

namespace kgp {

constexpr int WHITTLE_S = 3;

// k(tau) of a 1-D factor and its derivative w.r.t. hyper-parameter t (covSEiso: [log ell, log sf]; covSM.m:124-153 with
// D = 1: [log w (Q); log m (Q); log sqrt v (Q)])
__device__ __forceinline__ double factor_k_lag(int kind, int Q, const double* __restrict__ hyp, double tau) {
  if (kind == 0) {
    const double il = exp(-hyp[0]), sf2 = exp(2.0 * hyp[1]), u = tau * il;
    return sf2 * exp(-0.5 * u * u);
  }
  const double T = 2.0 * M_PI * tau;
  double val = 0.0;
  for (int c = 0; c < Q; ++c) {
    const double w = exp(hyp[c]), m = exp(hyp[Q + c]), v = exp(2.0 * hyp[2 * Q + c]);
    val += w * exp(-0.5 * T * T * v) * cos(T * m);
  }
  return val;
}
__device__ __forceinline__ double factor_dk_lag(int kind, int Q, const double* __restrict__ hyp, double tau, int t) {
  if (kind == 0) {
    const double il = exp(-hyp[0]), sf2 = exp(2.0 * hyp[1]), u = tau * il, k = sf2 * exp(-0.5 * u * u);
    return t == 0 ? k * u * u : 2.0 * k;
  }
  const int c = t % Q, which = t / Q;
  const double T = 2.0 * M_PI * tau;
  const double w = exp(hyp[c]), m = exp(hyp[Q + c]), v = exp(2.0 * hyp[2 * Q + c]);
  const double E = exp(-0.5 * T * T * v);
  if (which == 0) return w * E * cos(T * m);
  if (which == 1) return -w * E * sin(T * m) * T * m;
  return -w * E * cos(T * m) * T * T * v;
}

// e[t] = max(real(fft(c_w))[t], 0): one CTA, dynamic shared memory n doubles
__global__ void whittle_eig_kernel(int kind, int Q, const double* __restrict__ hyp, int n, double h, double* __restrict__ e) {
  extern __shared__ double cw[];
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    double s = 0.0;
    for (int r = -WHITTLE_S; r < WHITTLE_S; ++r) s += factor_k_lag(kind, Q, hyp, (double)(j + r * n) * h);
    cw[j] = s;
  }
  __syncthreads();
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    double acc = 0.0;
    for (int j = 0; j < n; ++j) {
      const long jt = ((long)j * t) % n;
      acc = fma(cw[j], cospi(2.0 * (double)jt / (double)n), acc);
    }
    e[t] = fmax(acc, 0.0);
  }
}

// dhyp[t] = sum over the 2 s n Whittle lags of real(fft(z))[l mod n] dk(l h)/dhyp_t : one CTA per hyper-parameter t
__global__ void toep_fiedler_dirder_kernel(int kind, int Q, const double* __restrict__ hyp, int n, double h,
                                           const double* __restrict__ z, double* __restrict__ dhyp) {
  extern __shared__ double Fz[];
  __shared__ double sh[32];
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    double acc = 0.0;
    for (int j = 0; j < n; ++j) {
      const long jt = ((long)j * t) % n;
      acc = fma(z[j], cospi(2.0 * (double)jt / (double)n), acc);
    }
    Fz[t] = acc;
  }
  __syncthreads();
  const int t = blockIdx.x;
  double acc = 0.0;
  for (int q = threadIdx.x; q < 2 * WHITTLE_S * n; q += blockDim.x) {
    const int l = q - WHITTLE_S * n;                       // lag index, -s n .. s n - 1
    acc = fma(Fz[q % n], factor_dk_lag(kind, Q, hyp, (double)l * h, t), acc);
  }
  const double tot = block_sum(acc, sh);
  if (threadIdx.x == 0) dhyp[t] = tot;
}

}  // namespace kgp
