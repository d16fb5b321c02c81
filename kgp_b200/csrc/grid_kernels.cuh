// Device kernels of the KISS-GP / MSGP (infGrid) path.  Reference arithmetic reproduced:
// GPML/cov/apxGrid.m:495-539 (eqinterp, kcub, dkcub: cubic convolution weights with boundary
// snapping), :440-490 (interpgrid: first grid axis fastest), :302-329 (kronmvm),
// GPML/cov/apx.m:281-302 (conjgrad), :312-318 (logdet_fiedler), GPML/cov/apxGrid.m:631-652
// (dirder), covSEiso / GPML/cov/covSM.m:124-156 factor kernels, kgp/backend/utils/deriv_x.m.
//
// The interpolation matrix Mx (n x N, 4^p non-zeros per row) is never stored: every pass
// recomputes the 4 weights per dimension from x in registers, so an operator application
// streams only x and the vectors (HBM bound), while the grid (N doubles) lives in L2 / shared.
#pragma once
#include "exact_kernels.cuh"

namespace kgp {

constexpr int GRID_MAXP = 4;     // grid dimensions supported (KGP uses 1 and 2)

struct GridDesc {
  int p; int ng[GRID_MAXP]; long stride[GRID_MAXP]; long N;
  const double* ax[GRID_MAXP];           // device axis values
  double s0[GRID_MAXP], ds[GRID_MAXP], range[GRID_MAXP];
  int deg[GRID_MAXP];                    // opt.deg on an equispaced axis: 3 cubic, 1 linear, 0 nearest node (apxGrid.m:497-500)
};

__device__ __forceinline__ double kcub(double x) {
  x = fabs(x);
  if (x <= 1.0) return ((1.5 * x - 2.5) * x) * x + 1.0;
  if (x <= 2.0) return ((-0.5 * x + 2.5) * x - 4.0) * x + 2.0;
  return 0.0;
}
__device__ __forceinline__ double dkcub(double x) {
  double s = (x > 0.0) ? 1.0 : ((x < 0.0) ? -1.0 : 0.0);
  x = fabs(x);
  if (x <= 1.0) return s * (4.5 * x - 5.0) * x;
  if (x <= 2.0) return s * ((-1.5 * x + 5.0) * x - 4.0);
  return 0.0;
}

// eqinterp for one coordinate: first node j (0-based), weights c[4], optional derivative weights
// dc[4] (already multiplied by ns-1 as in apxGrid.m:512).
// neqinterp for one coordinate (apxGrid.m:547-557): piecewise linear on an increasing, non-equispaced axis, expressed as a
// 4-node stencil with two zero weights so that every kernel keeps its 4^p structure.  Interval ii from
// histc(t, [-inf; s(2:end-1); inf]); outside the axis the weight 1 goes to the end node and the derivative is 0.  The derivative
// weights are per unit of x (deriv_x.m:11-15 divides by w_i = 1 for such an axis: GridDesc.range = 1).
template <bool DERIV>
__device__ __forceinline__ void neqinterp1(const double* __restrict__ ax, int ns, double t, int& j, double c[4], double dc[4]) {
  int lo = 0, hi = ns - 2;
  while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (ax[mid] <= t) lo = mid; else hi = mid - 1; }
  const int ii = lo;
  double d0 = t - ax[ii], d1 = ax[ii + 1] - t;
  const bool out = (d0 < 0.0) || (d1 < 0.0);
  d0 = fmax(d0, 0.0); d1 = fmax(d1, 0.0);
  const double inv = 1.0 / (d1 + d0);
  int jb = ii - 1; if (jb < 0) jb = 0; if (jb > ns - 4) jb = ns - 4;
  const int s0 = ii - jb;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    c[q] = (q == s0) ? d1 * inv : ((q == s0 + 1) ? d0 * inv : 0.0);
    if (DERIV) dc[q] = out ? 0.0 : ((q == s0) ? -inv : ((q == s0 + 1) ? inv : 0.0));
  }
  j = jb;
}

// eqinterp with deg = 1 (linear) or 0 (nearest node), apxGrid.m:498-499 with it = -1:0, nc = 2: nodes j, j+1 with weights
// k(w), k(w-1), k(x) = max(1-|x|,0) resp. [-0.5 < x <= 0.5]; boundary rule of apxGrid.m:515-522 on two nodes.  Returned inside the
// 4-node stencil (two zero weights), like neqinterp1.
template <bool DERIV>
__device__ __forceinline__ void eqinterp_low(const double* __restrict__ ax, int ns, double s0, double ds, double t, int deg,
                                             int& j, double c[4], double dc[4]) {
  const double u = (t - s0) / ds;
  double jf = floor(u) + 1.0;                         // 1-based index of the closest smaller grid point (it(nc) = 0)
  const double w = u - jf + 1.0;
  double c0, c1, d0 = 0.0, d1 = 0.0;
  if (deg == 1) {
    c0 = fmax(1.0 - fabs(w), 0.0); c1 = fmax(1.0 - fabs(w - 1.0), 0.0);
    if (DERIV) {                                      // dk(x) = -(|x| <= 1) sign(x), times ns - 1 (apxGrid.m:512)
      const double f = (double)(ns - 1), w1 = w - 1.0;
      d0 = (fabs(w) <= 1.0) ? ((w > 0.0) ? -f : ((w < 0.0) ? f : 0.0)) : 0.0;
      d1 = (fabs(w1) <= 1.0) ? ((w1 > 0.0) ? -f : ((w1 < 0.0) ? f : 0.0)) : 0.0;
    }
  } else {
    c0 = (-0.5 < w && w <= 0.5) ? 1.0 : 0.0; c1 = (-0.5 < w - 1.0 && w - 1.0 <= 0.5) ? 1.0 : 0.0;
  }
  if (jf < 1.0) {                                     // lower boundary: weight 1 on the nearer of s(1:2) (min takes the first on a tie)
    const bool first = fabs(ax[0] - t) <= fabs(ax[1] - t);
    c0 = first ? 1.0 : 0.0; c1 = first ? 0.0 : 1.0; d0 = d1 = 0.0; jf = 1.0;
  }
  const double v = (double)(ns - 1);
  if (jf > v) {                                       // upper boundary
    const bool first = fabs(ax[ns - 2] - t) <= fabs(ax[ns - 1] - t);
    c0 = first ? 1.0 : 0.0; c1 = first ? 0.0 : 1.0; d0 = d1 = 0.0; jf = v;
  }
  const int jn = (int)jf - 1;                         // 0-based first node of the pair
  int jb = jn - 1; if (jb < 0) jb = 0; if (jb > ns - 4) jb = ns - 4;
  const int sft = jn - jb;
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    c[q] = (q == sft) ? c0 : ((q == sft + 1) ? c1 : 0.0);
    if (DERIV) dc[q] = (q == sft) ? d0 : ((q == sft + 1) ? d1 : 0.0);
  }
  j = jb;
}

template <bool DERIV>
__device__ __forceinline__ void eqinterp1(const double* __restrict__ ax, int ns, double s0, double ds, double t,
                                          int& j, double c[4], double dc[4], int deg = 3) {
  if (ds <= 0.0) { neqinterp1<DERIV>(ax, ns, t, j, c, dc); return; }       // non-equispaced axis (grid_setup sets ds = 0)
  if (deg != 3) { eqinterp_low<DERIV>(ax, ns, s0, ds, t, deg, j, c, dc); return; }
  const double u = (t - s0) / ds;
  double jf = floor(u) + 1.0;
  const double w = u - jf + 1.0;
  jf -= 1.0;                                          // j - it(nc), it(nc) = 1
  c[0] = kcub(w + 1.0); c[1] = kcub(w); c[2] = kcub(w - 1.0); c[3] = kcub(w - 2.0);
  if (DERIV) {
    const double f = (double)(ns - 1);
    dc[0] = dkcub(w + 1.0) * f; dc[1] = dkcub(w) * f; dc[2] = dkcub(w - 1.0) * f; dc[3] = dkcub(w - 2.0) * f;
  }
  if (jf < 2.0) {                                     // lower boundary: weight 1 on the nearest of s(1:4)
    int best = 0; double bd = fabs(ax[0] - t);
#pragma unroll
    for (int i = 1; i < 4; ++i) { double d = fabs(ax[i] - t); if (d < bd) { bd = d; best = i; } }
#pragma unroll
    for (int i = 0; i < 4; ++i) { c[i] = (i == best) ? 1.0 : 0.0; if (DERIV) dc[i] = 0.0; }
    jf = 1.0;
  }
  const double v = (double)(ns - 3);
  if (jf > v) {                                       // upper boundary
    int best = 0; double bd = fabs(ax[ns - 4] - t);
#pragma unroll
    for (int i = 1; i < 4; ++i) { double d = fabs(ax[ns - 4 + i] - t); if (d < bd) { bd = d; best = i; } }
#pragma unroll
    for (int i = 0; i < 4; ++i) { c[i] = (i == best) ? 1.0 : 0.0; if (DERIV) dc[i] = 0.0; }
    jf = v;
  }
  j = (int)jf - 1;
}

// ---- g += Mx' (scale .* v)   (scatter; g must be zeroed before).  One thread per point.
// N*8 bytes of dynamic shared memory when use_smem (block-private grid copy, flushed with one
// global atomic per node), else global atomics directly.
template <int P>
__global__ void interp_scatter_kernel(GridDesc G, const double* __restrict__ X, long n, int D,
                                      const double* __restrict__ v, double scale, double* __restrict__ g,
                                      int use_smem) {
  extern __shared__ double sg[];
  if (use_smem) {
    for (long q = threadIdx.x; q < G.N; q += blockDim.x) sg[q] = 0.0;
    __syncthreads();
  }
  double* tgt = use_smem ? sg : g;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    int j[P]; double c[P][4], dummy[4];
#pragma unroll
    for (int d = 0; d < P; ++d) eqinterp1<false>(G.ax[d], G.ng[d], G.s0[d], G.ds[d], X[i * D + d], j[d], c[d], dummy, G.deg[d]);
    const double vi = v[i] * scale;
    if (P == 1) {
#pragma unroll
      for (int a = 0; a < 4; ++a) if (c[0][a] != 0.0) atomicAdd(tgt + j[0] + a, c[0][a] * vi);
    } else if (P == 2) {
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const double wb = c[1][b] * vi;
        if (wb == 0.0) continue;
        double* row = tgt + (long)(j[1] + b) * G.stride[1] + j[0];
#pragma unroll
        for (int a = 0; a < 4; ++a) if (c[0][a] != 0.0) atomicAdd(row + a, c[0][a] * wb);
      }
    } else {
      int tot = 1;
      for (int d = 0; d < P; ++d) tot *= 4;
      for (int q = 0; q < tot; ++q) {
        double w = vi; long idx = 0; int r = q;
        for (int d = 0; d < P; ++d) { int a = r & 3; r >>= 2; w *= c[d][a]; idx += (long)(j[d] + a) * G.stride[d]; }
        if (w != 0.0) atomicAdd(tgt + idx, w);
      }
    }
  }
  if (use_smem) {
    __syncthreads();
    for (long q = threadIdx.x; q < G.N; q += blockDim.x) if (sg[q] != 0.0) atomicAdd(g + q, sg[q]);
  }
}

// value of (Mx u)_i, and optionally of (dMx_k u)_i for every k, for one point
template <int P, bool DERIV>
__device__ __forceinline__ double interp_gather_point(const GridDesc& G, const double* __restrict__ xi,
                                                      const double* __restrict__ u, double dout[P]) {
  int j[P]; double c[P][4], dc[P][4];
#pragma unroll
  for (int d = 0; d < P; ++d) eqinterp1<DERIV>(G.ax[d], G.ng[d], G.s0[d], G.ds[d], xi[d], j[d], c[d], dc[d], G.deg[d]);
  double acc = 0.0;
  if (DERIV) {
#pragma unroll
    for (int d = 0; d < P; ++d) dout[d] = 0.0;
  }
  int tot = 1;
#pragma unroll
  for (int d = 0; d < P; ++d) tot *= 4;
  for (int q = 0; q < tot; ++q) {
    long idx = 0; int r = q; int a[P];
#pragma unroll
    for (int d = 0; d < P; ++d) { a[d] = r & 3; r >>= 2; idx += (long)(j[d] + a[d]) * G.stride[d]; }
    const double uv = u[idx];
    double w = 1.0;
#pragma unroll
    for (int d = 0; d < P; ++d) w *= c[d][a[d]];
    acc = fma(w, uv, acc);
    if (DERIV) {
#pragma unroll
      for (int k = 0; k < P; ++k) {
        double wk = 1.0;
#pragma unroll
        for (int d = 0; d < P; ++d) wk *= (d == k) ? dc[d][a[d]] : c[d][a[d]];
        dout[k] = fma(wk, uv, dout[k]);
      }
    }
  }
  return acc;
}

// z = sw2 * (Mx u) + d      (B d = d + sW Mx Kg Mx' sW d, with u = Kg Mx' (sW d), sw2 = sW);
// also per-block partial of d.z  -> part[blockIdx.x]
template <int P>
__global__ void interp_gather_bz_kernel(GridDesc G, const double* __restrict__ X, long n, int D,
                                        const double* __restrict__ u, double sw, const double* __restrict__ dvec,
                                        double* __restrict__ z, double* __restrict__ part) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    double xi[P], dd[P];
#pragma unroll
    for (int d = 0; d < P; ++d) xi[d] = X[i * D + d];
    const double mu = interp_gather_point<P, false>(G, xi, u, dd);
    const double di = dvec ? dvec[i] : 0.0;
    const double zi = sw * mu + di;
    z[i] = zi;
    acc = fma(di, zi, acc);
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0 && part) part[blockIdx.x] = t;
}

// out[i][k] = -alpha_i * ( (dMx_k beta)_i / range_k + dmean_k )     (deriv_x.m:22-23)
template <int P>
__global__ void interp_dx_kernel(GridDesc G, const double* __restrict__ X, long n, int D,
                                 const double* __restrict__ beta, const double* __restrict__ alpha,
                                 const double* __restrict__ dmean /* D values or null */, double* __restrict__ out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double xi[P], dd[P];
#pragma unroll
  for (int d = 0; d < P; ++d) xi[d] = X[i * D + d];
  interp_gather_point<P, true>(G, xi, beta, dd);
#pragma unroll
  for (int k = 0; k < P; ++k) out[i * D + k] = -alpha[i] * (dd[k] / G.range[k] + (dmean ? dmean[k] : 0.0));
}

// ---- Kronecker factor application along one grid mode:
// out[o][a][in] = sum_b K[a][b] * in_[o][b][in],  inner = prod ng[<mode], outer = prod ng[>mode]
__global__ void mode_product_kernel(const double* __restrict__ K, int ng, const double* __restrict__ in,
                                    double* __restrict__ out, long inner, long outer) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long tot = inner * ng * outer;
  if (q >= tot) return;
  long ii = q % inner; long r = q / inner; int a = (int)(r % ng); long o = r / ng;
  const double* src = in + o * ng * inner + ii;
  const double* krow = K + (long)a * ng;
  double acc = 0.0;
  for (int b = 0; b < ng; ++b) acc = fma(krow[b], src[(long)b * inner], acc);
  out[q] = acc;
}

// P[a][b] = sum_{o,in} A[o][a][in] * B[o][b][in]   (contract every mode but one; ng x ng output)
__global__ void mode_gram_kernel(const double* __restrict__ A, const double* __restrict__ B, int ng, long inner,
                                 long outer, double* __restrict__ Pm) {
  int a = blockIdx.x, b = blockIdx.y;
  __shared__ double sh[32];
  double acc = 0.0;
  const long tot = inner * outer;
  for (long q = threadIdx.x; q < tot; q += blockDim.x) {
    long ii = q % inner, o = q / inner;
    acc = fma(A[(o * ng + a) * inner + ii], B[(o * ng + b) * inner + ii], acc);
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) Pm[(long)a * ng + b] = t;
}

// ---- factor covariance matrices on an axis --------------------------------------------------
// kind 0: covSEiso, hyp = [log ell, log sf]; kind 2: covSM with Q components, hyp = [log w; log m; log s]
__global__ void factor_build_kernel(int kind, int Q, const double* __restrict__ hyp, const double* __restrict__ ax,
                                    int ng, double axmean, double* __restrict__ K) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)ng * ng) return;
  int a = (int)(q / ng), b = (int)(q % ng);
  double val;
  if (kind == 0) {
    const double il = exp(-hyp[0]), sf2 = exp(2.0 * hyp[1]);
    const double xa = (ax[a] - axmean) * il, xb = (ax[b] - axmean) * il;        // covMaha.m:138-143
    const double d2 = fmax(xa * xa + (xb * xb - 2.0 * xa * xb), 0.0);
    val = sf2 * exp(-0.5 * d2);
  } else {
    const double T = 2.0 * M_PI * (ax[a] - ax[b]);
    val = 0.0;
    for (int c = 0; c < Q; ++c) {
      const double w = exp(hyp[c]), m = exp(hyp[Q + c]), v = exp(2.0 * hyp[2 * Q + c]);
      val += w * exp(-0.5 * T * T * v) * cos(T * m);
    }
  }
  K[q] = val;
}

// dhyp[t] (+)= sign * sum_ab Pm[a][b] * dK_t[a][b]; one block per hyper-parameter t of the factor
__global__ void factor_dirder_kernel(int kind, int Q, const double* __restrict__ hyp, const double* __restrict__ ax,
                                     int ng, double axmean, const double* __restrict__ Pm, double sign,
                                     double* __restrict__ dhyp, int accumulate) {
  __shared__ double sh[32];
  const int t = blockIdx.x;
  double acc = 0.0;
  for (long q = threadIdx.x; q < (long)ng * ng; q += blockDim.x) {
    int a = (int)(q / ng), b = (int)(q % ng);
    double d;
    if (kind == 0) {
      const double il = exp(-hyp[0]), sf2 = exp(2.0 * hyp[1]);
      const double xa = (ax[a] - axmean) * il, xb = (ax[b] - axmean) * il;
      const double d2 = fmax(xa * xa + (xb * xb - 2.0 * xa * xb), 0.0);
      const double k = sf2 * exp(-0.5 * d2);
      d = (t == 0) ? k * d2 : 2.0 * k;
    } else {
      const int c = t % Q, which = t / Q;
      const double T = 2.0 * M_PI * (ax[a] - ax[b]);
      const double w = exp(hyp[c]), m = exp(hyp[Q + c]), v = exp(2.0 * hyp[2 * Q + c]);
      const double E = exp(-0.5 * T * T * v);
      if (which == 0) d = E * cos(T * m) * w;
      else if (which == 1) d = -E * sin(T * m) * T * w * m;
      else d = -E * cos(T * m) * T * T * w * v;
    }
    acc = fma(Pm[q], d, acc);
  }
  double s = block_sum(acc, sh);
  if (threadIdx.x == 0) dhyp[t] = (accumulate ? dhyp[t] : 0.0) + sign * s;
}

}  // namespace kgp
