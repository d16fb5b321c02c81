// KISS-GP path, part 5: stochastic (Chebyshev-Hutchinson) log-determinant, logdet_sample of GPML/cov/apx.m:336-391.
//
//   ldB2 = n log(a)/2 + 1/(2m) sum_i v_i' p(Bs) v_i ,  Bs = sf (I + sW K sW) - I/(1-2 dl),  p = degree-d Chebyshev
// interpolant of log on the scaled interval, evaluated in the monomial basis exactly like the reference
// (apx.m:365-372).  Probes v_i (Rademacher) and emax are explicit inputs (the reference's seed is
// mis-wired); when none are injected they are generated from opt.ldB2_cheby_seed by splitmix64 and emax by power
// iteration.
//
// Derivatives (apx.m:373-389) are bilinear forms in the Krylov vectors.  Because W is uniform (likGauss) every
// one of them collapses to GRID space: with T_j = Mx'(sW Bs^j v) and U_j = Kg T_j (both by-products of the
// recurrence) the hyper-derivative is  sum_{j,k} a_kj dK(T_{j-k}, T_{k-1})  and  sum(dW) = sum_j sf p_j /(4 m sW^2)
// sum_l (T_{j-1-l}.U_l + T_l.U_{j-1-l}), so no n-vector other than the current Krylov pair is ever stored.
// Probes are independent: with several ranks each one handles a contiguous share and the grid-sized
// accumulators are all-reduced once.
#pragma once
#include "grid_cg.cuh"

namespace kgp {

// out = c1 * bv + c2 * (Mx u) ;  part[block] = sum v .* out
template <int P>
__global__ void cheby_step_kernel(GridDesc G, const double* __restrict__ X, long n, int D, const double* __restrict__ u,
                                  const double* __restrict__ bv, const double* __restrict__ v, double c1, double c2,
                                  double* __restrict__ out, double* __restrict__ part) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    double xi[P], dd[P];
#pragma unroll
    for (int d = 0; d < P; ++d) xi[d] = X[i * D + d];
    const double mu = interp_gather_point<P, false>(G, xi, u, dd);
    const double o = fma(c2, mu, c1 * bv[i]);
    out[i] = o;
    acc = fma(v[i], o, acc);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}

// ---- the same step on the CELL-SORTED point set of the CG path (cg_prepare_sorted: encoded coordinates U, permutation) ----
// The Krylov vectors of a probe live in sorted order for the whole recurrence (dot products do not care), so the
// scatter is the segmented one of cg_scatter_kernel -- one flush per run of equal cells instead of 16 atomics per point on
// 4900 hot nodes -- and the gather reads a cell's 4^P node values once per run.  Per step at n = 1M: ~200 us -> ~35 us.
// vs[k] = v[perm[k]] (k < n), 0 on the padding
// (blockIdx.y = probe of the batch in all three kernels: vectors v + col * vstride, grid vectors g + col * gstride)
__global__ void permute_in_kernel(const double* __restrict__ v, const int* __restrict__ perm, long n, long npad, double* __restrict__ vs) {
  v += (long)blockIdx.y * n; vs += (long)blockIdx.y * npad;
  for (long k = (long)blockIdx.x * blockDim.x + threadIdx.x; k < npad; k += (long)gridDim.x * blockDim.x) vs[k] = (k < n) ? v[perm[k]] : 0.0;
}
// g0 += Mx'(sw v), v in sorted order (padding entries are zero)
template <int P>
__global__ void __launch_bounds__(256, 3) sorted_scatter_kernel(GridDesc G, const double* __restrict__ U, long npad,
    const double* __restrict__ v, double sw, double* __restrict__ g0, long vstride, long gstride) {
  constexpr int NW = CgNodes<P>::NW;
  const int lane = threadIdx.x & 31;
  v += (long)blockIdx.y * vstride; g0 += (long)blockIdx.y * gstride;
  const long nunits = npad / (32 * CG_PTS), nwarps = (long)gridDim.x * (blockDim.x >> 5);
  for (long unit = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < nunits; unit += nwarps) {
    const long base = unit * (32 * CG_PTS) + (long)lane * CG_PTS;
    double dn[CG_PTS], uu[CG_PTS * P];
    {
      const double2 d0 = *(const double2*)(v + base), d1 = *(const double2*)(v + base + 2);
      dn[0] = d0.x; dn[1] = d0.y; dn[2] = d1.x; dn[3] = d1.y;
#pragma unroll
      for (int q = 0; q < CG_PTS * P / 2; ++q) {
        const double2 t = __ldg((const double2*)(U + base * P) + q);
        uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
      }
    }
    double acc[NW];
#pragma unroll
    for (int w = 0; w < NW; ++w) acc[w] = 0.0;
    int cur = -1;
#pragma unroll
    for (int q = 0; q < CG_PTS; ++q) {
      int j[P]; double c[P][4];
#pragma unroll
      for (int dd = 0; dd < P; ++dd) decode_u(uu[q * P + dd], G.ng[dd], j[dd], c[dd]);
      const int cell = (P == 1) ? j[0] : j[0] + j[P - 1] * G.ng[0];
      if (cell != cur) {
        if (cur >= 0) {
          flush_nodes<P>(G, g0, cur, acc);
#pragma unroll
          for (int w = 0; w < NW; ++w) acc[w] = 0.0;
        }
        cur = cell;
      }
      const double vi = sw * dn[q];
      if (P == 1) {
#pragma unroll
        for (int a = 0; a < 4; ++a) acc[a] = fma(c[0][a], vi, acc[a]);
      } else {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          const double wb = c[P - 1][b] * vi;
#pragma unroll
          for (int a = 0; a < 4; ++a) acc[b * 4 + a] = fma(c[0][a], wb, acc[b * 4 + a]);
        }
      }
    }
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {              // segmented reduction over the warp (cells are contiguous)
      const int oc = __shfl_down_sync(0xffffffffu, cur, off);
      const bool take = (lane + off < 32) && (oc == cur);
#pragma unroll
      for (int w = 0; w < NW; ++w) {
        const double o = __shfl_down_sync(0xffffffffu, acc[w], off);
        if (take) acc[w] += o;
      }
    }
    const int pc = __shfl_up_sync(0xffffffffu, cur, 1);
    if (lane == 0 || pc != cur) flush_nodes<P>(G, g0, cur, acc);
  }
}

// out = c1 * bv + c2 * (Mx u) ;  part[block] = sum v .* out      (all n-vectors in sorted order, padding stays zero)
template <int P>
__global__ void __launch_bounds__(256, 3) cheby_step_sorted_kernel(GridDesc G, const double* __restrict__ U, long n, long npad,
    const double* __restrict__ ug, const double* __restrict__ bv, const double* __restrict__ v, double c1, double c2,
    double* __restrict__ out, double* __restrict__ part, long vstride, long gstride) {
  constexpr int NW = CgNodes<P>::NW;
  __shared__ double sh[32];
  double acc = 0.0;
  ug += (long)blockIdx.y * gstride; bv += (long)blockIdx.y * vstride; v += (long)blockIdx.y * vstride; out += (long)blockIdx.y * vstride;
  part += (long)blockIdx.y * gridDim.x;
  const long nunits = npad / (32 * CG_PTS), nwarps = (long)gridDim.x * (blockDim.x >> 5);
  for (long unit = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < nunits; unit += nwarps) {
    const long base = unit * (32 * CG_PTS) + (long)(threadIdx.x & 31) * CG_PTS;
    double bq[CG_PTS], vq[CG_PTS], uu[CG_PTS * P], ov[CG_PTS];
    {
      const double2 b0 = *(const double2*)(bv + base), b1 = *(const double2*)(bv + base + 2);
      const double2 v0 = *(const double2*)(v + base), v1 = *(const double2*)(v + base + 2);
      bq[0] = b0.x; bq[1] = b0.y; bq[2] = b1.x; bq[3] = b1.y; vq[0] = v0.x; vq[1] = v0.y; vq[2] = v1.x; vq[3] = v1.y;
#pragma unroll
      for (int q = 0; q < CG_PTS * P / 2; ++q) {
        const double2 t = __ldg((const double2*)(U + base * P) + q);
        uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
      }
    }
    double nv[NW];
    int cur = -1;
#pragma unroll
    for (int q = 0; q < CG_PTS; ++q) {
      int j[P]; double c[P][4];
#pragma unroll
      for (int dd = 0; dd < P; ++dd) decode_u(uu[q * P + dd], G.ng[dd], j[dd], c[dd]);
      const int cell = (P == 1) ? j[0] : j[0] + j[P - 1] * G.ng[0];
      if (cell != cur) {
        cur = cell;
        if (P == 1) {
#pragma unroll
          for (int a = 0; a < 4; ++a) nv[a] = ug[j[0] + a];
        } else {
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            const double* row = ug + (long)(j[P - 1] + b) * G.stride[1] + j[0];
#pragma unroll
            for (int a = 0; a < 4; ++a) nv[b * 4 + a] = row[a];
          }
        }
      }
      double mu = 0.0;
      if (P == 1) {
#pragma unroll
        for (int a = 0; a < 4; ++a) mu = fma(c[0][a], nv[a], mu);
      } else {
#pragma unroll
        for (int b = 0; b < 4; ++b) {
          double rb = 0.0;
#pragma unroll
          for (int a = 0; a < 4; ++a) rb = fma(c[0][a], nv[b * 4 + a], rb);
          mu = fma(c[P - 1][b], rb, mu);
        }
      }
      const double o = (base + q < n) ? fma(c2, mu, c1 * bq[q]) : 0.0;
      ov[q] = o; acc = fma(vq[q], o, acc);
    }
    *(double2*)(out + base) = make_double2(ov[0], ov[1]); *(double2*)(out + base + 2) = make_double2(ov[2], ov[3]);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}

// part[block] = sum over the block's grid-stride share of a .* b  (a one-CTA dot over 10^6..10^7 entries costs 0.1 - 1 ms)
__global__ void dot_part_kernel(const double* __restrict__ a, const double* __restrict__ b, long n, double* __restrict__ part) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) acc = fma(a[i], b[i], acc);
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}
// out[0] = sum_i S[i mod cnt? no: S is [reps][cnt]] : out = sum_r sum_k S[r][k] C[k]   (bilinear form summed over the batch)
__global__ void dot_tiled_kernel(const double* __restrict__ S, const double* __restrict__ C, int cnt, int reps, double* __restrict__ out) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (int q = threadIdx.x; q < cnt * reps; q += blockDim.x) acc = fma(S[q], C[q % cnt], acc);
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) out[0] = t;
}

// Rademacher probes from a counter-based generator: V[col][i] = +-1 from splitmix64(seed, col*n + i)
__global__ void rademacher_kernel(unsigned long long seed, long n, int m, double* __restrict__ V) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * m) return;
  unsigned long long z = seed + 0x9E3779B97F4A7C15ULL * (unsigned long long)(q + 1);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL; z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL; z ^= z >> 31;
  V[q] = (z & 1ULL) ? 1.0 : -1.0;
}

// S[a][b] = T_a . U_b   (d x d dot products of N-vectors); one CTA per (a, b)
__global__ void gram_small_kernel(const double* __restrict__ T, const double* __restrict__ U, long N, int d, double* __restrict__ S) {
  __shared__ double sh[32];
  const int a = blockIdx.x, b = blockIdx.y;
  T += (long)blockIdx.z * d * N; U += (long)blockIdx.z * d * N; S += (long)blockIdx.z * d * d;      // blockIdx.z = probe of the batch
  double acc = 0.0;
  for (long q = threadIdx.x; q < N; q += blockDim.x) acc = fma(T[(long)a * N + q], U[(long)b * N + q], acc);
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) S[a * d + b] = t;
}

// Y[s][q] = sum_u C[u][s] T[u][q]   (mix the Krylov grid vectors with the a_kj coefficients)
__global__ void mix_small_kernel(const double* __restrict__ T, const double* __restrict__ C, long N, int d, double* __restrict__ Y) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const int s = blockIdx.y;
  if (q >= N) return;
  T += (long)blockIdx.z * d * N; Y += (long)blockIdx.z * d * N;                                      // blockIdx.z = probe of the batch
  double acc = 0.0;
  for (int u = 0; u < d; ++u) acc = fma(C[u * d + s], T[(long)u * N + q], acc);
  Y[(long)s * N + q] = acc;
}

// Pm[a][b] += sum_{o,in} A[o][a][in] * B[o][b][in]   (mode_gram_kernel with accumulation)
__global__ void mode_gram_acc_kernel(const double* __restrict__ A, const double* __restrict__ B, int ng, long inner,
                                     long outer, double* __restrict__ Pm) {
  int a = blockIdx.x, b = blockIdx.y;
  __shared__ double sh[32];
  double acc = 0.0;
  const long tot = inner * outer;
  for (long q = threadIdx.x; q < tot; q += blockDim.x) {
    long ii = q % inner, o = q / inner;
    acc = fma(A[(o * ng + a) * inner + ii], B[(o * ng + b) * inner + ii], acc);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) Pm[(long)a * ng + b] += t;
}

// Lanczos three-term update: w -= a q + beta qp  (a read from the device scalar the dot-product reduction wrote);
// part[block] = sum w.w
__global__ void lanczos_update_kernel(long n, const double* __restrict__ a_dev, double beta, const double* __restrict__ q,
                                      const double* __restrict__ qp, double* __restrict__ w, double* __restrict__ part) {
  __shared__ double sh[32];
  const double a = a_dev[0];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    const double wi = w[i] - a * q[i] - beta * qp[i];
    w[i] = wi; acc = fma(wi, wi, acc);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}

// Full re-orthogonalisation of the Lanczos vector (classical Gram-Schmidt against the k stored vectors Q[t][.]):
//   part[t * gridDim.x + block] = sum over the block's contiguous rows of Q[t][i] w[i]
__global__ void multi_dot_kernel(long n, int k, const double* __restrict__ Q, const double* __restrict__ w, double* __restrict__ part) {
  __shared__ double sh[32];
  const long per = (n + gridDim.x - 1) / gridDim.x, i0 = (long)blockIdx.x * per, i1 = (i0 + per < n) ? i0 + per : n;
  for (int t = 0; t < k; ++t) {
    const double* q = Q + (long)t * n;
    double acc = 0.0;
    for (long i = i0 + threadIdx.x; i < i1; i += blockDim.x) acc = fma(q[i], w[i], acc);
    const double s = block_sum(acc, sh);
    if (threadIdx.x == 0) part[(long)t * gridDim.x + blockIdx.x] = s;
    __syncthreads();
  }
}
// h[t] = sum_b part[t * nb + b]   (one CTA per t)
__global__ void sum_rows_kernel(const double* __restrict__ part, int nb, double* __restrict__ h) {
  __shared__ double sh[32];
  const double t = sum_partials(part + (long)blockIdx.x * nb, nb, sh);
  if (threadIdx.x == 0) h[blockIdx.x] = t;
}
//   w -= sum_t h[t] Q[t][.] ;  part2[block] = sum w.w
__global__ void multi_axpy_kernel(long n, int k, const double* __restrict__ Q, const double* __restrict__ h, double* __restrict__ w,
                                  double* __restrict__ part2) {
  __shared__ double sh[32];
  __shared__ double hs[512];
  for (int t = threadIdx.x; t < k; t += blockDim.x) hs[t] = h[t];
  __syncthreads();
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    double wi = w[i];
    for (int t = 0; t < k; ++t) wi = fma(-hs[t], Q[(long)t * n + i], wi);
    w[i] = wi; acc = fma(wi, wi, acc);
  }
  const double s = block_sum(acc, sh);
  if (threadIdx.x == 0) part2[blockIdx.x] = s;
}

// y = a*x + b*y  /  copy / scale helpers on n-vectors
__global__ void vec_scale_copy_kernel(long n, double a, const double* __restrict__ x, double* __restrict__ y) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = a * x[i];
}

}  // namespace kgp
