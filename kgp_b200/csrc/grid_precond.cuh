// KISS-GP path, part 6: grid-space (Woodbury) preconditioner for the CG of apx.m:281-302 -- SURVEY a21(i), no
// reference counterpart.  With uniform sW = s and K = F F' (F = V E^1/2 from the factor eigen-decompositions the
// Fiedler bound computes anyway):
//     B^-1 = (I + s^2 M K M')^-1 = I - s^2 M F C^-1 F' M' ,   C = I + s^2 E^1/2 V'(M'M)V E^1/2   (N x N, SPD)
// G = M'M depends only on (data, grid): accumulated once per update_data from the cell-sorted points; per
// hyper-parameter set C is formed with four small mode products and factored by the exact path's blocked FP64
// Cholesky + triangular inverse (DMMA).  Used as the preconditioner of the SAME conjugate-gradient iteration, it
// leaves the reference's stopping rule (relres < cg_tol) in charge and only changes the iteration count
// (thousands -> 2-3); the un-preconditioned solve stays the default and the parity reference.
// Rank truncation: Kronecker modes k with s^2 e_k n/N <= tau (default 1e-3) perturb B by a relative O(tau) and are
// left out: C is built on the r kept modes only (the exact inverse of I + s^2 M F_r F_r' M'), which keeps the
// preconditioned condition number at 1 + O(tau) while r is typically a few hundred of the N modes of a smooth kernel
// (bench problem: 322 of 4900) -- the N^3 factorisation becomes negligible.
#pragma once
#include "grid_cheby.cuh"

namespace kgp {

// G += sum_i m_i m_i'   over sorted points (U encoded coordinates, cells contiguous).  One warp per 32 points:
// lanes decode one point each, then the warp walks the 32 points; every lane owns 8 (2-D: 256/32) or, for 1-D,
// lanes 0-15 own one of the 16 entries of the (4^P x 4^P) block, flushed with atomics when the cell changes.
template <int P>
__global__ void __launch_bounds__(256) gram_accum_kernel(GridDesc G, const double* __restrict__ U, long n, long npad,
                                                        double* __restrict__ Gm) {
  constexpr int NW = (P == 1) ? 4 : 16, EPL = (P == 1) ? 1 : 8;         // node weights per point, entries per lane
  const int lane = threadIdx.x & 31;
  const long warp = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarps = ((long)gridDim.x * blockDim.x) >> 5;
  const long N = G.N;
  for (long base = warp * 32; base < npad; base += nwarps * 32) {
    const long i = base + lane;
    int j[P]; double c[P][4];
#pragma unroll
    for (int d = 0; d < P; ++d) decode_u(U[i * P + d], G.ng[d], j[d], c[d]);
    const bool valid = i < n;
    const int off = (P == 1) ? j[0] : j[P - 1] * G.ng[0] + j[0];          // node offset of the cell's first node
    const int er = (P == 1) ? (lane >> 2) : (lane >> 1);                   // row entry of this lane inside the block
    const int ec0 = (P == 1) ? (lane & 3) : (lane & 1) * 8;                // first column entry
    double acc[EPL];
#pragma unroll
    for (int t = 0; t < EPL; ++t) acc[t] = 0.0;
    int cur = __shfl_sync(0xffffffffu, off, 0);
    auto flush = [&](int o) {
      if (P == 1 && lane >= 16) return;
      const long row = (P == 1) ? (long)(o + er) : (long)(o + (er >> 2) * G.ng[0] + (er & 3));
#pragma unroll
      for (int t = 0; t < EPL; ++t) {
        const int ec = ec0 + t;
        const long col = (P == 1) ? (long)(o + ec) : (long)(o + (ec >> 2) * G.ng[0] + (ec & 3));
        if (acc[t] != 0.0) atomicAdd(Gm + row * N + col, acc[t]);
        acc[t] = 0.0;
      }
    };
    for (int q = 0; q < 32; ++q) {
      const int oq = __shfl_sync(0xffffffffu, off, q);
      const int vq = __shfl_sync(0xffffffffu, (int)valid, q);
      double w[NW];
      if (P == 1) {
#pragma unroll
        for (int a = 0; a < 4; ++a) w[a] = __shfl_sync(0xffffffffu, c[0][a], q);
      } else {
        double c0[4], c1[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { c0[a] = __shfl_sync(0xffffffffu, c[0][a], q); c1[a] = __shfl_sync(0xffffffffu, c[P - 1][a], q); }
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
          for (int a = 0; a < 4; ++a) w[b * 4 + a] = c1[b] * c0[a];
      }
      if (!vq) continue;
      if (oq != cur) { flush(cur); cur = oq; }
      double wr = 0.0;
#pragma unroll
      for (int e = 0; e < NW; ++e) if (e == er) wr = w[e];                 // (static indexing keeps w[] in registers)
#pragma unroll
      for (int t = 0; t < EPL; ++t) {
        double wc = 0.0;
#pragma unroll
        for (int e = 0; e < NW; ++e) if (e == ec0 + t) wc = w[e];
        acc[t] = fma(wr, wc, acc[t]);
      }
    }
    flush(cur);
  }
}

// Register-blocked mode product for the N x N arrays of the preconditioner set-up:
//   out[o][a][ii] = sum_j Mat[a][j] in[o][j][ii]; a thread forms R outputs a for one (o, ii), so every loaded input
// value feeds R FMAs; Mat sits in shared memory (rows padded with zeros to a multiple of R).
template <int R>
__global__ void __launch_bounds__(256) mode_product_rb_kernel(const double* __restrict__ Mat, int ng, const double* __restrict__ in,
                                                             double* __restrict__ out, long inner, long outer) {
  extern __shared__ double smat[];
  const int nab = (ng + R - 1) / R;
  for (int q = threadIdx.x; q < nab * R * ng; q += blockDim.x) smat[q] = (q < ng * ng) ? Mat[q] : 0.0;
  __syncthreads();
  const long total = outer * nab * inner;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (long)gridDim.x * blockDim.x) {
    const long ii = q % inner, r = q / inner; const int ab = (int)(r % nab); const long o = r / nab;
    const double* src = in + o * ng * inner + ii;
    const double* m0 = smat + (long)ab * R * ng;
    double acc[R];
#pragma unroll
    for (int t = 0; t < R; ++t) acc[t] = 0.0;
#pragma unroll 2
    for (int j = 0; j < ng; ++j) {
      const double v = src[(long)j * inner];
#pragma unroll
      for (int t = 0; t < R; ++t) acc[t] = fma(m0[t * ng + j], v, acc[t]);
    }
    double* dst = out + (o * ng + (long)ab * R) * inner + ii;
#pragma unroll
    for (int t = 0; t < R; ++t) if (ab * R + t < ng) dst[(long)t * inner] = acc[t];
  }
}

// se[k] = sqrt(kron eigenvalue k)  (first axis fastest)
__global__ void sqrt_kron_eig_kernel(GridDesc G, const double* __restrict__ ee, const int* __restrict__ eoff, double* __restrict__ se) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= G.N) return;
  long r = q; double v = 1.0;
  for (int d = 0; d < G.p; ++d) { v *= ee[eoff[d] + (int)(r % G.ng[d])]; r /= G.ng[d]; }
  se[q] = sqrt(fmax(v, 0.0));
}

// C = I + s2 * se se' .* H  into the lower triangle of the (np x np, identity padded) Cholesky workspace
__global__ void assemble_c_kernel(const double* __restrict__ H, const double* __restrict__ se, double s2, long N, long np,
                                  double* __restrict__ C) {
  const long col = (long)blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y;
  if (col >= np || col > row) return;
  double v = (row == col) ? 1.0 : 0.0;
  if (row < N && col < N) v += s2 * se[row] * se[col] * 0.5 * (H[row * N + col] + H[col * N + row]);
  C[row * np + col] = v;
}

// C over the SELECTED Kronecker modes sel[0..r): C = I + s2 * se se' .* H restricted to sel (lower triangle, identity padded)
__global__ void assemble_c_sel_kernel(const double* __restrict__ H, const double* __restrict__ se, double s2, long N,
                                      const int* __restrict__ sel, long r, long np, double* __restrict__ C) {
  const long col = (long)blockIdx.x * blockDim.x + threadIdx.x, row = blockIdx.y;
  if (col >= np || col > row) return;
  double v = (row == col) ? 1.0 : 0.0;
  if (row < r && col < r) {
    const long a = sel[row], b = sel[col];
    v += s2 * se[a] * se[b] * 0.5 * (H[a * N + b] + H[b * N + a]);
  }
  C[row * np + col] = v;
}
// out[q] = t[sel[q]] * se[sel[q]] for q < r, 0 up to np
__global__ void gather_sel_kernel(const double* __restrict__ t, const double* __restrict__ se, const int* __restrict__ sel, long r, long np,
                                  double* __restrict__ out) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < np) out[q] = (q < r) ? t[sel[q]] * se[sel[q]] : 0.0;
}
// out (N, zeroed before) [sel[q]] = a[q] * se[sel[q]]
__global__ void scatter_sel_kernel(const double* __restrict__ a, const double* __restrict__ se, const int* __restrict__ sel, long r,
                                   double* __restrict__ out) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < r) out[sel[q]] = a[q] * se[sel[q]];
}

// out[k] = a[k] * b[k] for k < N, 0 for N <= k < np
__global__ void mul_pad_kernel(const double* __restrict__ a, const double* __restrict__ b, long N, long np, double* __restrict__ out) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < np) out[q] = (q < N) ? a[q] * b[q] : 0.0;
}

// x += a d ; r -= a z ; part[block] = sum r.r
__global__ void pcg_update_kernel(long n, double a, double* __restrict__ x, double* __restrict__ r, const double* __restrict__ d,
                                  const double* __restrict__ z, double* __restrict__ part) {
  __shared__ double sh[32];
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) {
    x[i] = fma(a, d[i], x[i]);
    const double ri = fma(-a, z[i], r[i]);
    r[i] = ri; acc = fma(ri, ri, acc);
  }
  const double t = block_sum(acc, sh);
  if (threadIdx.x == 0) part[blockIdx.x] = t;
}
// d = s + beta d
__global__ void pcg_dir_kernel(long n, double beta, const double* __restrict__ s, double* __restrict__ d) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) d[i] = fma(beta, d[i], s[i]);
}

}  // namespace kgp
