// KISS-GP path, part 3: symmetric eigen-decomposition of the dense Kronecker factors (cyclic
// two-sided Jacobi, parallel round-robin ordering, one CTA per factor) and the Fiedler log-det
// bound with its derivatives (GPML/cov/apxGrid.m:571-588, GPML/cov/apx.m:246-269,312-318).
#pragma once
#include "grid_kernels2.cuh"

namespace kgp {

// A (ng x ng, symmetric; destroyed) -> eigenvalues ev[ng] (unsorted), eigenvectors V (columns).
// Works out of shared memory when 2*ng*ng*8 bytes fit (use_smem), otherwise in place in global.
constexpr int EIG_THREADS = 512;
struct EigBatch { double* A[GRID_MAXP]; double* V[GRID_MAXP]; double* ev[GRID_MAXP]; int ng[GRID_MAXP]; int use_smem[GRID_MAXP];
                  int warm[GRID_MAXP]; };   // warm: V holds the previous eigenvectors and A = V' K V (nearly diagonal)
// C = A B (transA = 0) or A' B (transA = 1), n x n row-major; the warm start forms V'(K V) with two launches
__global__ void matmul_small_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C, int n, int transA) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * n) return;
  const int i = q / n, j = q - i * n;
  double acc = 0.0;
  for (int k = 0; k < n; ++k) acc = fma(transA ? A[(long)k * n + i] : A[(long)i * n + k], B[(long)k * n + j], acc);
  C[q] = acc;
}
__global__ void __launch_bounds__(EIG_THREADS) jacobi_eig_kernel(const EigBatch eb, int max_sweeps) {
  double* __restrict__ Ag = eb.A[blockIdx.x]; double* __restrict__ Vg = eb.V[blockIdx.x];
  double* __restrict__ ev = eb.ev[blockIdx.x];
  const int ng = eb.ng[blockIdx.x], use_smem = eb.use_smem[blockIdx.x], warm = eb.warm[blockIdx.x];
  if (ng == 0) return;                           // Toeplitz factor: Whittle eigenvalues instead (grid_toeplitz.cuh)
  extern __shared__ double sm[];
  __shared__ double sh[32];
  __shared__ double rc[1024], rs[1024];          // rotation of pair k: cos, sin
  __shared__ int rp[1024], rq[1024];
  const int ld = use_smem ? (ng | 1) : ng;       // odd pitch in shared memory: conflict-free rows AND columns
  double* A = use_smem ? sm : Ag;
  double* V = use_smem ? sm + (long)ng * ld : Vg;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (long q = tid; q < (long)ng * ng; q += EIG_THREADS) {
    int r = (int)(q / ng), c = (int)(q % ng);
    if (use_smem) A[(long)r * ld + c] = 0.5 * (Ag[q] + Ag[(long)c * ng + r]);   // Kj = (Kj+Kj')/2   (apxGrid.m:583)
    if (warm) { if (use_smem) V[(long)r * ld + c] = Vg[q]; }
    else V[(long)r * ld + c] = (r == c) ? 1.0 : 0.0;
  }
  __syncthreads();
  if (!use_smem) {                                              // symmetrise in place, second pass
    for (long q = tid; q < (long)ng * ng; q += EIG_THREADS) {
      int r = (int)(q / ng), c = (int)(q % ng);
      if (c > r) { double v = 0.5 * (A[q] + A[(long)c * ng + r]); A[q] = v; A[(long)c * ng + r] = v; }
    }
    __syncthreads();
  }
  const int m = (ng + 1) & ~1, half = m / 2;
  for (int sweep = 0; sweep < max_sweeps; ++sweep) {
    // convergence: off-diagonal Frobenius norm against the diagonal
    double off = 0.0, dia = 0.0;
    for (int r = warp; r < ng; r += EIG_THREADS / 32)
      for (int c = lane; c < ng; c += 32) {
        double v = A[(long)r * ld + c];
        if (r == c) dia = fma(v, v, dia); else off = fma(v, v, off);
      }
    off = block_sum(off, sh); dia = block_sum(dia, sh);
    // squared norms: 1e-26 = relative off-diagonal norm 1e-13; rounding keeps the residue of a 70 x 70 factor near
    // 70 eps^2 = 3e-30, so a tighter test would never fire and all max_sweeps would run
    if (off <= 1e-26 * dia || off == 0.0) break;
    for (int round = 0; round < m - 1; ++round) {
      // pairs of this round + their rotation angles
      for (int k = tid; k < half; k += EIG_THREADS) {
        int p, q;
        if (k == 0) { p = m - 1; q = round; }
        else { p = (round + k) % (m - 1); q = (round - k + m - 1) % (m - 1); }
        if (p > q) { int t = p; p = q; q = t; }
        double c = 1.0, s = 0.0;
        if (q < ng) {
          const double apq = A[(long)p * ld + q];
          if (apq != 0.0) {
            const double tau = (A[(long)q * ld + q] - A[(long)p * ld + p]) / (2.0 * apq);
            const double t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
            c = 1.0 / sqrt(1.0 + t * t); s = t * c;
          }
        } else q = -1;
        rp[k] = p; rq[k] = q; rc[k] = c; rs[k] = s;
      }
      __syncthreads();
      // A <- J' A J in ONE pass: the index pairs partition the matrix into 2x2 blocks (pair a x pair b), each
      // rotated from both sides by one thread; V <- V J by (row, pair) items.  Two barriers per round.
      const int nblk2 = half * half, nitems = nblk2 + ng * half;
      for (int t = tid; t < nitems; t += EIG_THREADS) {
        if (t < nblk2) {
          const int a = t / half, b2 = t - a * half;
          const int pa = rp[a], qa = rq[a], pb = rp[b2], qb = rq[b2];
          const double ca = rc[a], sa = rs[a], cb = rc[b2], sb = rs[b2];
          const bool va = qa >= 0, vb = qb >= 0;
          const double b00 = A[(long)pa * ld + pb], b01 = vb ? A[(long)pa * ld + qb] : 0.0;
          const double b10 = va ? A[(long)qa * ld + pb] : 0.0, b11 = (va && vb) ? A[(long)qa * ld + qb] : 0.0;
          const double tpp = cb * b00 - sb * b01, tpq = sb * b00 + cb * b01;
          const double tqp = cb * b10 - sb * b11, tqq = sb * b10 + cb * b11;
          double npp = ca * tpp - sa * tqp, nqp = sa * tpp + ca * tqp, npq = ca * tpq - sa * tqq, nqq = sa * tpq + ca * tqq;
          if (a == b2) { npq = 0.0; nqp = 0.0; }              // the annihilated pair: exact zeros, not rounding residue
          A[(long)pa * ld + pb] = npp;
          if (vb) A[(long)pa * ld + qb] = npq;
          if (va) A[(long)qa * ld + pb] = nqp;
          if (va && vb) A[(long)qa * ld + qb] = nqq;
        } else {
          const int u = t - nblk2, r = u / half, k = u - r * half;
          const int pp = rp[k], qq = rq[k];
          if (qq >= 0) {
            const double c = rc[k], s2 = rs[k];
            const double vp = V[(long)r * ld + pp], vq = V[(long)r * ld + qq];
            V[(long)r * ld + pp] = c * vp - s2 * vq; V[(long)r * ld + qq] = s2 * vp + c * vq;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = tid; i < ng; i += EIG_THREADS) ev[i] = fmax(A[(long)i * ld + i], 0.0);   // apxGrid.m:586
  if (use_smem)
    for (long q = tid; q < (long)ng * ng; q += EIG_THREADS) Vg[q] = V[(q / ng) * ld + q % ng];
}

// Fiedler bound pieces for every Kronecker eigenvalue product e (all W equal to w):
//   k = n/N e ; kw1 = k w + 1 ; ub += log(kw1)/2 ; sdW += k/(2 kw1) ; de = [e>0] w / (N 2/n kw1)
// `keep` (optional, N flags) restricts to the n largest when n < N  (apx.m:313-318)
__global__ void fiedler_kernel(GridDesc G, const double* __restrict__ ee /* concatenated */, const int* __restrict__ eoff,
                               double n_over_N, double w, const unsigned char* __restrict__ keep,
                               double* __restrict__ de, double* __restrict__ part_ub, double* __restrict__ part_dw) {
  __shared__ double sh[32];
  double ub = 0.0, dw = 0.0;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < G.N; q += (long)gridDim.x * blockDim.x) {
    long r = q; double e = 1.0;
    for (int d = 0; d < G.p; ++d) { e *= ee[eoff[d] + (int)(r % G.ng[d])]; r /= G.ng[d]; }
    double dev = 0.0;
    if (!keep || keep[q]) {
      const double k = n_over_N * e, kw1 = k * w + 1.0;
      ub += log(kw1) * 0.5; dw += k / (2.0 * kw1);
      if (e > 0.0) dev = w / (2.0 / n_over_N * kw1);
    }
    de[q] = dev;
  }
  double t = block_sum(ub, sh);
  if (threadIdx.x == 0) part_ub[blockIdx.x] = t;
  t = block_sum(dw, sh);
  if (threadIdx.x == 0) part_dw[blockIdx.x] = t;
}

// Kronecker eigenvalue products into a flat array (for the top-n selection when n < N)
__global__ void kron_eig_kernel(GridDesc G, const double* __restrict__ ee, const int* __restrict__ eoff,
                                double* __restrict__ e, int* __restrict__ idx, long Npad) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Npad) return;
  double v = -1.0;                                    // padding sorts last (eigenvalues are >= 0)
  if (q < G.N) {
    long r = q; v = 1.0;
    for (int d = 0; d < G.p; ++d) { v *= ee[eoff[d] + (int)(r % G.ng[d])]; r /= G.ng[d]; }
  }
  e[q] = v; idx[q] = (int)q;
}
// one compare-exchange stage of a bitonic sort (descending) on (key, index) pairs
__global__ void bitonic_stage_kernel(double* __restrict__ key, int* __restrict__ idx, long Npad, long j, long k) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Npad) return;
  long l = i ^ j;
  if (l > i) {
    const bool desc = (i & k) == 0;
    double a = key[i], b = key[l];
    const bool sw = desc ? (a < b) : (a > b);
    if (sw) { key[i] = b; key[l] = a; int t = idx[i]; idx[i] = idx[l]; idx[l] = t; }
  }
}
__global__ void mark_keep_kernel(const int* __restrict__ idx, long nkeep, long N, unsigned char* __restrict__ keep) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < nkeep && idx[q] < N) keep[idx[q]] = 1;
}

// z_i[a] = sum over the other modes of de[..a..] * prod_{j != i} ee_j[..]    (apx.m:249-250)
__global__ void fiedler_z_kernel(GridDesc G, int mode, const double* __restrict__ ee, const int* __restrict__ eoff,
                                 const double* __restrict__ de, double* __restrict__ z) {
  __shared__ double sh[32];
  const int a = blockIdx.x;
  double acc = 0.0;
  for (long q = threadIdx.x; q < G.N; q += blockDim.x) {
    long r = q; double w = 1.0; bool mine = true;
    for (int d = 0; d < G.p; ++d) {
      int id = (int)(r % G.ng[d]); r /= G.ng[d];
      if (d == mode) { if (id != a) { mine = false; break; } }
      else w *= ee[eoff[d] + id];
    }
    if (mine) acc = fma(w, de[q], acc);
  }
  double t = block_sum(acc, sh);
  if (threadIdx.x == 0) z[a] = t;
}
// Q = V diag(z) V'   (apx.m:252)
__global__ void vdv_kernel(const double* __restrict__ V, const double* __restrict__ z, int ng, double* __restrict__ Qm) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)ng * ng) return;
  int a = (int)(q / ng), b = (int)(q % ng);
  double acc = 0.0;
  for (int k = 0; k < ng; ++k) acc = fma(V[(long)a * ng + k] * z[k], V[(long)b * ng + k], acc);
  Qm[q] = acc;
}
__global__ void sum_small_kernel(const double* __restrict__ part, int cnt, double* __restrict__ out) {
  __shared__ double sh[32];
  double t = sum_partials(part, cnt, sh);
  if (threadIdx.x == 0) out[0] = t;
}
__global__ void axpy_scale_kernel(long n, double a, const double* __restrict__ x, double* __restrict__ y) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = a * x[i];
}

}  // namespace kgp
