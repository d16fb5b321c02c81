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
                  int warm[GRID_MAXP];      // warm: V holds the previous eigenvectors (split = 0: and A = V' K V, nearly diagonal)
                  int split[GRID_MAXP]; };  // centrosymmetric factor: two half-order problems (jacobi_eig_kernel)
// C = A B (transA = 0) or A' B (transA = 1), n x n row-major; the warm start forms V'(K V) with two launches
__global__ void matmul_small_kernel(const double* __restrict__ A, const double* __restrict__ B, double* __restrict__ C, int n, int transA) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= n * n) return;
  const int i = q / n, j = q - i * n;
  double acc = 0.0;
  for (int k = 0; k < n; ++k) acc = fma(transA ? A[(long)k * n + i] : A[(long)i * n + k], B[(long)k * n + j], acc);
  C[q] = acc;
}
// Cyclic two-sided Jacobi sweeps on A (n x n, pitch ld; destroyed: ends diagonal) with V <- V J accumulated.
// Round-robin ordering: every round rotates n/2 disjoint index pairs.  The pairs partition the matrix into 2x2 blocks
// (pair a x pair b), each rotated from both sides by one thread; V <- V J by (row, pair) items.  Both kinds of item go
// through one flat index (no division: a float reciprocal is exact at these sizes) and are handled TWO at a time with all
// loads ahead of the stores -- a round is a chain of dependent shared-memory round trips, not a throughput problem.
// The rotation comes from two rsqrt instead of div / sqrt / div / sqrt:  with d = aqq - app, b = 2 apq, r = hypot(d, b):
//   cos 2t = |d| / r,  c = sqrt((1 + cos 2t) / 2),  s = sign(d) b / (2 r c)      (|t| <= pi/4, the classical choice).
// Two barriers per round.
struct JacItem { int i0, i1, i2, i3; double c0, s0, c1, s1; double v00, v01, v10, v11; int kind; };
__device__ __forceinline__ void jac_load(JacItem& it, int t, int nblk2, int nitems, int half, float inv_half, int ld,
                                         const double* __restrict__ A, const double* __restrict__ V,
                                         const double2* __restrict__ rcs, const int2* __restrict__ rpq) {
  it.kind = 0;
  if (t >= nitems) return;
  if (t < nblk2) {
    const int a = (int)(((float)t + 0.5f) * inv_half), b2 = t - a * half;
    const int2 pqa = rpq[a], pqb = rpq[b2]; const double2 csa = rcs[a], csb = rcs[b2];
    const bool va = pqa.y >= 0, vb = pqb.y >= 0;
    it.kind = 1 | (va ? 2 : 0) | (vb ? 4 : 0) | (a == b2 ? 8 : 0);
    it.i0 = pqa.x * ld; it.i1 = (va ? pqa.y : pqa.x) * ld; it.i2 = pqb.x; it.i3 = vb ? pqb.y : pqb.x;
    it.c0 = csa.x; it.s0 = csa.y; it.c1 = csb.x; it.s1 = csb.y;
    it.v00 = A[it.i0 + it.i2]; it.v01 = vb ? A[it.i0 + it.i3] : 0.0;
    it.v10 = va ? A[it.i1 + it.i2] : 0.0; it.v11 = (va && vb) ? A[it.i1 + it.i3] : 0.0;
  } else {
    const int u = t - nblk2, r = (int)(((float)u + 0.5f) * inv_half), k = u - r * half;
    const int2 pq = rpq[k];
    if (pq.y < 0) return;
    const double2 cs = rcs[k];
    it.kind = 16; it.i0 = r * ld + pq.x; it.i1 = r * ld + pq.y; it.c0 = cs.x; it.s0 = cs.y;
    it.v00 = V[it.i0]; it.v01 = V[it.i1];
  }
}
__device__ __forceinline__ void jac_store(const JacItem& it, double* __restrict__ A, double* __restrict__ V) {
  if (it.kind & 1) {
    const double ca = it.c0, sa = it.s0, cb = it.c1, sb = it.s1;
    const double tpp = cb * it.v00 - sb * it.v01, tpq = sb * it.v00 + cb * it.v01;
    const double tqp = cb * it.v10 - sb * it.v11, tqq = sb * it.v10 + cb * it.v11;
    double npp = ca * tpp - sa * tqp, nqp = sa * tpp + ca * tqp, npq = ca * tpq - sa * tqq, nqq = sa * tpq + ca * tqq;
    if (it.kind & 8) { npq = 0.0; nqp = 0.0; }                    // the annihilated pair: exact zeros, not rounding residue
    const bool va = it.kind & 2, vb = it.kind & 4;
    A[it.i0 + it.i2] = npp;
    if (vb) A[it.i0 + it.i3] = npq;
    if (va) A[it.i1 + it.i2] = nqp;
    if (va && vb) A[it.i1 + it.i3] = nqq;
  } else if (it.kind & 16) {
    V[it.i0] = it.c0 * it.v00 - it.s0 * it.v01; V[it.i1] = it.s0 * it.v00 + it.c0 * it.v01;
  }
}
__device__ void jacobi_sweeps(double* __restrict__ A, double* __restrict__ V, const int ng, const int ld, const int max_sweeps,
                              double* sh, double2* rcs, int2* rpq) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = EIG_THREADS / 32;
  const int m = (ng + 1) & ~1, half = m / 2;
  const float inv_half = 1.0f / (float)half;
  const int nblk2 = half * half, nitems = nblk2 + ng * half;
  for (int sweep = 0; sweep < max_sweeps; ++sweep) {
    // convergence: off-diagonal Frobenius norm against the diagonal
    double off = 0.0, dia = 0.0;
    for (int r = warp; r < ng; r += NW)
      for (int c = lane; c < ng; c += 32) {
        const double v = A[r * ld + c];
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
        else { p = round + k; if (p >= m - 1) p -= m - 1; q = round - k; if (q < 0) q += m - 1; }
        if (p > q) { int t = p; p = q; q = t; }
        double c = 1.0, s = 0.0;
        if (q < ng) {
          const double apq = A[p * ld + q];
          if (apq != 0.0) {
            const double d = A[q * ld + q] - A[p * ld + p], b = 2.0 * apq, ss = fma(d, d, b * b);
            if (ss > 1e-290 && ss < 1e290) {
              const double inv_r = rsqrt(ss), h2 = fma(0.5 * fabs(d), inv_r, 0.5), ic = rsqrt(h2);
              c = h2 * ic; s = (d >= 0.0 ? 0.5 : -0.5) * b * inv_r * ic;
            } else {                                              // out-of-range magnitudes: the division form
              const double tau = d / b, t = (tau >= 0.0 ? 1.0 : -1.0) / (fabs(tau) + sqrt(1.0 + tau * tau));
              c = 1.0 / sqrt(1.0 + t * t); s = t * c;
            }
          }
        } else q = -1;
        rpq[k] = make_int2(p, q); rcs[k] = make_double2(c, s);
      }
      __syncthreads();
      for (int t = tid; t < nitems; t += 2 * EIG_THREADS) {       // A <- J' A J and V <- V J in one pass
        JacItem i0, i1;
        jac_load(i0, t, nblk2, nitems, half, inv_half, ld, A, V, rcs, rpq);
        jac_load(i1, t + EIG_THREADS, nblk2, nitems, half, inv_half, ld, A, V, rcs, rpq);
        jac_store(i0, A, V);
        jac_store(i1, A, V);
      }
      __syncthreads();
    }
  }
}

// One CTA pair per factor (blockIdx.x = 2 d + h).
// split[d] = 0: CTA h = 0 decomposes the whole ng x ng factor (shared memory when it fits, else in place in global).
// split[d] = 1 (the grid axis is mirror-symmetric -- every equispaced axis is -- and the factor kernel stationary, so the
//   factor is symmetric Toeplitz, hence centrosymmetric, J T J = T): with Q = [I I; J -J]/sqrt 2 (odd ng: a middle row)
//   Q' T Q = diag(C+, C-),  C+-[i][j] = T[i][j] +- T[i][ng-1-j]  (odd: C+ bordered by sqrt 2 T[i][m], T[m][m]),
//   two independent problems of HALF the order: a quarter of the work per round, half the rounds per sweep, and the two
//   halves run on two SMs.  Eigenvectors [u; (a); J u]/sqrt 2 and [w; (0); -J w]/sqrt 2.  ng = 100: 7.9 ms -> cold / 1.07 ms
//   warm of the full-order kernel were the whole evaluation time of the reference's msgp_lstm_actuator-sized problems.
//   Warm start inside the kernel: the previous half-eigenvectors are read back out of V's columns, C <- U' C U.
__global__ void __launch_bounds__(EIG_THREADS) jacobi_eig_kernel(const EigBatch eb, int max_sweeps) {
  const int d = blockIdx.x >> 1, h = blockIdx.x & 1;
  double* __restrict__ Ag = eb.A[d]; double* __restrict__ Vg = eb.V[d];
  double* __restrict__ ev = eb.ev[d];
  const int ng = eb.ng[d], use_smem = eb.use_smem[d], warm = eb.warm[d];
  if (ng == 0) return;                           // Toeplitz factor: Whittle eigenvalues instead (grid_toeplitz.cuh)
  extern __shared__ double sm[];
  __shared__ double sh[32];
  __shared__ double2 rcs[1024];                  // rotation of pair k: cos, sin
  __shared__ int2 rpq[1024];
  const int tid = threadIdx.x;
  if (!eb.split[d]) {
    if (h) return;
    const int ld = use_smem ? (ng | 1) : ng;     // odd pitch in shared memory: conflict-free rows AND columns
    double* A = use_smem ? sm : Ag;
    double* V = use_smem ? sm + (long)ng * ld : Vg;
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
    jacobi_sweeps(A, V, ng, ld, max_sweeps, sh, rcs, rpq);
    for (int i = tid; i < ng; i += EIG_THREADS) ev[i] = fmax(A[(long)i * ld + i], 0.0);   // apxGrid.m:586
    if (use_smem)
      for (long q = tid; q < (long)ng * ng; q += EIG_THREADS) Vg[q] = V[(q / ng) * ld + q % ng];
    return;
  }
  // ---- centrosymmetric split: half h (0: symmetric eigenvectors, 1: skew-symmetric)
  const int m = ng >> 1, odd = ng & 1, nh = h == 0 ? m + odd : m, col0 = h == 0 ? 0 : m + odd;
  if (nh == 0) return;
  const int ld = nh | 1;
  double* C = sm; double* U = sm + nh * ld; double* T = U + nh * ld;
  const double sgn = h == 0 ? 1.0 : -1.0, r2 = sqrt(2.0), ir2 = 1.0 / r2;
  auto Ks = [&](int i, int j) { return 0.5 * (Ag[(long)i * ng + j] + Ag[(long)j * ng + i]); };
  for (int q = tid; q < nh * nh; q += EIG_THREADS) {
    const int i = q / nh, j = q - i * nh;
    double v;
    if (i < m && j < m) v = 0.5 * ((Ks(i, j) + sgn * Ks(i, ng - 1 - j)) + (Ks(j, i) + sgn * Ks(j, ng - 1 - i)));
    else if (i == m && j == m) v = Ks(m, m);
    else v = r2 * Ks(i < m ? i : j, m);                            // the border of C+ (odd ng)
    C[i * ld + j] = v;
    double u = (i == j) ? 1.0 : 0.0;
    if (warm) u = (i < m ? r2 : 1.0) * Vg[(long)i * ng + col0 + j];   // previous half-eigenvectors out of V's columns
    U[i * ld + j] = u;
  }
  __syncthreads();
  if (warm) {                                                     // C <- U' (C U): nearly diagonal start
    for (int q = tid; q < nh * nh; q += EIG_THREADS) {
      const int i = q / nh, j = q - i * nh;
      double acc = 0.0;
      for (int k = 0; k < nh; ++k) acc = fma(C[i * ld + k], U[k * ld + j], acc);
      T[i * ld + j] = acc;
    }
    __syncthreads();
    for (int q = tid; q < nh * nh; q += EIG_THREADS) {
      const int i = q / nh, j = q - i * nh;
      double acc = 0.0;
      if (j >= i) {
        for (int k = 0; k < nh; ++k) acc = fma(U[k * ld + i], T[k * ld + j], acc);
        C[i * ld + j] = acc;
      }
    }
    __syncthreads();
    for (int q = tid; q < nh * nh; q += EIG_THREADS) {            // exact symmetry from the upper triangle
      const int i = q / nh, j = q - i * nh;
      if (j < i) C[i * ld + j] = C[j * ld + i];
    }
    __syncthreads();
  }
  jacobi_sweeps(C, U, nh, ld, max_sweeps, sh, rcs, rpq);
  for (int k = tid; k < nh; k += EIG_THREADS) ev[col0 + k] = fmax(C[k * ld + k], 0.0);   // apxGrid.m:586
  for (int q = tid; q < nh * nh; q += EIG_THREADS) {
    const int i = q / nh, k = q - i * nh;
    const double u = U[i * ld + k];
    if (i < m) { Vg[(long)i * ng + col0 + k] = u * ir2; Vg[(long)(ng - 1 - i) * ng + col0 + k] = sgn * u * ir2; }
    else Vg[(long)m * ng + col0 + k] = u;
  }
  if (odd && h == 1) for (int k = tid; k < nh; k += EIG_THREADS) Vg[(long)m * ng + col0 + k] = 0.0;
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
