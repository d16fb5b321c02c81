// KISS-GP path, part 4: the conjugate-gradient solve in CELL-SORTED point order.
//
// Reference: conjgrad / linsolve (GPML/cov/apx.m:275-302) on B = I + sW Mx Kg Mx' sW with
// per-column convergence masks (apx.m:296-298) and the `iter = 2..maxit` count (apx.m:289).
//
// Layout: points are sorted once per (data, grid) by the interpolation cell they fall into (the
// clamped first node of eqinterp, apxGrid.m:505-523).  Per point the kernels stream only the
// pre-divided grid coordinate u = (x - s0)/ds (same FP64 division as eqinterp, done once), so an
// iteration touches  U (8P B, twice) + the CG vectors: n*(16P + 88) bytes -- SURVEY 8(d).
// One iteration = 3 streaming kernels + the Kronecker MVM:
//   scatter : d = r + beta d (fused), g0 += Mx'(sW d)  -- per-thread runs of 4 consecutive sorted
//             points accumulate their 4^P node weights in registers, a warp-level segmented
//             reduction (shuffles) merges equal cells, one RED.F64 per node and segment
//   kron    : u = (K_2 kron K_1) g0                      (grid-sized, L2 resident)
//   gather  : z = sW Mx u + d, d.z                       (node values cached per cell in registers)
//   update  : x += a d, r -= a z, r.r, zero g0
// All scalars (a, beta, relres, iteration count, convergence flag) live in a device control block
// written by the LAST CTA of the producing kernel (ticket counter), so the whole loop replays as a
// CUDA graph with no host involvement; the host polls the flag once per graph launch.
// blockIdx.y = right-hand-side column (each column has its own control block -> the per-column
// masks of apx.m:296-298).
#pragma once
#include "grid_kernels3.cuh"

namespace kgp {

enum { CT_A = 0, CT_BETA, CT_DONE, CT_ITER, CT_RELRES, CT_NB, CT_RRCUR, CT_ACC_RR, CT_ACC_DZ, CT_TOL, CT_MAXIT, CT_SW, CT_PASSES,
       CT_STRIDE = 16 };
constexpr int CG_PTS = 4;                     // consecutive sorted points per thread
constexpr int CG_SUB = 256 * CG_PTS;          // points per CTA pass

// ---- cubic convolution weights on a pre-divided coordinate --------------------------------------
// kcub on its two branches with the operation order of kcub() (grid_kernels.cuh)
__device__ __forceinline__ double kcub_in(double x) { return fma(fma(1.5, x, -2.5) * x, x, 1.0); }           // |x| <= 1
__device__ __forceinline__ double kcub_out(double x) { return fma(fma(-0.5, x, 2.5) * x - 4.0, x, 2.0); }    // 1 <= |x| <= 2

// encoded coordinate: up >= 0 -> u itself (regular point, floor(u) in [2, ns-3]);
// up < 0 -> snapped to one node: code = -up-1, bits 0-1 = which of the 4 edge nodes, bit 2 = upper edge
__device__ __forceinline__ double encode_u(const double* __restrict__ ax, int ns, double s0, double ds, double t) {
  const double u = (t - s0) / ds;
  double jf = floor(u);
  if (jf < 2.0) {
    int best = 0; double bd = fabs(ax[0] - t);
#pragma unroll
    for (int i = 1; i < 4; ++i) { double d = fabs(ax[i] - t); if (d < bd) { bd = d; best = i; } }
    jf = 1.0;
    if (!(jf > (double)(ns - 3))) return -(double)(best + 1);
  }
  if (jf > (double)(ns - 3)) {
    int best = 0; double bd = fabs(ax[ns - 4] - t);
#pragma unroll
    for (int i = 1; i < 4; ++i) { double d = fabs(ax[ns - 4 + i] - t); if (d < bd) { bd = d; best = i; } }
    return -(double)(4 + best + 1);
  }
  return u;
}
__device__ __forceinline__ void decode_u(double up, int ns, int& j, double c[4]) {
  // branch-free: both forms are evaluated and selected, so the decodes of a thread's points interleave
  const bool reg = up >= 0.0;
  const double ur = reg ? up : 2.0;
  const double fl = floor(ur);
  const double w = ur - (fl + 1.0) + 1.0;               // eqinterp: w = u - j + 1 with j = floor(u)+1
  const double c0 = kcub_out(w + 1.0), c1 = kcub_in(w), c2 = kcub_in(fabs(w - 1.0)), c3 = kcub_out(fabs(w - 2.0));
  const int code = reg ? 0 : (int)(-up) - 1, a = code & 3;
  j = reg ? (int)fl - 1 : ((code & 4) ? ns - 4 : 0);
  c[0] = reg ? c0 : (a == 0 ? 1.0 : 0.0);
  c[1] = reg ? c1 : (a == 1 ? 1.0 : 0.0);
  c[2] = reg ? c2 : (a == 2 ? 1.0 : 0.0);
  c[3] = reg ? c3 : (a == 3 ? 1.0 : 0.0);
}

// first node only (cheap form of decode_u)
__device__ __forceinline__ int decode_j(double up, int ns) {
  const bool reg = up >= 0.0;
  const int code = reg ? 0 : (int)(-up) - 1;
  return reg ? (int)floor(up) - 1 : ((code & 4) ? ns - 4 : 0);
}

// ---- set-up: sort key, sorted coordinates ---------------------------------------------------------
// key = cell * Npad + i (all distinct -> the descending bitonic sort gives one deterministic order)
template <int P>
__global__ void cg_cellkey_kernel(GridDesc G, const double* __restrict__ X, long n, int D, double* __restrict__ key,
                                  int* __restrict__ idx, long Npad) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Npad) return;
  double k = -1.0;
  if (i < n) {
    long cell = 0, mul = 1;
#pragma unroll
    for (int d = 0; d < P; ++d) {
      int j; double c[4];
      decode_u(encode_u(G.ax[d], G.ng[d], G.s0[d], G.ds[d], X[i * D + d]), G.ng[d], j, c);
      cell += (long)j * mul; mul *= (long)(G.ng[d] - 3);
    }
    k = (double)cell * (double)Npad + (double)i;
  }
  key[i] = k; idx[i] = (int)i;
}
template <int P>
__global__ void cg_sorted_coords_kernel(GridDesc G, const double* __restrict__ X, long n, int D,
                                        const int* __restrict__ perm, long npad, double* __restrict__ U) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npad) return;
  const long src = perm[i < n ? i : n - 1];             // padding repeats the last point (its vector entries stay 0)
#pragma unroll
  for (int d = 0; d < P; ++d) U[i * P + d] = encode_u(G.ax[d], G.ng[d], G.s0[d], G.ds[d], X[src * D + d]);
}

// ---- last-CTA bookkeeping --------------------------------------------------------------------------
// adds this CTA's partial to *acc; returns true in thread 0 of the CTA that arrives last (then *acc is complete)
__device__ __forceinline__ bool cta_accumulate_last(double partial, double* acc, unsigned* counter, unsigned nctas) {
  if (threadIdx.x != 0) return false;
  atomicAdd(acc, partial);
  __threadfence();
  const unsigned t = atomicAdd(counter, 1u);
  if (t != nctas - 1) return false;
  __threadfence();
  *counter = 0u;
  return true;
}

// b = sw * rhs[perm] -> r, d ; x = 0 ; scalars of apx.m:284-287
__global__ void __launch_bounds__(256) cg_init_sorted_kernel(const double* __restrict__ rhs, long rhs_stride,
    const int* __restrict__ perm, long n, long npad, double sw, double tol, int maxit, double* __restrict__ x,
    double* __restrict__ r, double* __restrict__ d, double* __restrict__ ctl, unsigned* __restrict__ cnt) {
  __shared__ double sh[32];
  const int col = blockIdx.y;
  double* ct = ctl + col * CT_STRIDE;
  rhs += (long)col * rhs_stride; x += (long)col * npad; r += (long)col * npad; d += (long)col * npad;
  double acc = 0.0;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npad; i += (long)gridDim.x * blockDim.x) {
    const double b = (i < n) ? sw * rhs[perm[i]] : 0.0;
    x[i] = 0.0; r[i] = b; d[i] = b; acc = fma(b, b, acc);
  }
  const double t = block_sum(acc, sh);
  if (cta_accumulate_last(t, ct + CT_ACC_RR, cnt + col * 4, gridDim.x)) {
    const double rr = atomicAdd(ct + CT_ACC_RR, 0.0);
    const double nb = sqrt(rr), rel = nb / nb;
    ct[CT_ACC_RR] = 0.0; ct[CT_ACC_DZ] = 0.0;
    ct[CT_NB] = nb; ct[CT_RRCUR] = rr; ct[CT_RELRES] = rel; ct[CT_ITER] = 1.0; ct[CT_BETA] = 0.0; ct[CT_A] = 0.0;
    ct[CT_TOL] = tol; ct[CT_MAXIT] = (double)maxit; ct[CT_SW] = sw;
    ct[CT_DONE] = (rel >= tol) ? ((2 > maxit) ? 2.0 : 0.0) : 1.0;   // NaN (b == 0): nothing to do, apx.m:286
  }
}

template <int P> struct CgNodes { static constexpr int NW = (P == 1) ? 4 : 16; };

template <int P>
__device__ __forceinline__ void flush_nodes(const GridDesc& G, double* __restrict__ g0, int cell, const double* acc) {
  if (P == 1) {
#pragma unroll
    for (int a = 0; a < 4; ++a) if (acc[a] != 0.0) atomicAdd(g0 + cell + a, acc[a]);
  } else {
    const int j1 = cell / G.ng[0], j0 = cell - j1 * G.ng[0];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      double* row = g0 + (long)(j1 + b) * G.stride[1] + j0;
#pragma unroll
      for (int a = 0; a < 4; ++a) if (acc[b * 4 + a] != 0.0) atomicAdd(row + a, acc[b * 4 + a]);
    }
  }
}

// ---- order-independent scatter accumulation (cg_persistent2_kernel) -------------------------------------------------
// Floating-point atomics make Mx'(sW d) depend on the order in which warps and CTAs reach a node, so two runs of the
// same solve differed in the last bits and could stop one iteration apart.  Integer atomics are associative: every
// flushed value v becomes hi = rint(v 2^e1) and is accumulated with ONE 64-bit integer RED; the sum hi 2^-e1 is the same
// for every arrival order.  e1 = 61 - ilogb(bound) with a bound that holds for every partial sum of a node:
//   |sum_i c_i sw d_i| <= sw sqrt(sum_i c_i^2) |d|_2 <= sw C2max |d|_2,   C2max^2 = max over nodes of sum_i c_i^2
// (Cauchy-Schwarz over the points that touch the node; C2max^2 is a property of data + grid, accumulated once in integer
// arithmetic by cg_node_weight2_kernel so that it is itself run-to-run identical) and |d|_2^2 from the recursion
// |d_k|^2 = |r_k|^2 + beta^2 |d_k-1|^2, which every CTA evaluates identically from the shared scalars (x16 safety).
// Resolution: 2^-61 of the bound.  With ~3000 points per node C2max is ~20, not sqrt(n) = 1000, so the absolute resolution
// is ~2^-55 of a typical node value -- the level of the FP64 atomics it replaces -- and normwise (what CG needs) far below.
// (Round 2 first shipped a two-word form, hi + lo with the sqrt(n) bound: two REDs per flush, 53 us per iteration at
// n = 1M against 38 with FP64 atomics; the tighter bound makes the second word unnecessary.)
struct FxScale { double s1, i1, s2, i2; };
__device__ __forceinline__ FxScale fx_scale(double bound) {
  int e1 = 0;
  if (bound > 0.0 && bound < 1e300) e1 = 61 - ilogb(bound);
  e1 = e1 < -1000 ? -1000 : (e1 > 960 ? 960 : e1);
  FxScale f; f.s1 = scalbn(1.0, e1); f.i1 = scalbn(1.0, -e1); f.s2 = 0.0; f.i2 = 0.0;
  return f;
}
// lo is only a mode flag now: non-null = fixed point in hi, null = the array holds plain doubles (FP64-atomics mode)
__device__ __forceinline__ void fx_add(long long* __restrict__ hi, long long* __restrict__ lo, long idx, double v, const FxScale& f) {
  atomicAdd((unsigned long long*)(hi + idx), (unsigned long long)__double2ll_rn(v * f.s1));
}
__device__ __forceinline__ double fx_load(const long long* __restrict__ hi, const long long* __restrict__ lo, long idx, const FxScale& f) {
  if (!lo) return __ldcg((const double*)hi + idx);      // FP64-atomics mode: the array holds plain doubles
  return (double)__ldcg(hi + idx) * f.i1;
}
// w2[node] += ceil(c^2 2^30) for every point and stencil node: integer sums, hence deterministic; ceil keeps it an upper bound
template <int P>
__global__ void cg_node_weight2_kernel(GridDesc G, const double* __restrict__ U, long npts, unsigned long long* __restrict__ w2) {
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < npts; i += (long)gridDim.x * blockDim.x) {
    int j[P]; double c[P][4];
#pragma unroll
    for (int d = 0; d < P; ++d) decode_u(U[i * P + d], G.ng[d], j[d], c[d]);
    if (P == 1) {
#pragma unroll
      for (int a = 0; a < 4; ++a) { const double w = c[0][a]; if (w != 0.0) atomicAdd(w2 + j[0] + a, (unsigned long long)ceil(w * w * 1073741824.0)); }
    } else {
#pragma unroll
      for (int b = 0; b < 4; ++b)
#pragma unroll
        for (int a = 0; a < 4; ++a) {
          const double w = c[0][a] * c[P - 1][b];
          if (w != 0.0) atomicAdd(w2 + (long)(j[P - 1] + b) * G.stride[1] + j[0] + a, (unsigned long long)ceil(w * w * 1073741824.0));
        }
    }
  }
}
__global__ void cg_umax_kernel(const unsigned long long* __restrict__ w2, long N, unsigned long long* __restrict__ out) {
  unsigned long long m = 0;
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < N; q += (long)gridDim.x * blockDim.x) m = w2[q] > m ? w2[q] : m;
  for (int o = 16; o; o >>= 1) { const unsigned long long t = __shfl_xor_sync(0xffffffffu, m, o); m = t > m ? t : m; }
  if ((threadIdx.x & 31) == 0 && m) atomicMax(out, m);
}
template <int P>
__device__ __forceinline__ void flush_nodes_fx(const GridDesc& G, long long* __restrict__ hi, long long* __restrict__ lo, int cell, const double* acc,
                                               const FxScale& f) {
  if (P == 1) {
#pragma unroll
    for (int a = 0; a < 4; ++a) if (acc[a] != 0.0) fx_add(hi, lo, cell + a, acc[a], f);
  } else {
    const int j1 = cell / G.ng[0], j0 = cell - j1 * G.ng[0];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
      const long row = (long)(j1 + b) * G.stride[1] + j0;
#pragma unroll
      for (int a = 0; a < 4; ++a) if (acc[b * 4 + a] != 0.0) fx_add(hi, lo, row + a, acc[b * 4 + a], f);
    }
  }
}

// d = r + beta d ;  g0 += Mx' (sw d)
template <int P>
__global__ void __launch_bounds__(256, 3) cg_scatter_kernel(GridDesc G, const double* __restrict__ U, long npad,
    const double* __restrict__ r, double* __restrict__ d, const double* __restrict__ ctl, double* __restrict__ g0) {
  constexpr int NW = CgNodes<P>::NW;
  const int col = blockIdx.y;
  const double* ct = ctl + col * CT_STRIDE;
  if (ct[CT_DONE] != 0.0) return;
  const double beta = ct[CT_BETA], sw = ct[CT_SW];
  r += (long)col * npad; d += (long)col * npad; g0 += (long)col * G.N;
  const int lane = threadIdx.x & 31;
  const long nunits = npad / (32 * CG_PTS), nwarps = (long)gridDim.x * (blockDim.x >> 5);
  for (long unit = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < nunits; unit += nwarps) {
    const long base = unit * (32 * CG_PTS) + (long)(threadIdx.x & 31) * CG_PTS;
    double dn[CG_PTS], uu[CG_PTS * P];
    {
      const double2 r0 = *(const double2*)(r + base), r1 = *(const double2*)(r + base + 2);
      const double2 d0 = *(const double2*)(d + base), d1 = *(const double2*)(d + base + 2);
      dn[0] = fma(beta, d0.x, r0.x); dn[1] = fma(beta, d0.y, r0.y); dn[2] = fma(beta, d1.x, r1.x); dn[3] = fma(beta, d1.y, r1.y);
      *(double2*)(d + base) = make_double2(dn[0], dn[1]); *(double2*)(d + base + 2) = make_double2(dn[2], dn[3]);
#pragma unroll
      for (int q = 0; q < CG_PTS * P / 2; ++q) {
        const double2 t = __ldg((const double2*)(U + base * P) + q);
        uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
      }
    }
    double acc[NW];
#pragma unroll
    for (int v = 0; v < NW; ++v) acc[v] = 0.0;
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
          for (int v = 0; v < NW; ++v) acc[v] = 0.0;
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
    // segmented reduction over the warp (cells are contiguous in sorted order)
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
      const int oc = __shfl_down_sync(0xffffffffu, cur, off);
      const bool take = (lane + off < 32) && (oc == cur);
#pragma unroll
      for (int v = 0; v < NW; ++v) {
        const double o = __shfl_down_sync(0xffffffffu, acc[v], off);
        if (take) acc[v] += o;
      }
    }
    const int pc = __shfl_up_sync(0xffffffffu, cur, 1);
    if (lane == 0 || pc != cur) flush_nodes<P>(G, g0, cur, acc);
  }
}

// z = sw (Mx u) + d ; d.z ; last CTA: a = r2 / (d.z)
template <int P>
__global__ void __launch_bounds__(256, 3) cg_gather_kernel(GridDesc G, const double* __restrict__ U, long n, long npad,
    const double* __restrict__ ug, const double* __restrict__ d, double* __restrict__ z, double* __restrict__ ctl,
    unsigned* __restrict__ cnt) {
  constexpr int NW = CgNodes<P>::NW;
  __shared__ double sh[32];
  const int col = blockIdx.y;
  double* ct = ctl + col * CT_STRIDE;
  if (ct[CT_DONE] != 0.0) return;
  const double sw = ct[CT_SW];
  d += (long)col * npad; z += (long)col * npad; ug += (long)col * G.N;
  double dz = 0.0;
  const long nunits = npad / (32 * CG_PTS), nwarps = (long)gridDim.x * (blockDim.x >> 5);
  for (long unit = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < nunits; unit += nwarps) {
    const long base = unit * (32 * CG_PTS) + (long)(threadIdx.x & 31) * CG_PTS;
    double dv[CG_PTS], uu[CG_PTS * P], zv[CG_PTS];
    {
      const double2 d0 = *(const double2*)(d + base), d1 = *(const double2*)(d + base + 2);
      dv[0] = d0.x; dv[1] = d0.y; dv[2] = d1.x; dv[3] = d1.y;
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
      const double zi = (base + q < n) ? fma(sw, mu, dv[q]) : 0.0;
      zv[q] = zi; dz = fma(dv[q], zi, dz);
    }
    *(double2*)(z + base) = make_double2(zv[0], zv[1]); *(double2*)(z + base + 2) = make_double2(zv[2], zv[3]);
  }
  const double t = block_sum(dz, sh);
  if (cta_accumulate_last(t, ct + CT_ACC_DZ, cnt + col * 4 + 1, gridDim.x)) {
    const double tot = atomicAdd(ct + CT_ACC_DZ, 0.0);
    ct[CT_A] = ct[CT_RRCUR] / tot; ct[CT_ACC_DZ] = 0.0;
  }
}

// x += a d ; r -= a z ; r.r ; g0 = 0 ; last CTA: relres, convergence, iteration count, beta  (apx.m:291-300)
__global__ void __launch_bounds__(256) cg_update_kernel(long npad, double* __restrict__ x, double* __restrict__ r,
    const double* __restrict__ d, const double* __restrict__ z, double* __restrict__ ctl, unsigned* __restrict__ cnt,
    double* __restrict__ g0, long N) {
  __shared__ double sh[32];
  const int col = blockIdx.y;
  double* ct = ctl + col * CT_STRIDE;
  if (ct[CT_DONE] != 0.0) return;
  const double a = ct[CT_A];
  x += (long)col * npad; r += (long)col * npad; d += (long)col * npad; z += (long)col * npad; g0 += (long)col * N;
  double rr = 0.0;
  const long nunits = npad / (32 * CG_PTS), nwarps = (long)gridDim.x * (blockDim.x >> 5);
  for (long unit = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); unit < nunits; unit += nwarps) {
    const long base = unit * (32 * CG_PTS) + (long)(threadIdx.x & 31) * CG_PTS;
#pragma unroll
    for (int h = 0; h < CG_PTS; h += 2) {
      double2 xv = *(const double2*)(x + base + h), rv = *(const double2*)(r + base + h);
      const double2 dv = *(const double2*)(d + base + h), zv = *(const double2*)(z + base + h);
      xv.x = fma(a, dv.x, xv.x); xv.y = fma(a, dv.y, xv.y);
      rv.x = fma(-a, zv.x, rv.x); rv.y = fma(-a, zv.y, rv.y);
      rr = fma(rv.x, rv.x, rr); rr = fma(rv.y, rv.y, rr);
      *(double2*)(x + base + h) = xv; *(double2*)(r + base + h) = rv;
    }
  }
  for (long q = (long)blockIdx.x * blockDim.x + threadIdx.x; q < N; q += (long)gridDim.x * blockDim.x) g0[q] = 0.0;
  const double t = block_sum(rr, sh);
  if (cta_accumulate_last(t, ct + CT_ACC_RR, cnt + col * 4 + 2, gridDim.x)) {
    const double rrn = atomicAdd(ct + CT_ACC_RR, 0.0);
    ct[CT_ACC_RR] = 0.0;
    const double it = ct[CT_ITER] + 1.0, rel = sqrt(rrn) / ct[CT_NB];
    ct[CT_ITER] = it; ct[CT_RELRES] = rel;
    if (!(rel >= ct[CT_TOL])) ct[CT_DONE] = 1.0;
    else if (it + 1.0 > ct[CT_MAXIT]) ct[CT_DONE] = 2.0;
    else { ct[CT_BETA] = rrn / ct[CT_RRCUR]; ct[CT_RRCUR] = rrn; }
  }
}

// out = (K2 kron K1) T for a 2-D grid in one launch: CTA b computes output row b,
//   W[a'] = sum_b' K2[b][b'] T[b'][a'] ;  out[b][a] = sum_a' K1[a][a'] W[a']   (K1T = K1 transposed)
constexpr int KRON2_THREADS = 1024;
__global__ void __launch_bounds__(KRON2_THREADS) kron2_kernel(const double* __restrict__ K1T, const double* __restrict__ K2, int ng0, int ng1,
    const double* __restrict__ T, double* __restrict__ out, const double* __restrict__ ctl, long col_stride) {
  extern __shared__ double ksh[];                      // [nparts][ng0] partials + [ng0] W
  const int col = blockIdx.y;
  if (ctl && ctl[col * CT_STRIDE + CT_DONE] != 0.0) return;
  T += (long)col * col_stride; out += (long)col * col_stride;   // column = blockIdx.y (right-hand sides of a batched solve, probes)
  const int b = blockIdx.x;
  int TA = (ng0 + 31) & ~31; if (TA > 256) TA = 256;
  const int nparts = KRON2_THREADS / TA, part = threadIdx.x / TA, ta = threadIdx.x - part * TA;
  double* W = ksh + (long)nparts * ng0;
  if (part < nparts) {
    for (int a = ta; a < ng0; a += TA) {
      double acc = 0.0;
#pragma unroll 8
      for (int bp = part; bp < ng1; bp += nparts) acc = fma(K2[(long)b * ng1 + bp], T[(long)bp * ng0 + a], acc);
      ksh[(long)part * ng0 + a] = acc;
    }
  }
  __syncthreads();
  for (int a = threadIdx.x; a < ng0; a += KRON2_THREADS) {
    double acc = 0.0;
    for (int p2 = 0; p2 < nparts; ++p2) acc += ksh[(long)p2 * ng0 + a];
    W[a] = acc;
  }
  __syncthreads();
  if (part < nparts) {
    for (int a = ta; a < ng0; a += TA) {
      double acc = 0.0;
#pragma unroll 8
      for (int ap = part; ap < ng0; ap += nparts) acc = fma(K1T[(long)ap * ng0 + a], W[ap], acc);
      ksh[(long)part * ng0 + a] = acc;
    }
  }
  __syncthreads();
  for (int a = threadIdx.x; a < ng0; a += KRON2_THREADS) {
    double acc = 0.0;
    for (int p2 = 0; p2 < nparts; ++p2) acc += ksh[(long)p2 * ng0 + a];
    out[(long)b * ng0 + a] = acc;
  }
}

__global__ void transpose_small_kernel(const double* __restrict__ A, int ng, double* __restrict__ At) {
  long q = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= (long)ng * ng) return;
  int a = (int)(q / ng), b = (int)(q % ng);
  At[(long)b * ng + a] = A[q];
}

// alpha[perm[i]] = scale * x[i]
__global__ void cg_unsort_kernel(const double* __restrict__ x, const int* __restrict__ perm, long n, double scale,
                                 double* __restrict__ out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[perm[i]] = scale * x[i];
}


// ======================================================================================================
// Persistent cooperative CG: the WHOLE solve (apx.m:281-302) in one launch, one CTA per SM.
// Every CTA owns a fixed contiguous range of 128-point units of the cell-sorted points.  LEVEL 1 keeps
// the coordinates U of its range in shared memory for the whole solve; LEVEL 2 (single right-hand side)
// also keeps d and z there, so an iteration moves only x and r (40 B/point) through L2/HBM instead of
// the 16P+88 B/point of the streaming formulation.  Phases are separated by grid-wide barriers; the
// dot products are per-CTA partials summed in a fixed order by every CTA (deterministic, identical on
// all CTAs), so no control data has to be broadcast.
// ======================================================================================================
constexpr int CGP_THREADS = 512;
constexpr int CG_MAXCOLS_DEV = 64;
constexpr int CGP_WSCR = 16 * 33 + 16;              // doubles of per-warp scratch: 16 node sums x 33 (padded) + 32 int offsets
struct CgPersist {
  GridDesc G;
  const double* U; long n, npad; int m, upc;          // upc = units (128 points) per CTA
  double *x, *r, *d, *z;                               // [m][npad], sorted order
  double *g0, *ug;                                     // [m][N] grid vectors: Mx'(sW d) and Kg g0
  const double *K0, *K0T, *K1;                         // factor matrices (K0 = first grid axis; K0T its transpose; K1 second axis)
  double *ctl; double *part_dz, *part_rr;              // control blocks [m][CT_STRIDE]; partials [m][gridDim]
  unsigned* bar;                                       // grid barrier counter
  long long* dbg;                                      // optional: clock64 ticks per phase (CTA 0), 8 slots
};

// sense-free monotonically counting grid barrier (all CTAs co-resident: cooperative launch)
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& phase) {
  __syncthreads();
  if (threadIdx.x == 0) {
    phase += gridDim.x;
    __threadfence();
    atomicAdd(bar, 1u);
    while (*(volatile unsigned*)bar < phase) { }
    __threadfence();
  }
  __syncthreads();
}

template <int P, int LEVEL>
__global__ void __launch_bounds__(CGP_THREADS, 1) cg_persistent_kernel(const CgPersist A) {
  constexpr int NW = CgNodes<P>::NW;
  extern __shared__ __align__(16) double psm[];
  __shared__ double s_red[32];
  __shared__ double s_beta[CG_MAXCOLS_DEV], s_a[CG_MAXCOLS_DEV], s_rr[CG_MAXCOLS_DEV], s_nb[CG_MAXCOLS_DEV], s_it[CG_MAXCOLS_DEV], s_rel[CG_MAXCOLS_DEV];
  __shared__ int s_done[CG_MAXCOLS_DEV];
  __shared__ int s_any;
  const GridDesc& G = A.G;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = CGP_THREADS >> 5;
  const int m = A.m, cta = blockIdx.x, nctas = gridDim.x;
  const long nunits = A.npad / (32 * CG_PTS);
  const long u_lo = (long)cta * A.upc, u_hi = (u_lo + A.upc < nunits) ? u_lo + A.upc : nunits;
  const int nu = (u_hi > u_lo) ? (int)(u_hi - u_lo) : 0;
  // shared layout: [kron scratch][sU nu*128*P][sd nu*128][sz nu*128]
  int TA = (G.ng[0] + 31) & ~31; if (TA > CGP_THREADS) TA = CGP_THREADS;
  const int kparts = CGP_THREADS / TA;
  double* kscr = psm;
  double* wscr = psm + (((long)(kparts + 1) * G.ng[0] + 1) & ~1L) + (long)warp * CGP_WSCR;   // this warp's reduction scratch
  int* woff = (int*)(wscr + NW * 33);
  double* sU = psm + (((long)(kparts + 1) * G.ng[0] + 1) & ~1L) + (long)nwarp * CGP_WSCR;
  double* sd = sU + (LEVEL >= 1 ? (long)A.upc * 128 * P : 0);
  double* sz = sd + (LEVEL >= 2 ? (long)A.upc * 128 : 0);
  unsigned phase = 0;
  const double sw = A.ctl[CT_SW], tol = A.ctl[CT_TOL], maxit = A.ctl[CT_MAXIT];
  if (tid < m) {
    const double* ct = A.ctl + tid * CT_STRIDE;
    s_beta[tid] = 0.0; s_a[tid] = 0.0; s_rr[tid] = ct[CT_RRCUR]; s_nb[tid] = ct[CT_NB]; s_it[tid] = ct[CT_ITER];
    s_rel[tid] = ct[CT_RELRES]; s_done[tid] = (int)ct[CT_DONE];
  }
  if (LEVEL >= 1)
    for (long q = tid; q < (long)nu * 128 * P; q += CGP_THREADS) sU[q] = A.U[u_lo * 128 * P + q];
  if (LEVEL >= 2)
    for (long q = tid; q < (long)nu * 128; q += CGP_THREADS) sd[q] = A.d[u_lo * 128 + q];
  __syncthreads();

  long long tk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t0 = 0;
#define CGP_TICK(slot) do { if (A.dbg && cta == 0 && tid == 0) { long long t1 = clock64(); tk[slot] += t1 - t0; t0 = t1; } } while (0)
  if (A.dbg && cta == 0 && tid == 0) t0 = clock64();
  while (true) {
    if (tid == 0) { int any = 0; for (int c = 0; c < m; ++c) any |= (s_done[c] == 0); s_any = any; }
    __syncthreads();
    if (!s_any) break;
    // ---- A: d = r + beta d ; g0 += Mx'(sw d)
    for (int col = 0; col < m; ++col) {
      if (s_done[col]) continue;
      const double beta = s_beta[col];
      const double* r = A.r + (long)col * A.npad; double* d = A.d + (long)col * A.npad; double* g0 = A.g0 + (long)col * G.N;
      for (int ul = warp; ul < nu; ul += nwarp) {
        const long base = (u_lo + ul) * 128 + lane * CG_PTS; const int lb = ul * 128 + lane * CG_PTS;
        double dn[CG_PTS], uu[CG_PTS * P];
        {
          const double2 r0 = *(const double2*)(r + base), r1 = *(const double2*)(r + base + 2);
          double2 d0, d1;
          if (LEVEL >= 2) { d0 = *(const double2*)(sd + lb); d1 = *(const double2*)(sd + lb + 2); }
          else { d0 = *(const double2*)(d + base); d1 = *(const double2*)(d + base + 2); }
          dn[0] = fma(beta, d0.x, r0.x); dn[1] = fma(beta, d0.y, r0.y); dn[2] = fma(beta, d1.x, r1.x); dn[3] = fma(beta, d1.y, r1.y);
          if (LEVEL >= 2) { *(double2*)(sd + lb) = make_double2(dn[0], dn[1]); *(double2*)(sd + lb + 2) = make_double2(dn[2], dn[3]); }
          else { *(double2*)(d + base) = make_double2(dn[0], dn[1]); *(double2*)(d + base + 2) = make_double2(dn[2], dn[3]); }
#pragma unroll
          for (int q = 0; q < CG_PTS * P / 2; ++q) {
            const double2 t = (LEVEL >= 1) ? *((const double2*)(sU + (long)lb * P) + q) : __ldg((const double2*)(A.U + base * P) + q);
            uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
          }
        }
        double acc[NW];
#pragma unroll
        for (int v = 0; v < NW; ++v) acc[v] = 0.0;
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
              for (int v = 0; v < NW; ++v) acc[v] = 0.0;
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
        // segmented reduction through this warp's scratch: lane (k, h) walks 16 source lanes of node k
        {
          const int off_cur = (P == 1) ? cur : (cur / G.ng[0]) * (int)G.stride[1] + (cur % G.ng[0]);
          __syncwarp();
          woff[lane] = off_cur;
#pragma unroll
          for (int v = 0; v < NW; ++v) wscr[v * 33 + lane] = acc[v];
          __syncwarp();
          constexpr int LPN = 32 / NW;                       // lanes per node value: 2 (P=2) or 8 (P=1)
          constexpr int SRC = 32 / LPN;                      // source lanes each of them walks
          const int k = lane % NW, h = lane / NW;
          const int node_off = (P == 1) ? k : (k >> 2) * (int)G.stride[1] + (k & 3);
          double run = 0.0; int po = woff[h * SRC];
#pragma unroll
          for (int i2 = 0; i2 < SRC; ++i2) {
            const int co = woff[h * SRC + i2];
            const double v = wscr[k * 33 + h * SRC + i2];
            if (co != po) { if (run != 0.0) atomicAdd(g0 + po + node_off, run); run = 0.0; po = co; }
            run += v;
          }
          if (run != 0.0) atomicAdd(g0 + po + node_off, run);
        }
      }
    }
    CGP_TICK(0);
    grid_barrier(A.bar, phase);
    CGP_TICK(1);
    // ---- B: ug = Kg g0
    for (int col = 0; col < m; ++col) {
      if (s_done[col]) continue;
      const double* T = A.g0 + (long)col * G.N; double* out = A.ug + (long)col * G.N;
      if (P == 1) {
        const int ng = G.ng[0];
        for (int a = cta * nwarp + warp; a < ng; a += nctas * nwarp) {
          double acc = 0.0;
          for (int b = lane; b < ng; b += 32) acc = fma(A.K0[(long)a * ng + b], __ldcg(T + b), acc);
          acc = warp_sum(acc);
          if (lane == 0) out[a] = acc;
        }
      } else {
        const int ng0 = G.ng[0], ng1 = G.ng[1];
        const int part = tid / TA, ta = tid - part * TA;
        double* W = kscr + (long)kparts * ng0;
        for (int b = cta; b < ng1; b += nctas) {
          if (part < kparts)
            for (int a = ta; a < ng0; a += TA) {
              double acc = 0.0;
#pragma unroll 8
              for (int bp = part; bp < ng1; bp += kparts) acc = fma(A.K1[(long)b * ng1 + bp], __ldcg(T + (long)bp * ng0 + a), acc);
              kscr[(long)part * ng0 + a] = acc;
            }
          __syncthreads();
          for (int a = tid; a < ng0; a += CGP_THREADS) {
            double acc = 0.0;
            for (int p2 = 0; p2 < kparts; ++p2) acc += kscr[(long)p2 * ng0 + a];
            W[a] = acc;
          }
          __syncthreads();
          if (part < kparts)
            for (int a = ta; a < ng0; a += TA) {
              double acc = 0.0;
#pragma unroll 8
              for (int ap = part; ap < ng0; ap += kparts) acc = fma(A.K0T[(long)ap * ng0 + a], W[ap], acc);
              kscr[(long)part * ng0 + a] = acc;
            }
          __syncthreads();
          for (int a = tid; a < ng0; a += CGP_THREADS) {
            double acc = 0.0;
            for (int p2 = 0; p2 < kparts; ++p2) acc += kscr[(long)p2 * ng0 + a];
            out[(long)b * ng0 + a] = acc;
          }
          __syncthreads();
        }
      }
    }
    CGP_TICK(2);
    grid_barrier(A.bar, phase);
    CGP_TICK(3);
    // ---- C: z = sw Mx ug + d ; d.z
    for (int col = 0; col < m; ++col) {
      if (s_done[col]) continue;
      const double* d = A.d + (long)col * A.npad; double* z = A.z + (long)col * A.npad; const double* ug = A.ug + (long)col * G.N;
      double dz = 0.0;
      for (int ul = warp; ul < nu; ul += nwarp) {
        const long base = (u_lo + ul) * 128 + lane * CG_PTS; const int lb = ul * 128 + lane * CG_PTS;
        double dv[CG_PTS], uu[CG_PTS * P], zv[CG_PTS];
        {
          double2 d0, d1;
          if (LEVEL >= 2) { d0 = *(const double2*)(sd + lb); d1 = *(const double2*)(sd + lb + 2); }
          else { d0 = *(const double2*)(d + base); d1 = *(const double2*)(d + base + 2); }
          dv[0] = d0.x; dv[1] = d0.y; dv[2] = d1.x; dv[3] = d1.y;
#pragma unroll
          for (int q = 0; q < CG_PTS * P / 2; ++q) {
            const double2 t = (LEVEL >= 1) ? *((const double2*)(sU + (long)lb * P) + q) : __ldg((const double2*)(A.U + base * P) + q);
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
              for (int a = 0; a < 4; ++a) nv[a] = __ldcg(ug + j[0] + a);
            } else {
#pragma unroll
              for (int b = 0; b < 4; ++b) {
                const double* row = ug + (long)(j[P - 1] + b) * G.stride[1] + j[0];
#pragma unroll
                for (int a = 0; a < 4; ++a) nv[b * 4 + a] = __ldcg(row + a);
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
          const double zi = (base + q < A.n) ? fma(sw, mu, dv[q]) : 0.0;
          zv[q] = zi; dz = fma(dv[q], zi, dz);
        }
        if (LEVEL >= 2) { *(double2*)(sz + lb) = make_double2(zv[0], zv[1]); *(double2*)(sz + lb + 2) = make_double2(zv[2], zv[3]); }
        else { *(double2*)(z + base) = make_double2(zv[0], zv[1]); *(double2*)(z + base + 2) = make_double2(zv[2], zv[3]); }
      }
      const double t = block_sum(dz, s_red);
      if (tid == 0) A.part_dz[(long)col * nctas + cta] = t;
    }
    CGP_TICK(4);
    grid_barrier(A.bar, phase);
    CGP_TICK(5);
    for (int col = warp; col < m; col += nwarp) {
      if (s_done[col]) continue;
      double acc = 0.0;
      for (int i = lane; i < nctas; i += 32) acc += __ldcg(A.part_dz + (long)col * nctas + i);
      acc = warp_sum(acc);
      if (lane == 0) s_a[col] = s_rr[col] / acc;
    }
    __syncthreads();
    // ---- D: x += a d ; r -= a z ; r.r ; g0 = 0
    for (int col = 0; col < m; ++col) {
      if (s_done[col]) continue;
      const double a = s_a[col];
      double* x = A.x + (long)col * A.npad; double* r = A.r + (long)col * A.npad;
      const double* d = A.d + (long)col * A.npad; const double* z = A.z + (long)col * A.npad;
      double rr = 0.0;
      for (int ul = warp; ul < nu; ul += nwarp) {
        const long base = (u_lo + ul) * 128 + lane * CG_PTS; const int lb = ul * 128 + lane * CG_PTS;
#pragma unroll
        for (int h = 0; h < CG_PTS; h += 2) {
          double2 xv = *(const double2*)(x + base + h), rv = *(const double2*)(r + base + h);
          const double2 dv = (LEVEL >= 2) ? *(const double2*)(sd + lb + h) : *(const double2*)(d + base + h);
          const double2 zv = (LEVEL >= 2) ? *(const double2*)(sz + lb + h) : *(const double2*)(z + base + h);
          xv.x = fma(a, dv.x, xv.x); xv.y = fma(a, dv.y, xv.y);
          rv.x = fma(-a, zv.x, rv.x); rv.y = fma(-a, zv.y, rv.y);
          rr = fma(rv.x, rv.x, rr); rr = fma(rv.y, rv.y, rr);
          *(double2*)(x + base + h) = xv; *(double2*)(r + base + h) = rv;
        }
      }
      double* g0 = A.g0 + (long)col * G.N;
      for (long q = (long)cta * CGP_THREADS + tid; q < G.N; q += (long)nctas * CGP_THREADS) g0[q] = 0.0;
      const double t = block_sum(rr, s_red);
      if (tid == 0) A.part_rr[(long)col * nctas + cta] = t;
    }
    CGP_TICK(6);
    grid_barrier(A.bar, phase);
    CGP_TICK(7);
    for (int col = warp; col < m; col += nwarp) {
      if (s_done[col]) continue;
      double acc = 0.0;
      for (int i = lane; i < nctas; i += 32) acc += __ldcg(A.part_rr + (long)col * nctas + i);
      acc = warp_sum(acc);
      if (lane == 0) {
        const double it = s_it[col] + 1.0, rel = sqrt(acc) / s_nb[col];
        s_it[col] = it; s_rel[col] = rel;
        if (!(rel >= tol)) s_done[col] = 1;                     // apx.m:295-298
        else if (it + 1.0 > maxit) s_done[col] = 2;            // loop `for iter=2:maxit` exhausted
        else { s_beta[col] = acc / s_rr[col]; s_rr[col] = acc; }
      }
    }
    __syncthreads();
  }
  if (A.dbg && cta == 0 && tid == 0) for (int q = 0; q < 8; ++q) A.dbg[q] = tk[q];
  if (cta == 0 && tid < m) {
    double* ct = A.ctl + tid * CT_STRIDE;
    ct[CT_DONE] = (double)s_done[tid]; ct[CT_ITER] = s_it[tid]; ct[CT_RELRES] = s_rel[tid];
  }
}


// ======================================================================================================
// Persistent CG, second formulation: TWO grid barriers per iteration.
//   F : x += a d ; r -= a z ; |r|^2 ; d = r + beta d ; g0 += Mx'(sw d)        (update fused into the scatter)
//   -- barrier --
//   C : u = (Kg g0) on the node rows this CTA's points touch (computed by the CTA itself from g0 staged in
//       shared memory: no separate Kronecker phase, node values read from shared memory) ;
//       z = sw Mx u + d ; d.z, r.z, z.z
//   -- barrier --   every CTA sums the per-CTA partials in the same order:
//       a = |r|^2 / d.z ;  |r_next|^2 = |r|^2 - 2 a r.z + a^2 z.z  (exact identity for r_next = r - a z) gives the
//       convergence test and beta = |r_next|^2 / |r|^2 BEFORE the update is applied, so the update can ride in the
//       next F.  |r|^2 itself is always the directly summed value from F (no drift).
// Shared memory per CTA: coordinates U (LEVEL>=1), d and z (LEVEL 2, single right-hand side) of its own point
// range in a bank-conflict-free [double2][lane] layout, the node window, and one staging region (aliased onto z
// at LEVEL 2) used in turn for the scatter's segmented reduction, g0 and the first-axis factor.
// ======================================================================================================
// Row-sharded solve over several GPUs (SURVEY 8(e) KISS-GP (ii)): every rank owns n/P of the points and the whole grid.
// The two reductions of a CG iteration cross the ranks INSIDE the persistent kernel, over peer memory (NVLink), as a
// one-shot all-reduce in the style of NCCL's LL protocol: every FP64 value travels as two 8-byte words
// {32 data bits, 32-bit tag}; an 8-byte store is atomic, so a word whose tag matches carries valid data and no fence,
// flag or second round trip is needed -- the receiver polls the data itself.
//   after the scatter:  CTA j stores this rank's partial grid vector (N values) into peer rank+1+j's window;
//                       the staging step of phase C sums the P partials in rank order (identical on all ranks);
//   after phase C:      CTA 0 stores the 4 partial sums into every peer's window; every CTA sums them in rank order.
// All ranks therefore take bit-identical decisions and leave the loop in the same pass.  Tags (epoch + pass) grow
// monotonically over all solves and the windows are double-buffered by pass parity; the all-reduce of |b|^2 at the
// start of every solve keeps any rank from running a whole solve ahead.
struct CgDist {
  int nranks, rank; unsigned long long epoch; long ncap;
  char* loc; char* peer[8]; int* abort;
};
// window layout (16-byte packets): grid partials [2][8][ncap], scalars [2][8][4], locally combined grid vector [2][ncap]
constexpr int DX_NPUSH = 4;        // CTAs that share the push of one peer's copy
constexpr int DX_NCOMB = 16;       // CTAs (from the top of the grid) that sum the P partials of a slice
__device__ __forceinline__ ulonglong2* dx_rx(char* base, long ncap, int par, int r) { return (ulonglong2*)base + ((long)par * 8 + r) * ncap; }
__device__ __forceinline__ ulonglong2* dx_sc(char* base, long ncap, int par, int r) { return (ulonglong2*)base + 16 * ncap + (par * 8 + r) * 4; }
__device__ __forceinline__ ulonglong2* dx_gc(char* base, long ncap, int par) { return (ulonglong2*)base + 16 * ncap + 64 + (long)par * ncap; }
__device__ __forceinline__ void dx_put(ulonglong2* dst, double v, unsigned tag) {
  const unsigned long long b = (unsigned long long)__double_as_longlong(v), t = (unsigned long long)tag << 32;
  ulonglong2 w; w.x = t | (b & 0xffffffffULL); w.y = t | (b >> 32);
  asm volatile("st.volatile.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(w.x), "l"(w.y) : "memory");
}
// poll one packet until both halves carry `tag`; a peer that never arrives traps the kernel after ~17 s instead of
// hanging the GPU
// ---- exchange helpers, kept out of line so their registers do not add to the persistent kernel's budget -------
__device__ __forceinline__ bool dx_ok(unsigned long long x, unsigned long long y, unsigned tag) {
  return (unsigned)(x >> 32) == tag && (unsigned)(y >> 32) == tag;
}
__device__ __forceinline__ void dx_ld(const ulonglong2* p, unsigned long long& x, unsigned long long& y) {
  asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(p) : "memory");
}
__device__ __forceinline__ double dx_val(unsigned long long x, unsigned long long y) {
  return __longlong_as_double((long long)((x & 0xffffffffULL) | (y << 32)));
}
// re-poll one packet until it carries `tag` (x, y hold the first attempt)
__device__ __forceinline__ double dx_finish(const ulonglong2* p, unsigned long long x, unsigned long long y, unsigned tag) {
  long long t0 = 0; bool timing = false;
  while (!dx_ok(x, y, tag)) {
    if (!timing) { t0 = clock64(); timing = true; }
    else if (clock64() - t0 > (1LL << 35)) __trap();
    dx_ld(p, x, y);
  }
  return dx_val(x, y);
}
// dst[q] = packet(src[q]) for q in [q0, q1): four loads in flight, then four 16-byte stores
__device__ __noinline__ void dx_push_slice(const long long* __restrict__ shi, const long long* __restrict__ slo, FxScale f, ulonglong2* dst, long q0, long q1,
                                          unsigned tag, int tid, int nthreads) {
  for (long qb = q0 + tid; qb < q1; qb += 4L * nthreads) {
    double v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const long q = qb + (long)k * nthreads; v[k] = (q < q1) ? fx_load(shi, slo, q, f) : 0.0; }
#pragma unroll
    for (int k = 0; k < 4; ++k) { const long q = qb + (long)k * nthreads; if (q < q1) dx_put(dst + q, v[k], tag); }
  }
}
// out[q] = own[q] (+) pk[q] summed in rank order for q in [0, N); one peer (two ranks): no combiner stage
__device__ __noinline__ void dx_stage_pair(const long long* __restrict__ ohi, const long long* __restrict__ olo, FxScale f, const ulonglong2* pk, bool own_first,
                                          double* out, long N, unsigned tag, int tid, int nthreads) {
  for (long qb = tid; qb < N; qb += 4L * nthreads) {
    unsigned long long x[4], y[4]; double o[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const long q = qb + (long)k * nthreads; if (q < N) { dx_ld(pk + q, x[k], y[k]); o[k] = fx_load(ohi, olo, q, f); } }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const long q = qb + (long)k * nthreads;
      if (q < N) { const double pv = dx_finish(pk + q, x[k], y[k], tag); out[q] = own_first ? o[k] + pv : pv + o[k]; }
    }
  }
}
// out[q] = value of packet pk[q] for q in [0, N)
__device__ __noinline__ void dx_stage_packets(const ulonglong2* pk, double* out, long N, unsigned tag, int tid, int nthreads) {
  for (long qb = tid; qb < N; qb += 4L * nthreads) {
    unsigned long long x[4], y[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const long q = qb + (long)k * nthreads; if (q < N) dx_ld(pk + q, x[k], y[k]); }
#pragma unroll
    for (int k = 0; k < 4; ++k) { const long q = qb + (long)k * nthreads; if (q < N) out[q] = dx_finish(pk + q, x[k], y[k], tag); }
  }
}
// gc[q] = packet(sum over ranks, rank order) for q in [q0, q1): polls the peers' packets of the slice
__device__ __noinline__ void dx_combine_slice(const long long* __restrict__ ohi, const long long* __restrict__ olo, FxScale f, char* loc, long ncap, int par,
                                             int nranks, int rank, long q0, long q1, unsigned tag, int tid, int nthreads) {
  ulonglong2* gc = (ulonglong2*)loc + 16 * ncap + 64 + (long)par * ncap;
  for (long q = q0 + tid; q < q1; q += nthreads) {
    unsigned long long x[8], y[8];
#pragma unroll
    for (int rk = 0; rk < 8; ++rk) if (rk < nranks && rk != rank) dx_ld((const ulonglong2*)loc + ((long)par * 8 + rk) * ncap + q, x[rk], y[rk]);
    const double o = fx_load(ohi, olo, q, f);
    double sacc = 0.0;
#pragma unroll
    for (int rk = 0; rk < 8; ++rk)
      if (rk < nranks) sacc += (rk == rank) ? o : dx_finish((const ulonglong2*)loc + ((long)par * 8 + rk) * ncap + q, x[rk], y[rk], tag);
    dx_put(gc + q, sacc, tag);
  }
}
__device__ __forceinline__ double dx_get(const ulonglong2* src, unsigned tag) {
  unsigned long long x, y; long long t0 = 0; bool timing = false;
  while (true) {
    asm volatile("ld.volatile.global.v2.u64 {%0, %1}, [%2];" : "=l"(x), "=l"(y) : "l"(src) : "memory");
    if ((unsigned)(x >> 32) == tag && (unsigned)(y >> 32) == tag) break;
    if (!timing) { t0 = clock64(); timing = true; }
    else if (clock64() - t0 > (1LL << 35)) __trap();
  }
  return __longlong_as_double((long long)((x & 0xffffffffULL) | (y << 32)));
}

struct CgPersist2 {
  CgDist X;
  GridDesc G;
  const double* U; long n, npad; int m, upc, wcap;     // wcap = node rows the window buffer can hold
  double *x, *r, *d, *z;                               // [m][npad], sorted order
  long long *fxa, *fxb;                                // [m][2][N] fixed-point scatter targets (hi, lo), alternating per iteration
  long npts_bound;                                     // points this launch scatters
  double c2max2;                                       // max over nodes of sum_i c_i^2 (bound of the accumulation, see fx_scale)
  const double *K0, *K0T, *K1;
  double *ctl; double *part;                           // partials [4][m][gridDim]: |r|^2, d.z, r.z, z.z
  unsigned* bar; long long* dbg;
  long stg_doubles;                                    // size of the staging region
};
#ifndef KGP_CGP2_THREADS
#define KGP_CGP2_THREADS 512
#endif
constexpr int CGP2_THREADS = KGP_CGP2_THREADS;
constexpr int CGP_VPP = 8;                             // node values per reduction pass (2-D); 1-D uses 4
constexpr int CGP_WSCR2 = CGP_VPP * 33 + 16;

__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;\n" ::: "memory"); }

// per-CTA node-row window [rlo, rhi] of a sorted point range (rows of the LAST grid axis; nodes for 1-D)
template <int P>
__global__ void cg_window_kernel(GridDesc G, const double* __restrict__ U, long npad, int upc, int* __restrict__ wmax) {
  const long nunits = npad / 128;
  const long u_lo = (long)blockIdx.x * upc, u_hi = (u_lo + upc < nunits) ? u_lo + upc : nunits;
  if (threadIdx.x != 0 || u_hi <= u_lo) return;
  int jf, jl; double c[4];
  decode_u(U[(u_lo * 128) * P + (P - 1)], G.ng[P - 1], jf, c);
  decode_u(U[(u_hi * 128 - 1) * P + (P - 1)], G.ng[P - 1], jl, c);
  atomicMax(wmax, jf + 3 - jl + 1);
}

#if KGP_CGP2_THREADS > 576
#define KGP_CGP2_BOUNDS __maxnreg__(96)
#elif KGP_CGP2_THREADS > 512
#define KGP_CGP2_BOUNDS __maxnreg__(112)
#else
#define KGP_CGP2_BOUNDS __launch_bounds__(CGP2_THREADS, 1)
#endif
// FX: order-independent fixed-point scatter (bit-reproducible solves, one integer RED per flush); !FX: FP64 atomics (results differ
// in the last bits from run to run).  Without FX the (hi) array holds plain doubles.
template <int P, int LEVEL, bool DBG, bool DIST, bool FX>
__global__ void KGP_CGP2_BOUNDS cg_persistent2_kernel(const CgPersist2 A) {
  constexpr int NW = CgNodes<P>::NW;
  constexpr int VPP = (P == 1) ? 4 : CGP_VPP, NPASS = NW / VPP;
  extern __shared__ __align__(16) double psm[];
  __shared__ double s_red[32];
  __shared__ double s_beta[CG_MAXCOLS_DEV], s_a[CG_MAXCOLS_DEV], s_rr[CG_MAXCOLS_DEV], s_nb[CG_MAXCOLS_DEV], s_it[CG_MAXCOLS_DEV], s_rel[CG_MAXCOLS_DEV];
  __shared__ int s_state[CG_MAXCOLS_DEV];              // 0 active, 3 final update pending, 1 converged, 2 exhausted
  __shared__ int s_final[CG_MAXCOLS_DEV];
  __shared__ int s_any;
  __shared__ double s_sum[4][CG_MAXCOLS_DEV];
  __shared__ double s_dn2[CG_MAXCOLS_DEV];             // bound of |d|^2 of the direction about to be scattered
  __shared__ FxScale s_fx[CG_MAXCOLS_DEV];
  const GridDesc& G = A.G;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nwarp = CGP2_THREADS >> 5;
  const int m = A.m, cta = blockIdx.x, nctas = gridDim.x;
  const long nunits = A.npad / 128;
  const long u_lo = (long)cta * A.upc, u_hi = (u_lo + A.upc < nunits) ? u_lo + A.upc : nunits;
  const int nu = (u_hi > u_lo) ? (int)(u_hi - u_lo) : 0;
  const int ng0 = G.ng[0], ngl = G.ng[P - 1];
  // shared layout: [uwin wcap*ng0 (2-D) | wcap (1-D)] [sU] [sd] [sz / staging]
  double* uwin = psm;
  double* sU = uwin + (((P == 1) ? (long)A.wcap : (long)A.wcap * ng0) + 1 & ~1L);
  double* sd = sU + (LEVEL >= 1 ? (long)A.upc * 128 * P : 0);
  double* sz = sd + (LEVEL >= 2 ? (long)A.upc * 128 : 0);
  double* stg = sz;                                    // staging aliases z (z is dead whenever staging is live)
  unsigned phase = 0;
  const double sw = A.ctl[CT_SW], tol = A.ctl[CT_TOL], maxit = A.ctl[CT_MAXIT];
  if (tid < m) {
    const double* ct = A.ctl + tid * CT_STRIDE;
    s_beta[tid] = 0.0; s_a[tid] = 0.0; s_rr[tid] = ct[CT_RRCUR]; s_nb[tid] = ct[CT_NB]; s_it[tid] = ct[CT_ITER];
    s_rel[tid] = ct[CT_RELRES]; s_state[tid] = (int)ct[CT_DONE]; s_final[tid] = 0;
    s_dn2[tid] = ct[CT_RRCUR];                         // first direction: d = r
  }
  // resident data in the [double2 index][lane] layout
  if (LEVEL >= 1)
    for (long q = tid; q < (long)nu * 128 * P; q += CGP2_THREADS) {
      const long ul = q / (128 * P); const int rem = (int)(q - ul * 128 * P), ln = rem / (CG_PTS * P), w = rem - ln * CG_PTS * P;
      sU[ul * 128 * P + ((w >> 1) * 32 + ln) * 2 + (w & 1)] = A.U[u_lo * 128 * P + q];
    }
  if (LEVEL >= 2)
    for (long q = tid; q < (long)nu * 128; q += CGP2_THREADS) {
      const long ul = q >> 7; const int rem = (int)(q & 127), ln = rem >> 2, w = rem & 3;
      sd[ul * 128 + ((w >> 1) * 32 + ln) * 2 + (w & 1)] = A.d[u_lo * 128 + q];
    }
  if (LEVEL >= 2)
    for (long q = tid; q < (long)nu * 128; q += CGP2_THREADS) sz[q] = 0.0;
  // node window of this CTA
  int rlo = 0, wr = 0;
  if (nu > 0) {
    int jf, jl; double c[4];
    decode_u(A.U[(u_lo * 128) * P + (P - 1)], ngl, jf, c);
    decode_u(A.U[(u_hi * 128 - 1) * P + (P - 1)], ngl, jl, c);
    rlo = jl; wr = jf + 3 - jl + 1;
  }
  __syncthreads();
  long long tk[DBG ? 8 : 1], t0 = 0;
  if (DBG) for (int q = 0; q < 8; ++q) tk[q] = 0;
#define CGP2_TICK(slot) do { if (DBG && cta == 0 && tid == 0) { long long t1 = clock64(); tk[slot] += t1 - t0; t0 = t1; } } while (0)
  if (DBG && cta == 0 && tid == 0) t0 = clock64();
  int par = 0;
  unsigned long long pass = 0;

  while (true) {
    if (tid == 0) { int any = 0; for (int c = 0; c < m; ++c) any |= (s_state[c] == 0 || s_state[c] == 3); s_any = any; }
    if (FX && tid < m) s_fx[tid] = fx_scale(4.0 * sw * sqrt(A.c2max2) * sqrt(fmax(s_dn2[tid], 0.0)));
    __syncthreads();
    if (!s_any) break;
    ++pass;
    long long* g0w = par ? A.fxb : A.fxa;              // scattered into now, read in C
    long long* g0z = par ? A.fxa : A.fxb;              // read in the previous C: zeroed in this C
    // ---- F1: x += a d ; r -= a z ; |r|^2 ; d = r + beta d     (streams x and r; z is dead afterwards)
    for (int col = 0; col < m; ++col) {
      const int state = s_state[col];
      if (state != 0 && state != 3) continue;
      const double beta = s_beta[col], a = s_a[col];
      double* x = A.x + (long)col * A.npad; double* r = A.r + (long)col * A.npad;
      double* d = A.d + (long)col * A.npad; const double* z = A.z + (long)col * A.npad;
      double rr = 0.0;
      for (int ul0 = warp; ul0 < nu; ul0 += 2 * nwarp) {
        double2 xv[2][2], rv[2][2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ul = ul0 + h * nwarp;
          if (ul < nu) {
            const long base = (u_lo + ul) * 128 + lane * CG_PTS;
            xv[h][0] = *(const double2*)(x + base); xv[h][1] = *(const double2*)(x + base + 2);
            rv[h][0] = *(const double2*)(r + base); rv[h][1] = *(const double2*)(r + base + 2);
          }
        }
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ul = ul0 + h * nwarp;
          if (ul >= nu) continue;
          const long base = (u_lo + ul) * 128 + lane * CG_PTS;
          double2* sdu = (double2*)(sd + (long)ul * 128); const double2* szu = (const double2*)(sz + (long)ul * 128);
          double2 x0 = xv[h][0], x1 = xv[h][1], r0 = rv[h][0], r1 = rv[h][1];
          double2 d0, d1, z0, z1;
          if (LEVEL >= 2) { d0 = sdu[lane]; d1 = sdu[32 + lane]; z0 = szu[lane]; z1 = szu[32 + lane]; }
          else { d0 = *(const double2*)(d + base); d1 = *(const double2*)(d + base + 2); z0 = *(const double2*)(z + base); z1 = *(const double2*)(z + base + 2); }
          x0.x = fma(a, d0.x, x0.x); x0.y = fma(a, d0.y, x0.y); x1.x = fma(a, d1.x, x1.x); x1.y = fma(a, d1.y, x1.y);
          r0.x = fma(-a, z0.x, r0.x); r0.y = fma(-a, z0.y, r0.y); r1.x = fma(-a, z1.x, r1.x); r1.y = fma(-a, z1.y, r1.y);
          rr = fma(r0.x, r0.x, rr); rr = fma(r0.y, r0.y, rr); rr = fma(r1.x, r1.x, rr); rr = fma(r1.y, r1.y, rr);
          *(double2*)(x + base) = x0; *(double2*)(x + base + 2) = x1;
          *(double2*)(r + base) = r0; *(double2*)(r + base + 2) = r1;
          if (state == 3) continue;                        // final update only
          d0.x = fma(beta, d0.x, r0.x); d0.y = fma(beta, d0.y, r0.y); d1.x = fma(beta, d1.x, r1.x); d1.y = fma(beta, d1.y, r1.y);
          if (LEVEL >= 2) { sdu[lane] = d0; sdu[32 + lane] = d1; }
          else { *(double2*)(d + base) = d0; *(double2*)(d + base + 2) = d1; }
        }
      }
      const double t = block_sum(rr, s_red);             // (its barriers also order F1's z reads before F2's scratch writes)
      if (tid == 0) A.part[((long)0 * m + col) * nctas + cta] = t;
    }
    CGP2_TICK(5);
    // ---- F2: g0 += Mx'(sw d)     (shared memory only at LEVEL 2; the reduction scratch aliases z)
    for (int col = 0; col < m; ++col) {
      if (s_state[col] != 0) continue;
      const double* d = A.d + (long)col * A.npad;
      long long* ghi = g0w + (long)col * 2 * G.N; long long* glo = ghi + G.N; const FxScale fx = s_fx[col];
      double* wscr = stg + (long)warp * CGP_WSCR2; int* woff = (int*)(wscr + VPP * 33);
      for (int ul = warp; ul < nu; ul += nwarp) {
        const long base = (u_lo + ul) * 128 + lane * CG_PTS;
        const double2* sdu = (const double2*)(sd + (long)ul * 128);
        double dn[CG_PTS], uu[CG_PTS * P];
        {
          double2 d0, d1;
          if (LEVEL >= 2) { d0 = sdu[lane]; d1 = sdu[32 + lane]; }
          else { d0 = *(const double2*)(d + base); d1 = *(const double2*)(d + base + 2); }
          dn[0] = d0.x; dn[1] = d0.y; dn[2] = d1.x; dn[3] = d1.y;
#pragma unroll
          for (int q = 0; q < CG_PTS * P / 2; ++q) {
            const double2 t = (LEVEL >= 1) ? ((const double2*)(sU + (long)ul * 128 * P))[q * 32 + lane] : __ldg((const double2*)(A.U + base * P) + q);
            uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
          }
        }
        double acc[NW];
#pragma unroll
        for (int v = 0; v < NW; ++v) acc[v] = 0.0;
        int cur;
        {
          int cellq[CG_PTS];
#pragma unroll
          for (int q = 0; q < CG_PTS; ++q) {
            const int j0 = decode_j(uu[q * P], G.ng[0]);
            cellq[q] = (P == 1) ? j0 : j0 + decode_j(uu[q * P + P - 1], G.ng[P - 1]) * ng0;
          }
          cur = cellq[CG_PTS - 1];
          // two points at a time (register budget); points of this thread that fall in an EARLIER cell than its
          // last one go to the grid directly (rare)
#pragma unroll
          for (int q0 = 0; q0 < CG_PTS; q0 += 2) {
            int jj[2][P]; double cc[2][P][4];
#pragma unroll
            for (int h = 0; h < 2; ++h)
#pragma unroll
              for (int dd = 0; dd < P; ++dd) decode_u(uu[(q0 + h) * P + dd], G.ng[dd], jj[h][dd], cc[h][dd]);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              const int q = q0 + h;
              const double vi = sw * dn[q];
              if (cellq[q] == cur) {
                if (P == 1) {
#pragma unroll
                  for (int a2 = 0; a2 < 4; ++a2) acc[a2] = fma(cc[h][0][a2], vi, acc[a2]);
                } else {
#pragma unroll
                  for (int b = 0; b < 4; ++b) {
                    const double wb = cc[h][P - 1][b] * vi;
#pragma unroll
                    for (int a2 = 0; a2 < 4; ++a2) acc[b * 4 + a2] = fma(cc[h][0][a2], wb, acc[b * 4 + a2]);
                  }
                }
              } else {
                double one[NW];
                if (P == 1) {
#pragma unroll
                  for (int a2 = 0; a2 < 4; ++a2) one[a2] = cc[h][0][a2] * vi;
                } else {
#pragma unroll
                  for (int b = 0; b < 4; ++b) {
                    const double wb = cc[h][P - 1][b] * vi;
#pragma unroll
                    for (int a2 = 0; a2 < 4; ++a2) one[b * 4 + a2] = cc[h][0][a2] * wb;
                  }
                }
                if (FX) flush_nodes_fx<P>(G, ghi, glo, cellq[q], one, fx);
                else flush_nodes<P>(G, (double*)ghi, cellq[q], one);
              }
            }
          }
        }
        // segmented reduction through this warp's scratch, VPP node values per pass
        const int off_cur = (P == 1) ? cur : (cur / ng0) * (int)G.stride[1] + (cur % ng0);
        const int k = lane % VPP, h = lane / VPP;
#pragma unroll
        for (int ps = 0; ps < NPASS; ++ps) {
          __syncwarp();
          if (ps == 0) woff[lane] = off_cur;
#pragma unroll
          for (int v = 0; v < VPP; ++v) wscr[v * 33 + lane] = acc[ps * VPP + v];
          __syncwarp();
          const int kv = ps * VPP + k;
          const int node_off = (P == 1) ? kv : (kv >> 2) * (int)G.stride[1] + (kv & 3);
          double run = 0.0; int po = woff[h * VPP];
#pragma unroll
          for (int i2 = 0; i2 < VPP; ++i2) {
            const int co = woff[h * VPP + i2];
            const double v = wscr[k * 33 + h * VPP + i2];
            if (co != po) { if (run != 0.0) { if (FX) fx_add(ghi, glo, po + node_off, run, fx); else atomicAdd((double*)ghi + po + node_off, run); } run = 0.0; po = co; }
            run += v;
          }
          if (run != 0.0) { if (FX) fx_add(ghi, glo, po + node_off, run, fx); else atomicAdd((double*)ghi + po + node_off, run); }
        }
      }
    }
    CGP2_TICK(0);
    grid_barrier(A.bar, phase);
    if (DIST) {
      const unsigned tag = (unsigned)(A.X.epoch + pass);
      // (a) push: CTA (j, s), counted from the TOP of the grid (the last CTAs own the fewest points and reach the
      //     barrier first), stores slice s of this rank's partial grid vector into peer rank+1+j's window
      const int npush = (A.X.nranks - 1) * DX_NPUSH, rcta = nctas - 1 - cta;
      if (rcta < npush) {
        const int pr = (A.X.rank + 1 + rcta / DX_NPUSH) % A.X.nranks, sl = rcta % DX_NPUSH;
        const long per = (G.N + DX_NPUSH - 1) / DX_NPUSH, q0 = sl * per, q1 = (q0 + per < G.N) ? q0 + per : G.N;
        dx_push_slice(g0w, FX ? g0w + G.N : nullptr, s_fx[0], dx_rx(A.X.peer[pr], A.X.ncap, par, A.X.rank), q0, q1, tag, tid, CGP2_THREADS);
      }
      // (b) more than two ranks: the last DX_NCOMB CTAs sum the P partials of their slice in rank order (polling the
      //     peers' packets) and publish the result as packets in the LOCAL window -- no second grid barrier
      if (A.X.nranks > 2 && rcta >= npush && rcta < npush + DX_NCOMB) {
        const int sl = rcta - npush;
        const long per = (G.N + DX_NCOMB - 1) / DX_NCOMB, q0 = sl * per, q1 = (q0 + per < G.N) ? q0 + per : G.N;
        dx_combine_slice(g0w, FX ? g0w + G.N : nullptr, s_fx[0], A.X.loc, A.X.ncap, par, A.X.nranks, A.X.rank, q0, q1, tag, tid, CGP2_THREADS);
      }
    }
    CGP2_TICK(1);
    // ---- C: node window of u = Kg g0, then z = sw Mx u + d and the three dot products
    for (int col = 0; col < m; ++col) {
      if (s_state[col] != 0) continue;
      const long long* Thi = g0w + (long)col * 2 * G.N; const long long* Tlo = FX ? Thi + G.N : nullptr; const FxScale fx = s_fx[col];
      if (P == 1) {
        if (DIST) {
          if (A.X.nranks == 2) dx_stage_pair(Thi, Tlo, fx, dx_rx(A.X.loc, A.X.ncap, par, 1 - A.X.rank), A.X.rank == 0, stg, ng0, (unsigned)(A.X.epoch + pass), tid, CGP2_THREADS);
          else dx_stage_packets(dx_gc(A.X.loc, A.X.ncap, par), stg, ng0, (unsigned)(A.X.epoch + pass), tid, CGP2_THREADS);
        } else for (int q = tid; q < ng0; q += CGP2_THREADS) stg[q] = fx_load(Thi, Tlo, q, fx);
        __syncthreads();
        for (int a2 = warp; a2 < wr; a2 += nwarp) {
          double acc2 = 0.0;
          for (int b = lane; b < ng0; b += 32) acc2 = fma(A.K0[(long)(rlo + a2) * ng0 + b], stg[b], acc2);
          acc2 = warp_sum(acc2);
          if (lane == 0) uwin[a2] = acc2;
        }
        __syncthreads();
      } else {
        const int ng1 = G.ng[1];
        const long tb = (G.N > (long)ng0 * ng0) ? G.N : (long)ng0 * ng0;
        double* sT = stg; double* sK1 = stg + ((tb + 1) & ~1L); double* sW = sK1 + (((long)A.wcap * ng1 + 1) & ~1L);
        if (DIST) {                                                    // the combined grid vector
          if (A.X.nranks == 2) dx_stage_pair(Thi, Tlo, fx, dx_rx(A.X.loc, A.X.ncap, par, 1 - A.X.rank), A.X.rank == 0, sT, G.N, (unsigned)(A.X.epoch + pass), tid, CGP2_THREADS);
          else dx_stage_packets(dx_gc(A.X.loc, A.X.ncap, par), sT, G.N, (unsigned)(A.X.epoch + pass), tid, CGP2_THREADS);
        } else {
          for (long q = tid; q < G.N; q += CGP2_THREADS) sT[q] = fx_load(Thi, Tlo, q, fx);      // fixed point -> FP64 (an unrolled form measured slower)
        }
        for (int q = tid; q < wr * ng1; q += CGP2_THREADS) sK1[q] = A.K1[(long)rlo * ng1 + q];
        cp_async_wait_all();
        __syncthreads();
        for (int q = tid; q < wr * ng0; q += CGP2_THREADS) {           // W[bw][a'] = sum_b' K1[rlo+bw][b'] T[b'][a']
          const int bw = q / ng0, ap = q - bw * ng0;
          double a0 = 0.0, a1 = 0.0;
          int bp = 0;
          for (; bp + 1 < ng1; bp += 2) {
            a0 = fma(sK1[bw * ng1 + bp], sT[(long)bp * ng0 + ap], a0);
            a1 = fma(sK1[bw * ng1 + bp + 1], sT[(long)(bp + 1) * ng0 + ap], a1);
          }
          if (bp < ng1) a0 = fma(sK1[bw * ng1 + bp], sT[(long)bp * ng0 + ap], a0);
          sW[q] = a0 + a1;
        }
        __syncthreads();
        double* sK0T = stg;                                           // overwrites T (dead); W lies beyond it
        const long nk = (long)ng0 * ng0;
        for (long q = tid; q < nk / 2; q += CGP2_THREADS) cp_async16(sK0T + 2 * q, A.K0T + 2 * q);
        if ((nk & 1) && tid == 0) sK0T[nk - 1] = A.K0T[nk - 1];
        cp_async_wait_all();
        __syncthreads();
        for (int q = tid; q < wr * ng0; q += CGP2_THREADS) {           // u[bw][a] = sum_a' K0T[a'][a] W[bw][a']
          const int bw = q / ng0, a2 = q - bw * ng0;
          double a0 = 0.0, a1 = 0.0;
          int ap = 0;
          for (; ap + 1 < ng0; ap += 2) {
            a0 = fma(sK0T[(long)ap * ng0 + a2], sW[bw * ng0 + ap], a0);
            a1 = fma(sK0T[(long)(ap + 1) * ng0 + a2], sW[bw * ng0 + ap + 1], a1);
          }
          if (ap < ng0) a0 = fma(sK0T[(long)ap * ng0 + a2], sW[bw * ng0 + ap], a0);
          uwin[q] = a0 + a1;
        }
        __syncthreads();
      }
      CGP2_TICK(6);
      const double* d = A.d + (long)col * A.npad; double* z = A.z + (long)col * A.npad; const double* r = A.r + (long)col * A.npad;
      double dz = 0.0, rz = 0.0, zz = 0.0;
      for (int ul = warp; ul < nu; ul += nwarp) {
        const long base = (u_lo + ul) * 128 + lane * CG_PTS;
        const double2* sdu = (const double2*)(sd + (long)ul * 128); double2* szu = (double2*)(sz + (long)ul * 128);
        double dv[CG_PTS], rv[CG_PTS], uu[CG_PTS * P], zv[CG_PTS];
        {
          const double2 r0 = *(const double2*)(r + base), r1 = *(const double2*)(r + base + 2);
          rv[0] = r0.x; rv[1] = r0.y; rv[2] = r1.x; rv[3] = r1.y;
          double2 d0, d1;
          if (LEVEL >= 2) { d0 = sdu[lane]; d1 = sdu[32 + lane]; }
          else { d0 = *(const double2*)(d + base); d1 = *(const double2*)(d + base + 2); }
          dv[0] = d0.x; dv[1] = d0.y; dv[2] = d1.x; dv[3] = d1.y;
#pragma unroll
          for (int q = 0; q < CG_PTS * P / 2; ++q) {
            const double2 t = (LEVEL >= 1) ? ((const double2*)(sU + (long)ul * 128 * P))[q * 32 + lane] : __ldg((const double2*)(A.U + base * P) + q);
            uu[2 * q] = t.x; uu[2 * q + 1] = t.y;
          }
        }
#pragma unroll
        for (int q = 0; q < CG_PTS; ++q) {
          int j[P]; double c[P][4];
#pragma unroll
          for (int dd = 0; dd < P; ++dd) decode_u(uu[q * P + dd], G.ng[dd], j[dd], c[dd]);
          double mu = 0.0;
          if (P == 1) {
            const double* nv = uwin + (j[0] - rlo);
#pragma unroll
            for (int a2 = 0; a2 < 4; ++a2) mu = fma(c[0][a2], nv[a2], mu);
          } else {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              const double* nv = uwin + (long)(j[P - 1] + b - rlo) * ng0 + j[0];
              double rb = 0.0;
#pragma unroll
              for (int a2 = 0; a2 < 4; ++a2) rb = fma(c[0][a2], nv[a2], rb);
              mu = fma(c[P - 1][b], rb, mu);
            }
          }
          const double zi = (base + q < A.n) ? fma(sw, mu, dv[q]) : 0.0;
          zv[q] = zi; dz = fma(dv[q], zi, dz); rz = fma(rv[q], zi, rz); zz = fma(zi, zi, zz);
        }
        if (LEVEL >= 2) { szu[lane] = make_double2(zv[0], zv[1]); szu[32 + lane] = make_double2(zv[2], zv[3]); }
        else { *(double2*)(z + base) = make_double2(zv[0], zv[1]); *(double2*)(z + base + 2) = make_double2(zv[2], zv[3]); }
      }
      long long* gz = g0z + (long)col * 2 * G.N;
      for (long q = (long)cta * CGP2_THREADS + tid; q < G.N; q += (long)nctas * CGP2_THREADS) gz[q] = 0;     // (only the first of the two words per node is in use)
      double t = block_sum(dz, s_red);
      if (tid == 0) A.part[((long)1 * m + col) * nctas + cta] = t;
      t = block_sum(rz, s_red);
      if (tid == 0) A.part[((long)2 * m + col) * nctas + cta] = t;
      t = block_sum(zz, s_red);
      if (tid == 0) A.part[((long)3 * m + col) * nctas + cta] = t;
    }
    CGP2_TICK(2);
    grid_barrier(A.bar, phase);
    CGP2_TICK(3);
    for (int job = warp; job < 4 * m; job += nwarp) {         // one warp per (quantity, column): fixed summation order
      const int w = job / m, col = job - w * m, state = s_state[col];
      if (!((state == 0) || (state == 3 && w == 0))) continue;
      double acc2 = 0.0;
      for (int i = lane; i < nctas; i += 32) acc2 += __ldcg(A.part + ((long)w * m + col) * nctas + i);
      acc2 = warp_sum(acc2);
      if (lane == 0) s_sum[w][col] = acc2;
    }
    __syncthreads();
    if (DIST) {                                                // the four sums cross the ranks (m == 1)
      const unsigned tag = (unsigned)(A.X.epoch + pass);
      if (cta == 0 && tid < 4 * A.X.nranks) {
        const int pr = tid >> 2, w = tid & 3;
        if (pr != A.X.rank) dx_put(dx_sc(A.X.peer[pr], A.X.ncap, par, A.X.rank) + w, s_sum[w][0], tag);
      }
      __syncthreads();
      if (tid < 4) {
        double tot = 0.0;
        for (int rk = 0; rk < A.X.nranks; ++rk) tot += (rk == A.X.rank) ? s_sum[tid][0] : dx_get(dx_sc(A.X.loc, A.X.ncap, par, rk) + tid, tag);
        s_sum[tid][0] = tot;
      }
      __syncthreads();
    }
    for (int col = tid; col < m; col += CGP2_THREADS) {
      const int state = s_state[col];
      if (state != 0 && state != 3) continue;
      const double rr = s_sum[0][col];
      if (state == 3) { s_rel[col] = sqrt(rr) / s_nb[col]; s_state[col] = s_final[col]; }
      else {
        const double a = rr / s_sum[1][col];
        const double rrn = fma(a * a, s_sum[3][col], fma(-2.0 * a, s_sum[2][col], rr));
        const double it = s_it[col] + 1.0, rel = sqrt(fmax(rrn, 0.0)) / s_nb[col];
        s_a[col] = a; s_it[col] = it; s_rel[col] = rel;
        if (!(rel >= tol)) { s_state[col] = 3; s_final[col] = 1; }            // apx.m:295-298
        else if (it + 1.0 > maxit) { s_state[col] = 3; s_final[col] = 2; }    // `for iter=2:maxit` exhausted
        else { const double beta = rrn / rr; s_beta[col] = beta; s_dn2[col] = fmax(rrn, 0.0) + beta * beta * s_dn2[col]; }
      }
    }
    par ^= 1;
    __syncthreads();
    CGP2_TICK(4);
  }
  if (DBG && cta == 0 && tid == 0) for (int q = 0; q < 8; ++q) A.dbg[q] = tk[q];
  if (cta == 0 && tid < m) {
    double* ct = A.ctl + tid * CT_STRIDE;
    ct[CT_DONE] = (double)s_state[tid]; ct[CT_ITER] = s_it[tid]; ct[CT_RELRES] = s_rel[tid];
    if (tid == 0) ct[CT_PASSES] = (double)pass;
  }
}

// after the cross-rank all-reduce of |b|^2 (in CT_RRCUR): redo the scalar set-up of cg_init_sorted_kernel
__global__ void cg_dist_fix_ctl_kernel(double* __restrict__ ct) {
  const double rr = ct[CT_RRCUR], nb = sqrt(rr), rel = nb / nb;
  ct[CT_NB] = nb; ct[CT_RELRES] = rel;
  ct[CT_DONE] = (rel >= ct[CT_TOL]) ? ((2.0 > ct[CT_MAXIT]) ? 2.0 : 0.0) : 1.0;
}

}  // namespace kgp
