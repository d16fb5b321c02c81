// Latency-oriented variant of the NT GEMM for the reference-sized problems (n <= 2048: every example the reference
// ships, kgp/configs + examples/*.py, runs the exact path at n = 512 ... 1024).
//
// One 128x128x128 tile is 4.2 MFLOP, i.e. 17 us of ONE SM's DMMA pipe; a launch that has only a handful of such
// tiles (the trailing update of a 512-order Cholesky: 6, 3, 1 tiles; Z Z^T: 10 tiles with K up to 512) leaves 140 SMs
// idle and costs 26 ... 74 us.  Here a work item is a 32x32 sub-tile x one K chunk of 128: hundreds of items of
// ~0.26 MFLOP spread over all SMs.  Items of one sub-tile whose K range spans several chunks leave their partial
// sums in a workspace; the LAST one to arrive (a counter per sub-tile) adds the partials in chunk order -- the result
// does not depend on the arrival order -- and runs the epilogue.  Semantics (shape, ktri, alpha/beta, Cdiag
// redirect) are those of dgemm_nt_kernel_t; batched / rank-sharded launches stay on the big kernel.
#pragma once
#include "dgemm_nt.cuh"

namespace kgp {

constexpr int SG_T = 32, SG_KC = 128, SG_THREADS = 128;
constexpr int SG_SMEM_BYTES = 2 * SG_T * SG_KC * 8;       // 64 KB: A and B chunk, [k4-step][row][4]

__global__ void __launch_bounds__(SG_THREADS) dgemm_nt_small_kernel(const GemmArgs g, const long ntiles, const int nchmax,
                                                                    double* __restrict__ ws, unsigned* __restrict__ cnt) {
  extern __shared__ __align__(16) double sg_smem[];
  __shared__ int s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gq = lane >> 2, tq = lane & 3;
  const int wm = warp >> 1, wn = warp & 1;
  const long item = blockIdx.x; const int sub = (int)(item & 15), chunk = blockIdx.y;
  GemmTile t;
  gemm_fetch_tile(g, item >> 4, ntiles, t);
  const int K = t.nk * GEMM_BK, nch = (K + SG_KC - 1) / SG_KC;
  if (chunk >= nch) return;
  const int k0 = chunk * SG_KC, klen = (K - k0 < SG_KC) ? K - k0 : SG_KC, cpr = klen >> 1;   // 16-byte chunks per row
  const int sm = sub >> 2, sn = sub & 3;
  double* sA = sg_smem; double* sB = sg_smem + SG_T * SG_KC;
  const double* Ag = t.Ag + (long)(sm * SG_T) * g.lda + k0;
  const double* Bg = t.Bg + (long)(sn * SG_T) * g.ldb + k0;
  for (int q = tid; q < SG_T * cpr; q += SG_THREADS) {
    const int row = q / cpr, ch = q - row * cpr;
    const int so = ((ch >> 1) * SG_T + row) * 4 + (ch & 1) * 2;
    cp_async16(sA + so, Ag + (long)row * g.lda + ch * 2);
    cp_async16(sB + so, Bg + (long)row * g.ldb + ch * 2);
  }
  cp_async_commit();
  cp_async_wait<0>();
  __syncthreads();
  double acc[2][2][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  const double* pa = sA + (wm * 16 + gq) * 4 + tq;
  const double* pb = sB + (wn * 16 + gq) * 4 + tq;
  const int nstep = klen >> 2;
#pragma unroll 4
  for (int s = 0; s < nstep; ++s) {
    const double a0 = pa[s * (SG_T * 4)], a1 = pa[s * (SG_T * 4) + 32];
    const double b0 = pb[s * (SG_T * 4)], b1 = pb[s * (SG_T * 4) + 32];
    dmma884(acc[0][0][0], acc[0][0][1], a0, b0);
    dmma884(acc[0][1][0], acc[0][1][1], a0, b1);
    dmma884(acc[1][0][0], acc[1][0][1], a1, b0);
    dmma884(acc[1][1][0], acc[1][1][1], a1, b1);
  }
  if (nch > 1) {
    double2* wp = reinterpret_cast<double2*>(ws + ((long)item * nchmax + chunk) * (SG_T * SG_T));
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) __stcg(wp + (i * 2 + j) * SG_THREADS + tid, make_double2(acc[i][j][0], acc[i][j][1]));
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const unsigned old = atomicAdd(cnt + item, 1u);
      s_last = (old == (unsigned)(nch - 1));
      if (s_last) cnt[item] = 0;                          // self-resetting: the next launch finds zeros
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const double2* rp = reinterpret_cast<const double2*>(ws + (long)item * nchmax * (SG_T * SG_T));
#pragma unroll
    for (int i = 0; i < 2; ++i)
#pragma unroll
      for (int j = 0; j < 2; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
    for (int c = 0; c < nch; ++c)                         // fixed order: independent of which chunk arrived last
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const double2 v = __ldcg(rp + (long)c * (SG_T * SG_T / 2) + (i * 2 + j) * SG_THREADS + tid);
          acc[i][j][0] += v.x; acc[i][j][1] += v.y;
        }
  }
  double* Cp; long ldc; long crow0, ccol0;
  if (t.in_diag) {
    const int blk = t.m0 / g.nb;
    Cp = g.Cdiag + (long)blk * g.nb * g.nb; ldc = g.nb; crow0 = t.m0 - blk * g.nb; ccol0 = t.n0 - blk * g.nb;
  } else { Cp = g.C + t.coff; ldc = g.ldc; crow0 = t.m0; ccol0 = t.n0; }
  double* cbase = Cp + (crow0 + sm * SG_T + wm * 16 + gq) * ldc + ccol0 + sn * SG_T + wn * 16 + tq * 2;
  const double alpha = g.alpha, beta = g.beta;
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      double2* p = reinterpret_cast<double2*>(cbase + (long)i * 8 * ldc + j * 8);
      double2 v = make_double2(alpha * acc[i][j][0], alpha * acc[i][j][1]);
      if (beta != 0.0) { const double2 cv = __ldcg(p); v.x = beta * cv.x + v.x; v.y = beta * cv.y + v.y; }
      *p = v;
    }
}

// workspace a launch needs (bytes) / whether the small kernel should take it
inline bool gemm_small_eligible(const GemmArgs& g, long nt, int& nchmax, size_t& ws_bytes) {
  if (g.nbatch > 0 || g.own_mod > 1 || g.tile_cnt > 0 || nt <= 0 || nt > 96) return false;
  nchmax = (g.kend - g.kbeg + SG_KC - 1) / SG_KC;
  if (nchmax <= 0 || nt * nchmax > 512) return false;
  ws_bytes = nchmax > 1 ? (size_t)nt * 16 * nchmax * SG_T * SG_T * 8 : 0;
  return true;
}
inline void gemm_small_init() {
  static bool done = false;
  if (done) return;
  cudaFuncSetAttribute(dgemm_nt_small_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SG_SMEM_BYTES);
  done = true;
}
// g must already carry band_base (gemm_prepare); cnt: >= nt*16 zeroed counters
inline void gemm_launch_small(const GemmArgs& g, long nt, int nchmax, double* ws, unsigned* cnt, cudaStream_t st) {
  gemm_small_init();
  dim3 grid((unsigned)(nt * 16), (unsigned)nchmax);
  dgemm_nt_small_kernel<<<grid, SG_THREADS, SG_SMEM_BYTES, st>>>(g, nt, nchmax, ws, cnt);
}

}  // namespace kgp
