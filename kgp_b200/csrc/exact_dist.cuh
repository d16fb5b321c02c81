// Multi-GPU exact GP: one process per GPU, NCCL over NVLink, ONE n x n problem sharded.
//
//   potrf : 1-D block-cyclic block COLUMNS (NB = 512).  The owner factors panel k on a panel stream, packs
//           it and ncclBroadcasts it; every rank updates the block columns it owns with DMMA GEMMs (one batched
//           launch per step).  Look-ahead 1: the owner of panel k+1 updates that block column first and factors
//           it while the rest of the trailing update runs.  Every rank unpacks the panels, so all hold the full L.
//   Z = L^-T : rows of Z are independent -> contiguous row ranges (cube-root split: equal flops), right-looking,
//           no communication; then one broadcast per rank gathers Z everywhere.
//   B^-1 = Z Z^T : lower 128-tiles dealt round-robin over ranks in the launch's tile order (own_mod/own_rem).
//   gradient pass : each rank reduces ITS tiles of B^-1 (symmetric: a tile feeds row i and row j sums),
//           then ncclAllReduce of the n x (D+3) row sums.  alpha, log-det, scalars: redundantly on every rank.
// NCCL is bound at run time (dlopen of the libnccl.so.2 that PyTorch ships); the library has no link-time
// dependency on it and single-GPU use never loads it.
#pragma once
#include "ctx.h"
#include <dlfcn.h>

namespace kgp {

typedef struct { char internal[128]; } kgp_ncclUniqueId;
struct NcclApi {
  void* h = nullptr; bool ok = false; std::string err;
  int (*GetUniqueId)(kgp_ncclUniqueId*) = nullptr;
  int (*CommInitRank)(void**, int, kgp_ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*Broadcast)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*AllReduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
};
constexpr int KGP_NCCL_DOUBLE = 8, KGP_NCCL_SUM = 0, KGP_NCCL_INT8 = 0, KGP_NCCL_INT32 = 2, KGP_NCCL_MIN = 3;   // ncclFloat64, ncclSum, ncclInt8, ncclInt32, ncclMin (nccl.h enum values)
// peer exchange window (see grid_cg.cuh CgDist), 16-byte packets: [2][8][PX_NCAP] grid partials, [2][8][4] scalars,
// [2][PX_NCAP] locally combined grid vector
constexpr long PX_NCAP = 32768;
constexpr size_t PX_BYTES = (size_t)(16 * PX_NCAP + 64 + 2 * PX_NCAP) * 16;

static NcclApi& nccl_api(const char* libpath) {
  static NcclApi api;
  if (api.ok) return api;
  const char* cands[] = {libpath, getenv("KGP_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
  for (const char* p : cands) {
    if (!p || !*p) continue;
    api.h = dlopen(p, RTLD_NOW | RTLD_GLOBAL);
    if (api.h) break;
  }
  if (!api.h) { api.err = std::string("cannot load libnccl: ") + (dlerror() ? dlerror() : "?"); return api; }
#define KGP_NCCL_SYM(field, name) *(void**)(&api.field) = dlsym(api.h, name); if (!api.field) { api.err = "missing symbol " name; return api; }
  KGP_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
  KGP_NCCL_SYM(CommInitRank, "ncclCommInitRank")
  KGP_NCCL_SYM(CommDestroy, "ncclCommDestroy")
  KGP_NCCL_SYM(Broadcast, "ncclBroadcast")
  KGP_NCCL_SYM(AllReduce, "ncclAllReduce")
  KGP_NCCL_SYM(GroupStart, "ncclGroupStart")
  KGP_NCCL_SYM(GroupEnd, "ncclGroupEnd")
  KGP_NCCL_SYM(GetErrorString, "ncclGetErrorString")
#undef KGP_NCCL_SYM
  api.ok = true;
  return api;
}

#define KGP_NCCL(ctx, api, call)                                                                   \
  do { int r__ = (call);                                                                           \
       if (r__ != 0) return (ctx)->fail(KGP_CUDA_ERR, "%s:%d NCCL %s: %s", __FILE__, __LINE__, #call, (api).GetErrorString(r__)); \
  } while (0)

// ---- ownership maps (pure host logic; exported through the C ABI for the CPU tests) ---------------------
// block column k of the Cholesky belongs to rank k mod P
inline int dist_col_owner(int block, int nranks) { return block % nranks; }
// Z row ranges: boundaries (in units of NB rows) that equalise sum_{i in range} (n-i)^2  ->  a_r = n (1 - (1 - r/P)^(1/3))
inline long dist_zrow_begin(int r, int nranks, long np, int nb) {
  if (r <= 0) return 0;
  if (r >= nranks) return np;
  const double f = 1.0 - cbrt(1.0 - (double)r / (double)nranks);
  long a = (long)llround(f * (double)np / (double)nb) * nb;
  if (a > np) a = np;
  return a;
}


// KISS-GP row-sharded CG: contiguous shares of the data rows; sharded only when every rank gets >= 4096 points
// (below that the exchange latency exceeds the point work) and the grid fits the peer window
inline long dist_row_begin(int r, int nranks, long n) { return r <= 0 ? 0 : (r >= nranks ? n : n * r / nranks); }
inline bool dist_rows_sharded(int nranks, long n, long N) { return nranks > 1 && nranks <= 8 && n >= 4096L * nranks && N <= PX_NCAP; }

// ---- gradient pass over the lower 128-tiles of B^-1 a rank owns ------------------------------------------
// Same arithmetic per element as reduce_grad_kernel (G = (B^-1/sn2 - a_i a_j) K_ij with K recomputed from xs),
// but a tile (tm > tn) feeds BOTH the row sums of its rows i (over its columns j) and, by symmetry, the row sums
// of its columns j (over its rows i).  Tiles are enumerated exactly like the Z Z^T launch that produced them.
// rows_out[i][0..DP+2] (+)= [sum g, sum g d2, sum B^-1 K, sum_j g (xs_i - xs_j)_d]   (atomics: partial per rank)
template <int DP>
__global__ void __launch_bounds__(128) reduce_grad_tile_kernel(const GemmArgs g, long ntiles, const double* __restrict__ W, long ld,
    long n, const double* __restrict__ xs, const double* __restrict__ alpha, double sf2, double inv_sn2,
    double* __restrict__ rows_out) {
  extern __shared__ double rsm[];
  constexpr int WD = DP + 3;
  double (*sB)[33] = (double (*)[33])rsm;                     // [128][33] strip of the tile
  double* sxi = rsm + 128 * 33;                               // [128][DP]
  double* sxj = sxi + 128 * DP;                               // [32][DP]
  double* sai = sxj + 32 * DP;                                // [128]
  double* saj = sai + 128;                                    // [32]
  double* scol = saj + 32;                                    // [4][32][WD]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    int tm, tn;
    gemm_tile_coords(g, g.own_mod > 1 ? t * g.own_mod + g.own_rem : t, tm, tn);
    const long m0 = (long)tm * 128, n0 = (long)tn * 128;
    const double* Tp; long tld;
    if (g.Cdiag && (m0 / g.nb) == (n0 / g.nb)) {
      const long blk = m0 / g.nb;
      Tp = g.Cdiag + blk * g.nb * g.nb + (m0 - blk * g.nb) * g.nb + (n0 - blk * g.nb); tld = g.nb;
    } else { Tp = W + m0 * ld + n0; tld = ld; }
    __syncthreads();
    for (int q = tid; q < 128 * DP; q += 128) sxi[q] = xs[m0 * DP + q];
    sai[tid] = alpha[m0 + tid];
    const long i = m0 + tid;
    double xi[DP], gd[DP];
#pragma unroll
    for (int d = 0; d < DP; ++d) { xi[d] = xs[i * DP + d]; gd[d] = 0.0; }
    const double ai = alpha[i];
    double r = 0.0, rd2 = 0.0, rk = 0.0;
    for (int s = 0; s < 4; ++s) {
      const long j0 = n0 + s * 32;
      __syncthreads();
      for (int rr = 0; rr < 32; ++rr) sB[warp * 32 + rr][lane] = Tp[(long)(warp * 32 + rr) * tld + s * 32 + lane];
      for (int q = tid; q < 32 * DP; q += 128) sxj[q] = xs[j0 * DP + q];
      if (tid < 32) saj[tid] = alpha[j0 + tid];
      __syncthreads();
      if (j0 < n) {
        const int jn = (n - j0 < 32) ? (int)(n - j0) : 32;
        if (i < n) {
          for (int j = 0; j < jn; ++j) {
            double df[DP], d2 = 0.0;
#pragma unroll
            for (int d = 0; d < DP; ++d) { df[d] = xi[d] - sxj[j * DP + d]; d2 = fma(df[d], df[d], d2); }
            const double k = sf2 * exp(-0.5 * d2);
            const double bij = sB[tid][j];
            const double gg = (bij * inv_sn2 - ai * saj[j]) * k;
            r += gg; rd2 = fma(gg, d2, rd2); rk = fma(bij, k, rk);
#pragma unroll
            for (int d = 0; d < DP; ++d) gd[d] = fma(gg, df[d], gd[d]);
          }
        }
        if (tm != tn) {                                       // symmetric part: column sums of this strip
          const int jc = lane, part = warp;
          double cr = 0.0, crd2 = 0.0, crk = 0.0, cgd[DP];
#pragma unroll
          for (int d = 0; d < DP; ++d) cgd[d] = 0.0;
          if (jc < jn) {
            double xj[DP];
#pragma unroll
            for (int d = 0; d < DP; ++d) xj[d] = sxj[jc * DP + d];
            const double aj = saj[jc];
            for (int ii = 0; ii < 32; ++ii) {
              const int il = part * 32 + ii;
              if (m0 + il >= n) break;
              double df[DP], d2 = 0.0;
#pragma unroll
              for (int d = 0; d < DP; ++d) { df[d] = xj[d] - sxi[il * DP + d]; d2 = fma(df[d], df[d], d2); }
              const double k = sf2 * exp(-0.5 * d2);
              const double bij = sB[il][jc];
              const double gg = (bij * inv_sn2 - aj * sai[il]) * k;
              cr += gg; crd2 = fma(gg, d2, crd2); crk = fma(bij, k, crk);
#pragma unroll
              for (int d = 0; d < DP; ++d) cgd[d] = fma(gg, df[d], cgd[d]);
            }
          }
          double* sc = scol + ((long)part * 32 + jc) * WD;
          sc[0] = cr; sc[1] = crd2; sc[2] = crk;
#pragma unroll
          for (int d = 0; d < DP; ++d) sc[3 + d] = cgd[d];
          __syncthreads();
          for (int q = tid; q < jn * WD; q += 128) {
            const int jj = q / WD, cc = q - jj * WD;
            const double v = scol[((long)0 * 32 + jj) * WD + cc] + scol[((long)1 * 32 + jj) * WD + cc] +
                             scol[((long)2 * 32 + jj) * WD + cc] + scol[((long)3 * 32 + jj) * WD + cc];
            atomicAdd(rows_out + (j0 + jj) * WD + cc, v);
          }
        }
      }
    }
    if (i < n) {
      double* p = rows_out + i * WD;
      atomicAdd(p + 0, r); atomicAdd(p + 1, rd2); atomicAdd(p + 2, rk);
#pragma unroll
      for (int d = 0; d < DP; ++d) atomicAdd(p + 3 + d, gd[d]);
    }
  }
}
template <int DP> constexpr size_t reduce_tile_smem() { return (size_t)(128 * 33 + 128 * DP + 32 * DP + 128 + 32 + 4 * 32 * (DP + 3)) * 8; }

}  // namespace kgp
