// FP64 tensor-core GEMM  C = beta*C + alpha * A * B^T   ("NT": both operands K-contiguous)
//
// This is the one dense-contraction kernel of the exact-GP path.  Every O(n^3)
// step of the blocked Cholesky / triangular inverse / inverse product
// (GPML/cov/apx.m:160-163: chol, solve_chol(L,eye(n))) is phrased so that both
// operands are read along their contiguous K axis out of row-major storage, so a
// single kernel serves them all (see exact.cu for the call sites).
//
// sm_100a has no FP64 kind for tcgen05.mma; the FP64 tensor primitive is the
// warp-level DMMA.8x8x4 (PTX mma.sync.aligned.m8n8k4.f64).  Layout of one CTA:
//   128x128 C tile, 256 threads = 8 warps as 2(m) x 4(n), warp tile 64x32 =
//   8x4 DMMA tiles, 64 FP64 accumulators per thread.
//   K is consumed in steps of BK=16 through a 4-stage cp.async ring in shared
//   memory (32 KB per stage).  A stage is stored [k4-step][row][4]: the four K
//   values of one DMMA k-step are contiguous per row, so a thread's fragment
//   element is sA[(step*128 + row)*4 + tq] and a warp's LDS.64 touches 8 rows x
//   4 consecutive doubles = 32 distinct banks-pairs (conflict-free, no padding).
// All extents are multiples of 128 (rows/cols) and 16 (K): the engine pads the
// matrix order to a multiple of 128 with an identity block, so there are no
// bounds checks in the hot loop.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

namespace kgp {

enum : int { GEMM_FULL = 0, GEMM_LOWER = 1 };
enum : int { KTRI_NONE = 0, KTRI_FROM_M0 = 1 };

constexpr int GEMM_MAX_BATCH = 96;
constexpr int GEMM_MAX_BANDS = 160;     // 160 bands x 8 rows x 128 = order 163 840
struct GemmArgs {
  const double* A; long lda;   // A[m][k], row m at A + m*lda
  const double* B; long ldb;   // B[n][k]
  double* C; long ldc;         // C[m][n]
  int ntm, ntn;                // tiles of 128 in m / n
  int kbeg, kend;              // K range (multiples of 16)
  int shape;                   // GEMM_FULL | GEMM_LOWER (only tiles with tn <= tm)
  int ktri;                    // KTRI_FROM_M0: k starts at kbeg + 128*tm (upper-triangular A operand)
  double alpha, beta;
  // optional redirect of C tiles that fall in a diagonal nb x nb block (ZZ^T step):
  double* Cdiag; int nb;       // if Cdiag && (m0/nb == n0/nb): C tile -> Cdiag[blk][m0%nb.., n0%nb..], ld = nb
  int cdiag_only;              // (repurposed) 1: prefetch the C tile into L2 at tile start
  int nbands; int band_base[GEMM_MAX_BANDS];   // LOWER shape: tile index at which each band of 8 tile-rows starts
  // multi-GPU sharding of one launch (exact_dist.cu):
  int own_mod, own_rem;        // own_mod > 1: this rank computes tiles idx = t*own_mod + own_rem only (cost-balanced: neighbours cost alike)
  long tile_off, tile_cnt;     // tile_cnt > 0: this launch covers only the tile_cnt owned tiles that follow global tile index tile_off
                               // (same global enumeration and ownership; staged launches of one problem: exact.cu zzt_reduce_dist)
  // batch of FULL problems that share ntn, K and the operand matrices but start at different rows
  // (trailing update of the block columns a rank owns): problem p covers rows >= bat_roff[p] of A and C and the
  // ntn tile columns whose B rows start at bat_roff[p]; C tile (tm,tn) of problem p sits at C[roff+tm*128][roff+tn*128]
  int nbatch; int bat_roff[GEMM_MAX_BATCH]; int bat_base[GEMM_MAX_BATCH + 1];
};

#ifndef KGP_GEMM_BK
#define KGP_GEMM_BK 16
#endif
#ifndef KGP_GEMM_STAGES
#define KGP_GEMM_STAGES 4
#endif
#ifndef KGP_GEMM_SPLIT
#define KGP_GEMM_SPLIT 1
#endif
constexpr int GEMM_BM = 128, GEMM_BN = 128, GEMM_BK = KGP_GEMM_BK, GEMM_STAGES = KGP_GEMM_STAGES;
constexpr int GEMM_KSTEPS = GEMM_BK / 4, GEMM_CHUNKS = GEMM_BK / 2;   // k4-steps / 16-byte chunks per row per stage
constexpr int GEMM_THREADS = 256;
constexpr int GEMM_STAGE_DOUBLES = (GEMM_BM + GEMM_BN) * GEMM_BK;           // 4096
constexpr int GEMM_SMEM_BYTES = GEMM_STAGES * GEMM_STAGE_DOUBLES * 8;       // 131072

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// number of tiles a launch must cover
inline long gemm_num_tiles(const GemmArgs& g) {
  if (g.nbatch > 0) {
    long t = 0;
    for (int p = 0; p < g.nbatch; ++p) t += (long)(g.ntm - g.bat_roff[p] / 128) * g.ntn;
    return t;
  }
  if (g.shape == GEMM_FULL) return (long)g.ntm * g.ntn;
  long t = 0;
  for (int tm = 0; tm < g.ntm; ++tm) t += (tm + 1 < g.ntn ? tm + 1 : g.ntn);
  return t;
}

// Tile rasterisation: tiles are issued in bands of GEMM_BAND tile-rows, column-major inside a
// band, so that the ~148 concurrently resident CTAs share a handful of A row-panels and walk the
// B row-panels together (L2 working set ~ (8+19) panels instead of ~150).  Bands go top to
// bottom, which also issues the longest tiles first when the K range is triangular.
constexpr int GEMM_BAND = 8;
__device__ __forceinline__ void gemm_tile_coords(const GemmArgs& g, long idx, int& tm, int& tn) {
  if (g.shape == GEMM_FULL) {
    const long per_band = (long)GEMM_BAND * g.ntn;
    const int b = (int)(idx / per_band);
    const long r = idx - (long)b * per_band;
    const int rows = (g.ntm - b * GEMM_BAND < GEMM_BAND) ? g.ntm - b * GEMM_BAND : GEMM_BAND;
    tn = (int)(r / rows); tm = b * GEMM_BAND + (int)(r % rows);
    return;
  }
  // LOWER: row tm holds min(tm+1, ntn) tiles; band_base[] (host-built) = tiles before each band
  int lo = 0, hi = g.nbands - 1;
  while (lo < hi) {                       // last band with band_base <= idx
    int mid = (lo + hi + 1) >> 1;
    if (g.band_base[mid] <= idx) lo = mid; else hi = mid - 1;
  }
  const int b = lo; const long base = g.band_base[b];
  const int r0 = b * GEMM_BAND;
  const int rows = (g.ntm - r0 < GEMM_BAND) ? g.ntm - r0 : GEMM_BAND;
  long r = idx - base;
  const int rect = (r0 < g.ntn) ? r0 : g.ntn;          // columns shared by every row of the band
  if (r < (long)rect * rows) { tn = (int)(r / rows); tm = r0 + (int)(r % rows); return; }
  r -= (long)rect * rows;
  for (int c = rect; c < g.ntn; ++c) {                  // triangular part: column c holds rows >= c
    const int first = (c > r0) ? c : r0;
    const int cnt = r0 + rows - first;
    if (cnt <= 0) break;
    if (r < cnt) { tn = c; tm = first + (int)r; return; }
    r -= cnt;
  }
  tm = r0; tn = 0;   // unreachable
}

// Tile sequence of a persistent CTA: round p covers tiles [p*G, (p+1)*G); even rounds hand tile p*G + b to CTA b,
// odd rounds run backwards (p*G + G-1-b).  Tile cost falls linearly along the issue order whenever the K range is
// triangular (Z-sweep, Z Z^T) -- with the plain stride CTA 0 would get the longest tile of EVERY round; the
// serpentine pairs a long tile with a short one, so all CTAs finish together (Z-sweep: 28.0 -> 31+ TFLOP/s).
__device__ __forceinline__ long gemm_seq_idx(long p) {
  return p * gridDim.x + ((p & 1) ? (long)(gridDim.x - 1 - blockIdx.x) : (long)blockIdx.x);
}

// One work item of the persistent loop: tile coordinates, K range and operand pointers.
struct GemmTile {
  int m0, n0, nk; bool in_diag;
  const double* Ag; const double* Bg;
  long coff;                   // offset of the problem's C origin (batched mode), else 0
  int arow, brow, kc;          // the same tile as coordinates relative to g.A / g.B (TMA path): first A row, first B row, first k
};
__device__ __forceinline__ bool gemm_fetch_tile(const GemmArgs& g, long idx, long ntiles, GemmTile& t) {
  if (idx < ntiles) {
    int tm, tn;
    if (g.nbatch > 0) {
      int lo = 0, hi = g.nbatch - 1;
      while (lo < hi) { int mid = (lo + hi + 1) >> 1; if (g.bat_base[mid] <= idx) lo = mid; else hi = mid - 1; }
      const int roff = g.bat_roff[lo], ntm = g.ntm - roff / 128;
      const long r0 = idx - g.bat_base[lo], per_band = (long)GEMM_BAND * g.ntn;
      const int b = (int)(r0 / per_band); const long r = r0 - (long)b * per_band;
      const int rows = (ntm - b * GEMM_BAND < GEMM_BAND) ? ntm - b * GEMM_BAND : GEMM_BAND;
      tn = (int)(r / rows); tm = b * GEMM_BAND + (int)(r % rows);
      t.m0 = tm * GEMM_BM; t.n0 = tn * GEMM_BN; t.in_diag = false;
      t.nk = (g.kend - g.kbeg) / GEMM_BK;
      t.Ag = g.A + (long)(roff + t.m0) * g.lda + g.kbeg;
      t.Bg = g.B + (long)(roff + t.n0) * g.ldb + g.kbeg;
      t.coff = (long)roff * g.ldc + roff;
      t.arow = roff + t.m0; t.brow = roff + t.n0; t.kc = g.kbeg;
      return true;
    }
    gemm_tile_coords(g, (g.own_mod > 1 ? idx * g.own_mod + g.own_rem : idx) + g.tile_off, tm, tn);
    t.coff = 0;
    t.m0 = tm * GEMM_BM; t.n0 = tn * GEMM_BN;
    t.in_diag = g.Cdiag && (t.m0 / g.nb) == (t.n0 / g.nb);
    const int kbeg = g.kbeg + (g.ktri == KTRI_FROM_M0 ? t.m0 : 0);
    t.nk = (g.kend - kbeg) / GEMM_BK;
    t.Ag = g.A + (long)t.m0 * g.lda + kbeg;
    t.Bg = g.B + (long)t.n0 * g.ldb + kbeg;
    t.arow = t.m0; t.brow = t.n0; t.kc = kbeg;
    return true;
  }
  return false;
}

// Persistent kernel: one CTA per SM walks its serpentine tile sequence (gemm_seq_idx).  The cp.async ring
// runs ahead ACROSS tile boundaries (the load cursor is up to 3 k-steps ahead of the compute
// cursor), so the pipeline fill of the next tile overlaps the epilogue of the current one.
// BN = 128: one 256-thread CTA per SM on 128x128 tiles.  BN = 64: the tile is split into two 128x64 halves, 128
// threads each, TWO CTAs per SM (same registers and warps per SM): while one CTA is in its epilogue (C read-modify-
// write, no DMMAs) the other keeps the tensor pipe busy.
template <int BN>
__global__ void __launch_bounds__(2 * BN, 128 / BN) dgemm_nt_kernel_t(const GemmArgs g, const long nitems) {
  constexpr int THREADS = 2 * BN, WN = BN / 32, HALVES = 128 / BN, STAGE_DOUBLES = (GEMM_BM + BN) * GEMM_BK, LPR = BN / 16;
  extern __shared__ __align__(16) double smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / WN, wn = warp % WN;       // 2 x WN warps
  const int gq = lane >> 2, tq = lane & 3;
  const long ntiles = nitems / HALVES;
  auto fetch = [&](long item, GemmTile& t) -> bool {
    if (item >= nitems) return false;
    gemm_fetch_tile(g, item / HALVES, ntiles, t);
    if (HALVES > 1) { const int h = (int)(item % HALVES); t.n0 += h * BN; t.Bg += (long)h * BN * g.ldb; }
    return true;
  };

  // --- load cursor
  GemmTile lt; long lpos = 0; int lkt = 0;
  bool lvalid = fetch(gemm_seq_idx(lpos), lt);
  auto load_next = [&](int stage) {
    if (lvalid) {
      double* sA = smem + stage * STAGE_DOUBLES;
      double* sB = sA + GEMM_BM * GEMM_BK;
      const long koff = (long)lkt * GEMM_BK;
#pragma unroll
      for (int i = 0; i < (GEMM_BM * GEMM_CHUNKS) / THREADS; ++i) {
        int q = tid + i * THREADS;
        int row = q / GEMM_CHUNKS, ch = q % GEMM_CHUNKS;
        int so = ((ch >> 1) * 128 + row) * 4 + (ch & 1) * 2;     // [k4-step][row][4]
        cp_async16(sA + so, lt.Ag + (long)row * g.lda + koff + ch * 2);
      }
#pragma unroll
      for (int i = 0; i < (BN * GEMM_CHUNKS) / THREADS; ++i) {
        int q = tid + i * THREADS;
        int row = q / GEMM_CHUNKS, ch = q % GEMM_CHUNKS;
        int so = ((ch >> 1) * BN + row) * 4 + (ch & 1) * 2;
        cp_async16(sB + so, lt.Bg + (long)row * g.ldb + koff + ch * 2);
      }
      if (++lkt == lt.nk) { lkt = 0; ++lpos; lvalid = fetch(gemm_seq_idx(lpos), lt); }
    }
    cp_async_commit();
  };
#pragma unroll
  for (int s = 0; s < GEMM_STAGES - 1; ++s) load_next(s);

  // --- compute cursor.  Fragments are double-buffered in registers at k4-step granularity: the
  // LDS.64 of step s+1 are issued before the 32 DMMAs of step s, and the block barrier sits before
  // the LAST step of a stage (where the first fragments of the next stage are fetched), so the
  // tensor pipe always has a full step of DMMAs queued while a warp waits.
  double fa[2][8], fb[2][4];
  auto lds = [&](int buf, const double* sA, int step) {
    const double* sB = sA + GEMM_BM * GEMM_BK;
#pragma unroll
    for (int i = 0; i < 8; ++i) fa[buf][i] = sA[(step * 128 + wm * 64 + i * 8 + gq) * 4 + tq];
#pragma unroll
    for (int j = 0; j < 4; ++j) fb[buf][j] = sB[(step * BN + wn * 32 + j * 8 + gq) * 4 + tq];
  };
  GemmTile ct; int it = 0;
  cp_async_wait<GEMM_STAGES - 2>();
  __syncthreads();
  lds(0, smem, 0);
  for (long cpos = 0; fetch(gemm_seq_idx(cpos), ct); ++cpos) {
    double* Cp; long ldc; long crow0, ccol0;
    if (ct.in_diag) {
      int blk = ct.m0 / g.nb;
      Cp = g.Cdiag + (long)blk * g.nb * g.nb; ldc = g.nb;
      crow0 = ct.m0 - blk * g.nb; ccol0 = ct.n0 - blk * g.nb;
    } else { Cp = g.C + ct.coff; ldc = g.ldc; crow0 = ct.m0; ccol0 = ct.n0; }
    const double alpha = g.alpha, beta = g.beta;
    const bool last_tile = gemm_seq_idx(cpos + 1) >= nitems;
    if (beta != 0.0 && g.cdiag_only) {   // pull the C tile towards L2 while the main loop runs
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int line = tid * 4 + i;
        const double* pp = Cp + (crow0 + line / LPR) * ldc + ccol0 + (line % LPR) * 16;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pp));
      }
    }
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int kt = 0; kt < ct.nk; ++kt, ++it) {
      const double* sA = smem + (it % GEMM_STAGES) * STAGE_DOUBLES;
#pragma unroll
      for (int step = 0; step < GEMM_KSTEPS; ++step) {
        const int cur = step & 1;
        if (step < GEMM_KSTEPS - 1) {
          lds(cur ^ 1, sA, step + 1);
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
        } else {
          // stage boundary: barrier, then half of this step's DMMAs are issued BEFORE the cp.async /
          // LDS bookkeeping so the tensor pipe has work queued while the warp does address math
          cp_async_wait<GEMM_STAGES - 3>();
          __syncthreads();
#if KGP_GEMM_SPLIT
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
#endif
          load_next((it + GEMM_STAGES - 1) % GEMM_STAGES);
          if (!(last_tile && kt + 1 == ct.nk))
            lds(cur ^ 1, smem + ((it + 1) % GEMM_STAGES) * STAGE_DOUBLES, 0);
#pragma unroll
          for (int i = (KGP_GEMM_SPLIT ? 4 : 0); i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
        }
      }
    }

    // --- epilogue (registers -> global; shared memory is not touched, the ring keeps filling)
    double* cbase = Cp + (crow0 + wm * 64 + gq) * ldc + ccol0 + wn * 32 + tq * 2;
    if (beta != 0.0) {
      // read-modify-write: batch the loads (16 x 16 B in flight per thread) before any store, otherwise
      // possible aliasing makes the compiler serialise 32 DRAM round trips per thread
#pragma unroll
      for (int ih = 0; ih < 2; ++ih) {
        double2 cv[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j)
            cv[i][j] = __ldcg(reinterpret_cast<const double2*>(cbase + (long)(ih * 4 + i) * 8 * ldc + j * 8));
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            double2 v;
            v.x = beta * cv[i][j].x + alpha * acc[ih * 4 + i][j][0];
            v.y = beta * cv[i][j].y + alpha * acc[ih * 4 + i][j][1];
            *reinterpret_cast<double2*>(cbase + (long)(ih * 4 + i) * 8 * ldc + j * 8) = v;
          }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
          *reinterpret_cast<double2*>(cbase + (long)i * 8 * ldc + j * 8) =
              make_double2(alpha * acc[i][j][0], alpha * acc[i][j][1]);
    }
  }
  cp_async_wait<0>();
}

static int g_gemm_num_sms = 148;
static int g_gemm_persist = 1, g_gemm_prefetch_c = 1, g_gemm_bn = 128;   // tuning knobs (env KGP_GEMM_PERSIST / KGP_GEMM_PREFETCHC / KGP_GEMM_BN)
inline cudaError_t gemm_init() {
  int dev = 0; cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&g_gemm_num_sms, cudaDevAttrMultiProcessorCount, dev);
  if (const char* e = getenv("KGP_GEMM_PERSIST")) g_gemm_persist = atoi(e);
  if (const char* e = getenv("KGP_GEMM_PREFETCHC")) g_gemm_prefetch_c = atoi(e);
  if (const char* e = getenv("KGP_GEMM_BN")) g_gemm_bn = atoi(e) == 64 ? 64 : 128;
  cudaError_t e1 = cudaFuncSetAttribute(dgemm_nt_kernel_t<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_STAGES * (GEMM_BM + 64) * GEMM_BK * 8);
  if (e1 != cudaSuccess) return e1;
  return cudaFuncSetAttribute(dgemm_nt_kernel_t<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES);
}

}  // namespace kgp
