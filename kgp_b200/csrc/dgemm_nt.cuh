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
//   memory (32 KB per stage).  A stage holds two "k-slabs" of 8 doubles per row,
//   stored [slab][row][8]; a thread fetches its A/B fragments with one LDS.128
//   that yields the operands of TWO consecutive DMMA k-steps: inside a slab the
//   K order is permuted (step0 uses k={0,2,4,6}, step1 uses k={1,3,5,7}) for A
//   and B alike, which leaves the dot product unchanged and makes the quarter-
//   warp LDS.128 pattern bank-conflict free without padding.
// All extents are multiples of 128 (rows/cols) and 16 (K): the engine pads the
// matrix order to a multiple of 128 with an identity block, so there are no
// bounds checks in the hot loop.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace kgp {

enum : int { GEMM_FULL = 0, GEMM_LOWER = 1 };
enum : int { KTRI_NONE = 0, KTRI_FROM_M0 = 1 };

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
  int cdiag_only;              // 1: only those tiles; -1: only the others; 0: all
};

constexpr int GEMM_BM = 128, GEMM_BN = 128, GEMM_BK = 16, GEMM_STAGES = 4;
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
  if (g.shape == GEMM_FULL) return (long)g.ntm * g.ntn;
  long t = 0;
  for (int tm = 0; tm < g.ntm; ++tm) t += (tm + 1 < g.ntn ? tm + 1 : g.ntn);
  return t;
}

__device__ __forceinline__ void gemm_tile_coords(const GemmArgs& g, long idx, int& tm, int& tn) {
  if (g.shape == GEMM_FULL) { tm = (int)(idx / g.ntn); tn = (int)(idx % g.ntn); return; }
  // rows tm < ntn hold tm+1 tiles, later rows hold ntn tiles
  long tri = (long)g.ntn * (g.ntn + 1) / 2;
  if (idx < tri) {
    int r = (int)((sqrt(8.0 * (double)idx + 1.0) - 1.0) * 0.5);
    while ((long)(r + 1) * (r + 2) / 2 <= idx) ++r;
    while ((long)r * (r + 1) / 2 > idx) --r;
    tm = r; tn = (int)(idx - (long)r * (r + 1) / 2);
  } else {
    long rem = idx - tri;
    tm = g.ntn + (int)(rem / g.ntn); tn = (int)(rem % g.ntn);
  }
}

__global__ void __launch_bounds__(GEMM_THREADS, 1) dgemm_nt_kernel(const GemmArgs g) {
  extern __shared__ __align__(16) double smem[];
  int tm, tn;
  gemm_tile_coords(g, blockIdx.x, tm, tn);
  const int m0 = tm * GEMM_BM, n0 = tn * GEMM_BN;

  double* Cp = g.C; long ldc = g.ldc;
  bool in_diag = false;
  if (g.Cdiag) {
    in_diag = (m0 / g.nb) == (n0 / g.nb);
    if (g.cdiag_only == 1 && !in_diag) return;
    if (g.cdiag_only == -1 && in_diag) return;
  }
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp >> 2, wn = warp & 3;        // 2 x 4 warps
  const int gq = lane >> 2, tq = lane & 3;

  int kbeg = g.kbeg + (g.ktri == KTRI_FROM_M0 ? m0 : 0);
  const int nk = (g.kend - kbeg) / GEMM_BK;

  double acc[8][4][2];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

  // --- global->shared copy mapping: 1024 16-byte chunks per operand per stage, 4 per thread
  const double* Ag = g.A + (long)m0 * g.lda + kbeg;
  const double* Bg = g.B + (long)n0 * g.ldb + kbeg;
  auto load_stage = [&](int stage, int kt) {
    double* sA = smem + stage * GEMM_STAGE_DOUBLES;
    double* sB = sA + GEMM_BM * GEMM_BK;
    const long koff = (long)kt * GEMM_BK;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int q = tid + i * GEMM_THREADS;
      int row = q >> 3, ch = q & 7;
      int so = ((ch >> 2) * 128 + row) * 8 + (ch & 3) * 2;
      cp_async16(sA + so, Ag + (long)row * g.lda + koff + ch * 2);
      cp_async16(sB + so, Bg + (long)row * g.ldb + koff + ch * 2);
    }
  };

#pragma unroll
  for (int s = 0; s < GEMM_STAGES - 1; ++s) {
    if (s < nk) load_stage(s, s);
    cp_async_commit();
  }

  for (int kt = 0; kt < nk; ++kt) {
    cp_async_wait<GEMM_STAGES - 2>();
    __syncthreads();
    {
      int nxt = kt + GEMM_STAGES - 1;
      if (nxt < nk) load_stage(nxt % GEMM_STAGES, nxt);
      cp_async_commit();
    }
    const double* sA = smem + (kt % GEMM_STAGES) * GEMM_STAGE_DOUBLES;
    const double* sB = sA + GEMM_BM * GEMM_BK;
#pragma unroll
    for (int slab = 0; slab < 2; ++slab) {
      double2 af[8], bf[4];
#pragma unroll
      for (int i = 0; i < 8; ++i)
        af[i] = *reinterpret_cast<const double2*>(sA + (slab * 128 + wm * 64 + i * 8 + gq) * 8 + tq * 2);
#pragma unroll
      for (int j = 0; j < 4; ++j)
        bf[j] = *reinterpret_cast<const double2*>(sB + (slab * 128 + wn * 32 + j * 8 + gq) * 8 + tq * 2);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i].x, bf[j].x);
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], af[i].y, bf[j].y);
    }
  }
  cp_async_wait<0>();

  // --- epilogue
  long crow0, ccol0;
  if (in_diag) {
    int blk = m0 / g.nb;
    Cp = g.Cdiag + (long)blk * g.nb * g.nb; ldc = g.nb;
    crow0 = m0 - blk * g.nb; ccol0 = n0 - blk * g.nb;
  } else { crow0 = m0; ccol0 = n0; }
  const double alpha = g.alpha, beta = g.beta;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    long r = crow0 + wm * 64 + i * 8 + gq;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      long c = ccol0 + wn * 32 + j * 8 + tq * 2;
      double2* p = reinterpret_cast<double2*>(Cp + r * ldc + c);
      double2 v;
      if (beta != 0.0) {
        v = *p;
        v.x = beta * v.x + alpha * acc[i][j][0];
        v.y = beta * v.y + alpha * acc[i][j][1];
      } else {
        v.x = alpha * acc[i][j][0]; v.y = alpha * acc[i][j][1];
      }
      *p = v;
    }
  }
}

inline cudaError_t gemm_init() {
  return cudaFuncSetAttribute(dgemm_nt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GEMM_SMEM_BYTES);
}

inline void gemm_launch(const GemmArgs& g, cudaStream_t st) {
  long nt = gemm_num_tiles(g);
  if (nt <= 0 || g.kend <= g.kbeg) return;
  dgemm_nt_kernel<<<(unsigned)nt, GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(g);
}

}  // namespace kgp
