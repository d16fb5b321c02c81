// FP64 tensor-core GEMM, TMA-fed:  C = beta*C + alpha * A * B^T  (same contract and tile scheduler as dgemm_nt.cuh).
//
// What changes against dgemm_nt_kernel_t (cp.async ring + one __syncthreads per stage):
//   * operand stages arrive by TMA (cp.async.bulk.tensor.2d, one elected thread, two instructions per 32 KB stage)
//     into a 128B-swizzled [row][16 doubles] layout; the per-thread LDGSTS + address arithmetic is gone;
//   * full/empty mbarriers per ring slot instead of block barriers: no warp ever waits for another warp, only for data;
//   * the two warps that share an SM sub-partition (wm = 0 / wm = 1) run SKEWED by a few stages: while one of them is
//     in its epilogue (C read-modify-write, no DMMAs) the other one still has a full queue of DMMAs, so the FP64
//     tensor pipe does not drain at tile boundaries.  The ring (7 x 32 KB) is deep enough to hold the skew plus the
//     prefetch distance.
// sm_100a has no FP64 kind for tcgen05.mma (ptxas rejects .kind::f64), so the math instruction stays the warp-level
// DMMA.8x8x4 (mma.sync.m8n8k4.f64) with register accumulators; TMA/mbarrier are the Blackwell parts that apply.
#pragma once
#include "dgemm_nt_small.cuh"
#include <cuda.h>

namespace kgp {

constexpr int TG_BK = 16;                                   // 16 doubles = 128 B = the swizzle span
// BM = 128: one 256-thread CTA per SM on 128x128 tiles, 7-slot ring.  BM = 64: 64x128 half tiles, 128 threads, TWO
// independent CTAs per SM (4-slot ring each) -- the shape cuBLAS picks for FP64 on this part (cutlass d884gemm 64x128).
template <int BM> struct TgCfg {
  static constexpr int THREADS = 2 * BM, WARPS = BM / 16, HALVES = 128 / BM;
  static constexpr int STAGES = BM == 128 ? 7 : 4;
  static constexpr int A_BYTES = BM * TG_BK * 8, B_BYTES = 128 * TG_BK * 8, STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 256;
};

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
               : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {}
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(c0), "r"(c1), "r"(bar) : "memory");
}
__device__ __forceinline__ double lds_f64(uint32_t addr) {
  double v;
  asm volatile("ld.shared.f64 %0, [%1];" : "=d"(v) : "r"(addr));
  return v;
}

// skew_ns: how long the wm = 1 warps wait at kernel start (they stay that far behind for the whole launch)
template <int BM>
__global__ void __launch_bounds__(2 * BM, 128 / BM)
dgemm_nt_tma_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const GemmArgs g,
                    const long nitems, const int skew_ns, const int depth, const int dbg_nomem) {
  using Cfg = TgCfg<BM>;
  constexpr int TG_STAGES = Cfg::STAGES, TG_STAGE_BYTES = Cfg::STAGE_BYTES, TG_OPER_BYTES = Cfg::A_BYTES, HALVES = Cfg::HALVES;
  const long ntiles = nitems / HALVES;
  extern __shared__ __align__(1024) unsigned char tg_smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp >> 2, wn = warp & 3;        // 2 x 4 warps; warps w and w+4 share a sub-partition
  const int gq = lane >> 2, tq = lane & 3;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(tg_smem);
  const uint32_t bars = sbase + TG_STAGES * TG_STAGE_BYTES;          // full[s] at bars + 8 s, empty[s] at bars + 8 (S + s)
  if (tid == 0) {
    if (sbase & 1023u) __trap();                                       // SWIZZLE_128B needs 1024-byte aligned stages
    for (int s = 0; s < TG_STAGES; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 8 * (TG_STAGES + s), Cfg::WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  __syncthreads();

  auto fetch = [&](long pos, GemmTile& t) -> bool {
    const long item = gemm_seq_idx(pos);
    if (item >= nitems) return false;
    gemm_fetch_tile(g, item / HALVES, ntiles, t);
    if (HALVES > 1) { const int h = (int)(item % HALVES); t.m0 += h * BM; t.arow += h * BM; }
    return true;
  };

  // ---- producer state (thread 0 only): ring slot + use parity of the next load, tile cursor
  GemmTile lt; long lpos = 0; int lkt = 0; int lslot = 0; uint32_t lpar = 1;   // first use of a slot: the empty wait passes at parity 1
  bool lvalid = false;
  if (tid == 0) lvalid = fetch(lpos, lt);
  auto produce = [&]() {                          // issue the loads of the next stage (thread 0)
    if (!lvalid) return;
    mbar_wait(bars + 8 * (TG_STAGES + lslot), lpar);                   // every warp has released the previous use
    const uint32_t full = bars + 8 * lslot, dst = sbase + lslot * TG_STAGE_BYTES;
    if (dbg_nomem) mbar_arrive(full);              // lab only: barrier protocol without the loads (stale operands)
    else {
      mbar_expect_tx(full, TG_STAGE_BYTES);
      const int kc = lt.kc + lkt * TG_BK;
      tma_load_2d(dst, &tmA, kc, lt.arow, full);
      tma_load_2d(dst + TG_OPER_BYTES, &tmB, kc, lt.brow, full);
    }
    if (++lslot == TG_STAGES) { lslot = 0; lpar ^= 1u; }
    if (++lkt == lt.nk) { lkt = 0; ++lpos; lvalid = fetch(lpos, lt); }
  };
  if (tid == 0) for (int s = 0; s < depth; ++s) produce();

  if (wm == 1 && skew_ns > 0) {                   // (experiment) fall behind the wm = 0 warps once; the lag persists
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    do { __nanosleep(256); asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1)); } while (t1 - t0 < (unsigned long long)skew_ns);
  }

  // per-thread fragment addresses inside stage 0 (128B swizzle: 16-byte chunk index XOR (row & 7), row & 7 == gq)
  const uint32_t tsw = (uint32_t)((tq >> 1) ^ gq);
  uint32_t offA[4], offB[4];
#pragma unroll
  for (int st = 0; st < 4; ++st) {
    const uint32_t sw = (((uint32_t)(2 * st) ^ tsw) << 4) + (uint32_t)(tq & 1) * 8u;
    offA[st] = sbase + (uint32_t)(wm * 64 + gq) * 128u + sw;
    offB[st] = sbase + TG_OPER_BYTES + (uint32_t)(wn * 32 + gq) * 128u + sw;
  }
  double fa[2][8], fb[2][4];
  auto lds = [&](int buf, uint32_t soff, int step) {              // soff = slot * TG_STAGE_BYTES
#pragma unroll
    for (int i = 0; i < 8; ++i) fa[buf][i] = lds_f64(soff + offA[step] + i * 1024);
#pragma unroll
    for (int j = 0; j < 4; ++j) fb[buf][j] = lds_f64(soff + offB[step] + j * 1024);
  };

  GemmTile ct;
  int slot = 0; uint32_t par = 0;                 // consumer position in the ring
  bool have = fetch(0, ct);
  if (have) { mbar_wait(bars, 0); lds(0, 0, 0); }
  for (long cpos = 0; have; ++cpos) {
    GemmTile nt; const bool have_next = fetch(cpos + 1, nt);
    double* Cp; long ldc; long crow0, ccol0;
    if (ct.in_diag) {
      int blk = ct.m0 / g.nb;
      Cp = g.Cdiag + (long)blk * g.nb * g.nb; ldc = g.nb;
      crow0 = ct.m0 - blk * g.nb; ccol0 = ct.n0 - blk * g.nb;
    } else { Cp = g.C + ct.coff; ldc = g.ldc; crow0 = ct.m0; ccol0 = ct.n0; }
    const double alpha = g.alpha, beta = g.beta;
    if (beta != 0.0 && g.cdiag_only) {            // pull the C tile towards L2 while the main loop runs
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        int line = tid * 4 + i;
        const double* pp = Cp + (crow0 + line / 8) * ldc + ccol0 + (line % 8) * 16;
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pp));
      }
    }
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;

    for (int kt = 0; kt < ct.nk; ++kt) {
      const uint32_t soff = (uint32_t)slot * TG_STAGE_BYTES;
      int nslot = slot + 1; uint32_t npar = par;
      if (nslot == TG_STAGES) { nslot = 0; npar ^= 1u; }
      const bool more = (kt + 1 < ct.nk) || have_next;
      bool ready = false;
      if (tid == 0) produce();                    // keeps `depth` stages in flight ahead of warp 0
#pragma unroll
      for (int step = 0; step < 4; ++step) {
        const int cur = step & 1;
        if (step < 3) {
          lds(cur ^ 1, soff, step + 1);
          if (step == 2 && more) ready = mbar_try_wait(bars + 8 * nslot, npar);   // probe early: its latency hides behind this step's DMMAs
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
        } else {
          // last k4-step of the stage: half of its DMMAs go out before the first fragments of the next stage are fetched
#pragma unroll
          for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
          if (more) {
            if (!ready) mbar_wait(bars + 8 * nslot, npar);
            lds(cur ^ 1, (uint32_t)nslot * TG_STAGE_BYTES, 0);
          }
#pragma unroll
          for (int i = 4; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
          __syncwarp();
          if (lane == 0) mbar_arrive(bars + 8 * (TG_STAGES + slot));   // this warp is done with the slot
        }
      }
      slot = nslot; par = npar;
    }

    // ---- epilogue (registers -> global), same as dgemm_nt_kernel_t
    double* cbase = Cp + (crow0 + wm * 64 + gq) * ldc + ccol0 + wn * 32 + tq * 2;
    if (beta != 0.0) {
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
    ct = nt; have = have_next;
  }
}

// ---- host side -------------------------------------------------------------------------------------------------
typedef CUresult (*tg_encode_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                 const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                 CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static tg_encode_fn g_tg_encode = nullptr;
static int g_tg_skew_ns = 0, g_tg_depth = 0, g_gemm_impl = 1, g_tg_nomem = 0, g_tg_bm = 128;   // env KGP_TG_SKEW_NS / KGP_TG_DEPTH / KGP_GEMM_IMPL (0: cp.async kernel) / KGP_TG_BM

inline cudaError_t gemm_tma_init() {
  if (const char* e = getenv("KGP_GEMM_IMPL")) g_gemm_impl = atoi(e);
  if (const char* e = getenv("KGP_TG_SKEW_NS")) g_tg_skew_ns = atoi(e);
  if (const char* e = getenv("KGP_TG_DEPTH")) g_tg_depth = atoi(e);
  if (const char* e = getenv("KGP_TG_BM")) g_tg_bm = atoi(e) == 64 ? 64 : 128;
  void* fn = nullptr; cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess) return e;
  if (!fn || qres != cudaDriverEntryPointSuccess) return cudaErrorNotSupported;
  g_tg_encode = (tg_encode_fn)fn;
  e = cudaFuncSetAttribute(dgemm_nt_tma_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, TgCfg<64>::SMEM_BYTES);
  if (e != cudaSuccess) return e;
  return cudaFuncSetAttribute(dgemm_nt_tma_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, TgCfg<128>::SMEM_BYTES);
}

// row-major FP64 matrix `base` (rows x kcols used, leading dimension ld) as a 2-D tensor map with 16 x box_rows boxes
inline bool tg_make_map(CUtensorMap* m, const double* base, long ld, long rows, long kcols, int box_rows) {
  cuuint64_t dims[2] = {(cuuint64_t)kcols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
  cuuint32_t box[2] = {TG_BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return g_tg_encode(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, (void*)base, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// same contract as gemm_launch (g already carries band / batch tables); returns false when TMA cannot serve the operands
inline bool gemm_launch_tma(const GemmArgs& g, long nt, bool persistent, cudaStream_t st) {
  if (!g_tg_encode) return false;
  if (((uintptr_t)g.A & 15) || ((uintptr_t)g.B & 15) || (g.lda & 1) || (g.ldb & 1)) return false;
  long rowsA = (long)g.ntm * 128, rowsB = (long)g.ntn * 128;
  if (g.nbatch > 0) { long mx = 0; for (int p = 0; p < g.nbatch; ++p) mx = g.bat_roff[p] > mx ? g.bat_roff[p] : mx; rowsB += mx; }
  CUtensorMap ma, mb;
  if (!tg_make_map(&ma, g.A, g.lda, rowsA, g.kend, g_tg_bm) || !tg_make_map(&mb, g.B, g.ldb, rowsB, g.kend, 128)) return false;
  if (g_tg_bm == 64) {
    const long items = 2 * nt, res = 2L * g_gemm_num_sms, grid = (persistent && items > res) ? res : items;
    const int depth = g_tg_depth > 0 && g_tg_depth < TgCfg<64>::STAGES ? g_tg_depth : TgCfg<64>::STAGES - 1;
    dgemm_nt_tma_kernel<64><<<(unsigned)grid, TgCfg<64>::THREADS, TgCfg<64>::SMEM_BYTES, st>>>(ma, mb, g, items, 0, depth, g_tg_nomem);
  } else {
    const long grid = (persistent && nt > g_gemm_num_sms) ? g_gemm_num_sms : nt;
    const int depth = g_tg_depth > 0 && g_tg_depth < TgCfg<128>::STAGES ? g_tg_depth : 4;
    dgemm_nt_tma_kernel<128><<<(unsigned)grid, TgCfg<128>::THREADS, TgCfg<128>::SMEM_BYTES, st>>>(ma, mb, g, nt, nt > grid ? g_tg_skew_ns : 0, depth, g_tg_nomem);
  }
  return true;
}

// tile_ctas: one CTA per tile instead of persistent CTAs -- a higher-priority stream's kernels (panel factorisation,
// NCCL) then get SMs at tile granularity (~80 us) instead of waiting for the whole launch
// workspace of the small-tile path (dgemm_nt_small.cuh); one per stream that may carry small launches
struct GemmSmallWs { double* ws; size_t ws_bytes; unsigned* cnt; long cnt_n; };
inline void gemm_launch(const GemmArgs& g_in, cudaStream_t st, bool tile_ctas = false, const GemmSmallWs* sw = nullptr) {
  GemmArgs g = g_in;
  long nt = gemm_num_tiles(g);
  if (g.shape == GEMM_LOWER) {
    long base = 0; int nb_ = 0;
    for (int r0 = 0; r0 < g.ntm && nb_ < GEMM_MAX_BANDS; r0 += 8) {
      g.band_base[nb_++] = (int)base;
      for (int r = r0; r < r0 + 8 && r < g.ntm; ++r) base += (r + 1 < g.ntn) ? r + 1 : g.ntn;
    }
    g.nbands = nb_;
  }
  if (g.nbatch > 0) {
    long base = 0;
    for (int p = 0; p < g.nbatch; ++p) { g.bat_base[p] = (int)base; base += (long)(g.ntm - g.bat_roff[p] / 128) * g.ntn; }
    g.bat_base[g.nbatch] = (int)base;
  }
  if (g.own_mod > 1) nt = (nt > g.own_rem) ? (nt - g.own_rem + g.own_mod - 1) / g.own_mod : 0;
  if (g.tile_cnt > 0) nt = g.tile_cnt;                 // a stage of a staged launch (tile_off already points at its first tile)
  if (nt <= 0 || g.kend <= g.kbeg) return;
  g.cdiag_only = g_gemm_prefetch_c;
  if (sw) {
    int nchmax = 0; size_t need = 0;
    if (gemm_small_eligible(g, nt, nchmax, need) && need <= sw->ws_bytes && nt * 16 <= sw->cnt_n) {
      gemm_launch_small(g, nt, nchmax, sw->ws, sw->cnt, st);
      return;
    }
  }
  if (g_gemm_bn == 64) {
    const long items = 2 * nt, res = 2L * g_gemm_num_sms;
    const long grid = (g_gemm_persist && !tile_ctas && items > res) ? res : items;
    dgemm_nt_kernel_t<64><<<(unsigned)grid, 128, GEMM_STAGES * (GEMM_BM + 64) * GEMM_BK * 8, st>>>(g, items);
    return;
  }
  if (g_gemm_impl == 1 && gemm_launch_tma(g, nt, g_gemm_persist && !tile_ctas, st)) return;
  long grid = (g_gemm_persist && !tile_ctas && nt > g_gemm_num_sms) ? g_gemm_num_sms : nt;
  dgemm_nt_kernel_t<128><<<(unsigned)grid, GEMM_THREADS, GEMM_SMEM_BYTES, st>>>(g, nt);
}

}  // namespace kgp
