// Development bench for the FP64 GEMM kernels (not part of the product library): DMMA issue-rate variants that
// bound what a 256-thread CTA can reach, bitwise comparison of dgemm_nt_tma_kernel with dgemm_nt_kernel_t over every
// launch mode the exact path uses, and timings at the n = 40k call-site shapes.
//   nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a lab/gemm_lab.cu -o lab/gemm_lab
#include "../kgp_b200/csrc/dgemm_nt_tma.cuh"
#include <stdio.h>
#include <string.h>
#include <vector>
using namespace kgp;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

// ---- issue-rate variants -------------------------------------------------------------------------------------
// V: 0 pure DMMA (8x4 accumulators, operands in registers) | 1 + fragment LDS.64 from shared memory (double-buffered)
//    2 + __syncthreads every 4 k-steps | 3 = 1 with 16 accumulators (4x4, for 512-thread CTAs)
template <int V, int NI, int NJ, int TH>
__global__ void __launch_bounds__(TH) rate_kernel(double* out, int iters) {
  extern __shared__ double sm[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, gq = lane >> 2, tq = lane & 3;
  for (int i = tid; i < 4 * 256 * 4; i += blockDim.x) sm[i] = 1.0 + 1e-9 * i;
  __syncthreads();
  double acc[NI][NJ][2];
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) acc[i][j][0] = acc[i][j][1] = 0.0;
  double fa[2][NI], fb[2][NJ];
  const int wm = warp & 1, wn = (warp >> 1) & 3;
  auto lds = [&](int buf, int step) {
#pragma unroll
    for (int i = 0; i < NI; ++i) fa[buf][i] = sm[(step * 256 + wm * 64 + i * 8 + gq) * 4 + tq];
#pragma unroll
    for (int j = 0; j < NJ; ++j) fb[buf][j] = sm[(step * 256 + 128 + wn * 32 + j * 8 + gq) * 4 + tq];
  };
  lds(0, 0);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int step = 0; step < 4; ++step) {
      const int cur = (V == 0) ? 0 : (step & 1);
      if (V >= 1) lds(cur ^ 1, (step + 1) & 3);
      if (V == 2 && step == 3) __syncthreads();
#pragma unroll
      for (int i = 0; i < NI; ++i)
#pragma unroll
        for (int j = 0; j < NJ; ++j) dmma884(acc[i][j][0], acc[i][j][1], fa[cur][i], fb[cur][j]);
    }
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int j = 0; j < NJ; ++j) s += acc[i][j][0] + acc[i][j][1];
  if (s == 123.456) out[0] = s;
}
template <int V, int NI, int NJ, int TH>
static void run_rate(const char* name, int threads, int bps, double* out) {
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int iters = 4000; cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  CK(cudaFuncSetAttribute(rate_kernel<V, NI, NJ, TH>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  rate_kernel<V, NI, NJ, TH><<<sms * bps, threads, 40 * 1024>>>(out, 16);
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    cudaEventRecord(a); rate_kernel<V, NI, NJ, TH><<<sms * bps, threads, 40 * 1024>>>(out, iters); cudaEventRecord(b);
    CK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  const double fl = (double)sms * bps * (threads / 32) * (double)iters * 4 * NI * NJ * 512.0;
  printf("rate %-44s threads %3d x %d CTA/SM: %6.2f TFLOP/s\n", name, threads, bps, fl / (best * 1e-3) / 1e12);
}

// ---- correctness: new kernel vs old kernel, bitwise ----------------------------------------------------------
__global__ void fill(double* p, long n, unsigned seed) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { unsigned h = (unsigned)(i * 2654435761u) ^ seed; h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15; p[i] = ((double)(h & 0xffff) / 65536.0 - 0.5); }
}
static long compare(const double* a, const double* b, long n) {
  std::vector<double> ha(n), hb(n);
  CK(cudaMemcpy(ha.data(), a, n * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hb.data(), b, n * 8, cudaMemcpyDeviceToHost));
  long bad = 0; for (long i = 0; i < n; ++i) if (memcmp(&ha[i], &hb[i], 8)) ++bad;
  return bad;
}
static void run_impl(int impl, const GemmArgs& g, bool tile_ctas = false) { g_gemm_impl = impl; gemm_launch(g, 0, tile_ctas); CK(cudaGetLastError()); }

static GemmArgs mk(const double* A, long lda, const double* B, long ldb, double* C, long ldc, int ntm, int ntn, int kbeg, int kend,
                   int shape, int ktri, double alpha, double beta) {
  GemmArgs g; memset(&g, 0, sizeof g);
  g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.C = C; g.ldc = ldc; g.ntm = ntm; g.ntn = ntn; g.kbeg = kbeg; g.kend = kend;
  g.shape = shape; g.ktri = ktri; g.alpha = alpha; g.beta = beta; g.nb = 512;
  return g;
}

int main(int argc, char** argv) {
  const bool quick = argc > 1 && !strcmp(argv[1], "quick");
  const bool prof = argc > 1 && !strcmp(argv[1], "prof");
  CK(cudaSetDevice(0)); CK(gemm_init()); CK(gemm_tma_init());
  double* out; CK(cudaMalloc(&out, 64));
  if (prof) {                                   // two launches of the TMA kernel at 8192^3 (for ncu)
    const long kk = 8192; double *Ap, *Cp2; CK(cudaMalloc(&Ap, 2 * kk * kk * 8)); CK(cudaMalloc(&Cp2, kk * kk * 8));
    fill<<<(unsigned)((2 * kk * kk + 255) / 256), 256>>>(Ap, 2 * kk * kk, 5);
    if (argc > 2) g_tg_bm = atoi(argv[2]);
    for (int r = 0; r < 2; ++r) run_impl(1, mk(Ap, kk, Ap + kk * kk, kk, Cp2, kk, 64, 64, 0, 8192, GEMM_FULL, KTRI_NONE, 1.0, 0.0));
    CK(cudaDeviceSynchronize()); printf("prof done\n"); return 0;
  }
  run_rate<0, 8, 4, 256>("pure DMMA, 32 acc/thread", 256, 1, out);
  run_rate<1, 8, 4, 256>("DMMA + fragment LDS (GEMM ratio 12:32)", 256, 1, out);
  run_rate<2, 8, 4, 256>("DMMA + LDS + __syncthreads per stage", 256, 1, out);
  run_rate<0, 4, 4, 512>("pure DMMA, 16 acc/thread", 512, 1, out);
  run_rate<1, 4, 4, 512>("DMMA + LDS (8:16), 16 acc/thread", 512, 1, out);
  run_rate<2, 4, 4, 512>("DMMA + LDS + sync, 16 acc/thread", 512, 1, out);
  run_rate<0, 4, 4, 512>("pure DMMA, 16 acc/thread", 512, 2, out);
  run_rate<0, 8, 4, 256>("pure DMMA, 32 acc/thread", 128, 1, out);

  // correctness over the launch modes of exact.cu
  const long n = 2048, ld = 2048; const long N2 = n * ld;
  double *A, *B, *C0, *C1, *C2, *D1, *D2;
  CK(cudaMalloc(&A, N2 * 8)); CK(cudaMalloc(&B, N2 * 8)); CK(cudaMalloc(&C0, N2 * 8)); CK(cudaMalloc(&C1, N2 * 8)); CK(cudaMalloc(&C2, N2 * 8));
  CK(cudaMalloc(&D1, n * 512 * 8)); CK(cudaMalloc(&D2, n * 512 * 8));
  fill<<<(unsigned)((N2 + 255) / 256), 256>>>(A, N2, 1); fill<<<(unsigned)((N2 + 255) / 256), 256>>>(B, N2, 2); fill<<<(unsigned)((N2 + 255) / 256), 256>>>(C0, N2, 3);
  auto reset = [&]() { CK(cudaMemcpy(C1, C0, N2 * 8, cudaMemcpyDeviceToDevice)); CK(cudaMemcpy(C2, C0, N2 * 8, cudaMemcpyDeviceToDevice));
                       CK(cudaMemset(D1, 0, n * 512 * 8)); CK(cudaMemset(D2, 0, n * 512 * 8)); };
  struct Case { const char* name; GemmArgs g; bool tile; };
  std::vector<Case> cases;
  const int nt = (int)(n / 128);
  cases.push_back({"FULL beta=0", mk(A, ld, B, ld, nullptr, ld, nt, nt, 0, (int)n, GEMM_FULL, KTRI_NONE, 1.0, 0.0), false});
  cases.push_back({"FULL beta=1 alpha=-1 k=512 offset", mk(A + 128, ld, B + 128, ld, nullptr, ld, nt - 1, nt - 3, 0, 512, GEMM_FULL, KTRI_NONE, -1.0, 1.0), false});
  cases.push_back({"LOWER syrk beta=1 k=512", mk(A, ld, A, ld, nullptr, ld, nt, nt, 0, 512, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), false});
  cases.push_back({"LOWER syrk k=128 tile_ctas", mk(A, ld, A, ld, nullptr, ld, nt, nt, 0, 128, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), true});
  cases.push_back({"FULL ktri (z-sweep)", mk(A, ld, B, ld, nullptr, ld, nt - 4, 4, 0, (int)(n - 512), GEMM_FULL, KTRI_FROM_M0, -1.0, 0.0), false});
  { GemmArgs g = mk(A, ld, A, ld, nullptr, ld, nt, nt, 0, (int)n, GEMM_LOWER, KTRI_FROM_M0, 1.0, 0.0); cases.push_back({"LOWER ktri + Cdiag (zzt)", g, false}); }
  { GemmArgs g = mk(A, ld, A, ld, nullptr, ld, nt, nt, 0, (int)n, GEMM_LOWER, KTRI_FROM_M0, 1.0, 0.0); g.own_mod = 3; g.own_rem = 1; cases.push_back({"LOWER ktri own_mod 3", g, false}); }
  { GemmArgs g = mk(A, 512, A, 512, nullptr, ld, nt, 4, 0, 512, GEMM_FULL, KTRI_NONE, -1.0, 1.0); g.nbatch = 3; g.bat_roff[0] = 0; g.bat_roff[1] = 512; g.bat_roff[2] = 1024;
    cases.push_back({"batched FULL (dist trailing update)", g, false}); }
  for (int bm : {128, 64})
  for (auto& cs : cases) {
    g_tg_bm = bm;
    reset();
    GemmArgs g1 = cs.g, g2 = cs.g;
    g1.C = C1 + (cs.g.A == A + 128 ? 128 * ld : 0); g2.C = C2 + (cs.g.A == A + 128 ? 128 * ld : 0);
    if (strstr(cs.name, "Cdiag")) { g1.Cdiag = D1; g2.Cdiag = D2; }
    run_impl(0, g1, cs.tile); run_impl(1, g2, cs.tile);
    CK(cudaDeviceSynchronize());
    long bad = compare(C1, C2, N2) + compare(D1, D2, n * 512);
    long same_as_in = compare(C2, C0, N2);
    printf("check BM=%3d %-40s mismatching elements: %ld   (elements changed vs input: %ld)\n", bm, cs.name, bad, same_as_in);
  }
  if (quick) return 0;

  // timings at call-site shapes
  cudaFree(A); cudaFree(B); cudaFree(C0); cudaFree(C1); cudaFree(C2);
  const long np = 40064, kmax = 8192;
  double *Ab, *Cb; CK(cudaMalloc(&Ab, np * kmax * 8)); CK(cudaMalloc(&Cb, np * np * 8));
  fill<<<(unsigned)((np * kmax + 255) / 256), 256>>>(Ab, np * kmax, 5); CK(cudaMemset(Cb, 0, np * np * 8));
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  auto timeit = [&](const char* name, GemmArgs g, double flops, int impl) {
    run_impl(impl, g);
    cudaEventRecord(e0); for (int r = 0; r < 3; ++r) run_impl(impl, g); cudaEventRecord(e1);
    CK(cudaDeviceSynchronize()); float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 3;
    printf("time %-34s impl %d bm %3d depth %d: %8.3f ms  %6.2f TFLOP/s\n", name, impl, g_tg_bm, g_tg_depth, ms, flops / (ms * 1e-3) / 1e12);
  };
  const int ntb = (int)(np / 128);
  auto tri = [&](int t) { return (double)t * (t + 1) / 2; };
  struct Cfg { int bm, skew, depth; };
  Cfg cfgs[] = {{128, 0, 4}, {128, 0, 3}, {128, 0, 5}, {64, 0, 2}, {64, 0, 3}, {64, 0, 1}};
  for (int impl = 0; impl < 2; ++impl)
    for (auto cf : cfgs) {
      if (impl == 0 && (cf.bm != 128 || cf.depth != 4)) continue;
      g_tg_skew_ns = cf.skew; g_tg_depth = cf.depth; g_tg_bm = cf.bm;
      timeit("SYRK n=40064 k=512 beta=1", mk(Ab, kmax, Ab, kmax, Cb, np, ntb, ntb, 0, 512, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), 2.0 * 128 * 128 * 512 * tri(ntb), impl);
      timeit("GEMM 8192^3 beta=0", mk(Ab, kmax, Ab + 8192 * kmax, kmax, Cb, np, 64, 64, 0, 8192, GEMM_FULL, KTRI_NONE, 1.0, 0.0), 2.0 * 8192.0 * 8192 * 8192, impl);
      timeit("SYRK n=20096 k=512 beta=1", mk(Ab, kmax, Ab, kmax, Cb, np, 157, 157, 0, 512, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), 2.0 * 128 * 128 * 512 * tri(157), impl);
      timeit("panel update n=40064 k=128", mk(Ab, kmax, Ab, kmax, Cb, np, ntb, 3, 0, 128, GEMM_FULL, KTRI_NONE, -1.0, 1.0), 2.0 * 128 * 128 * 128 * ntb * 3.0, impl);
    }
  g_tg_skew_ns = 0; g_tg_depth = 4; g_tg_bm = 128; g_tg_nomem = 1;
  timeit("GEMM 8192^3 NO LOADS (stale smem)", mk(Ab, kmax, Ab + 8192 * kmax, kmax, Cb, np, 64, 64, 0, 8192, GEMM_FULL, KTRI_NONE, 1.0, 0.0), 2.0 * 8192.0 * 8192 * 8192, 1);
  timeit("SYRK n=40064 k=512 NO LOADS", mk(Ab, kmax, Ab, kmax, Cb, np, ntb, ntb, 0, 512, GEMM_LOWER, KTRI_NONE, -1.0, 1.0), 2.0 * 128 * 128 * 512 * tri(ntb), 1);
  g_tg_bm = 64; g_tg_depth = 2;
  timeit("GEMM 8192^3 NO LOADS (stale smem)", mk(Ab, kmax, Ab + 8192 * kmax, kmax, Cb, np, 64, 64, 0, 8192, GEMM_FULL, KTRI_NONE, 1.0, 0.0), 2.0 * 8192.0 * 8192 * 8192, 1);
  g_tg_nomem = 0;
  return 0;
}
