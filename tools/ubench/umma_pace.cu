// Microbenchmark: steady-state cycles per tcgen05.mma (SS form: both operands in shared memory, 128-byte swizzle,
// K-major) by instruction shape, for one CTA and for a CTA pair.  One elected thread issues `iters` K blocks of four
// MMAs back to back into one accumulator and waits for a single commit; the operands are zeros already in shared
// memory, so what is measured is the tensor pipe's own pace including its operand reads.
//   umma_pace        the table by shape (1 SM busy, 148 SMs busy, with a commit after every K block)
//   umma_pace tma    the same pace while another warp streams bulk loads into the same SM's shared memory
// Issued under elect.sync like the GEMM: ptxas then knows one lane is active and emits bare UTCHMMAs (under a
// threadIdx test it wraps every one in an ELECT loop, which costs ~45 cycles per MMA by itself).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I tilelang_mesh_b200/csrc -o tools/ubench/_build/umma_pace tools/ubench/umma_pace.cu
#include <cstdio>
#include <cuda.h>
#include "ptx.cuh"
using namespace tlb200;

template <int kCta, int kKind>
__global__ void __launch_bounds__(128, 1) pace(uint32_t idesc, int iters, long long* cycles, int commit_each, const uint8_t* src,
                                               int load_bytes) {
  extern __shared__ uint8_t raw[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t bar;
  __shared__ __align__(8) uint64_t bar2;     // receives the per-K-block commits; nobody waits on it
  __shared__ __align__(8) uint64_t lbar[4];  // one per buffer of the concurrent loader
  __shared__ volatile uint32_t stop;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int warp = threadIdx.x >> 5;
  const uint32_t rank = kCta == 2 ? cluster_ctarank() : 0u;
  for (int i = threadIdx.x; i < (48 << 10) / 4; i += blockDim.x) asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u));
  if (threadIdx.x == 0) {
    mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&bar2), 1);
    for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&lbar[i]), 1);
    stop = 0;
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kCta>(smem_u32(&tptr), 512);
  fence_proxy_async_smem();
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tptr;
  if (rank == 0 && warp == 0 && elect_one()) {
    const uint32_t hi = (64u) | (1u << 14) | (2u << 29);  // SBO = 1024 B, version 1, 128-byte swizzle
    const uint32_t a_w = ((base >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t b_w = (((base + (16u << 10)) >> 4) & 0x3FFFu) | (1u << 16);
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      umma_ss_x4<kCta, kKind>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
      if (commit_each) umma_commit<kCta>(smem_u32(&bar2), 0x1);  // what a GEMM main loop does to hand the stage back
    }
    umma_commit<kCta>(smem_u32(&bar), 0x1);
    mbar_wait(smem_u32(&bar), 0, 0);
    const long long t1 = clock64();
    for (uint32_t r = 0; r < kCta; ++r)  // both CTAs of a pair run a loader
      asm volatile("st.shared::cluster.u32 [%0], %1;" ::"r"(kCta == 2 ? mapa(smem_u32((const void*)&stop), r) : smem_u32((const void*)&stop)),
                   "r"(1u)
                   : "memory");
    if (blockIdx.x == 0) cycles[0] = t1 - t0;
  }
  if (warp == 2 && load_bytes > 0 && elect_one()) {
    // concurrent bulk loads (L2-resident source) into four shared-memory buffers behind the operands, as fast as a
    // four-deep ring allows, until the MMA thread is done: do TMA writes and MMA operand reads share bandwidth?
    const uint8_t* mine = src + size_t(blockIdx.x) * 3 * load_bytes;
    long long loaded = 0;
    int j = 0;
    for (; !stop && j < 400000; ++j) {  // the cap bounds the kernel should the stop flag be lost
      const uint32_t b = smem_u32(&lbar[j & 3]);
      if (j >= 4) mbar_wait(b, ((j >> 2) - 1) & 1, 3);
      mbar_arrive_expect_tx(b, load_bytes);
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                       base + (48u << 10) + (j & 3) * load_bytes),
                   "l"(mine + (j % 3) * load_bytes), "r"(load_bytes), "r"(b)
                   : "memory");
      loaded += load_bytes;
    }
    for (int q = j < 4 ? 0 : j - 4; q < j; ++q) mbar_wait(smem_u32(&lbar[q & 3]), (q >> 2) & 1, 3);  // drain
    if (blockIdx.x == 0) cycles[1] = loaded;
  }
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) tmem_dealloc<kCta>(tmem, 512);
}

static uint32_t idesc_of(int m, int n, bool fp8) {
  uint32_t d = 0;
  d |= 1u << 4;               // fp32 accumulator
  d |= (fp8 ? 0u : 1u) << 7;  // A: e4m3 / bf16
  d |= (fp8 ? 0u : 1u) << 10; // B
  d |= (uint32_t(n >> 3) & 0x3fu) << 17;
  d |= (uint32_t(m >> 4) & 0x1fu) << 24;
  return d;
}

static const uint8_t* g_src = nullptr;
static long long g_loaded = 0;

template <int kCta, int kKind>
static long long run(int m, int n, bool fp8, int iters, long long* cyc, int ctas, int commit_each = 0, int load_bytes = 0) {
  auto kern = pace<kCta, kKind>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 180 << 10);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(ctas);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = 180 << 10;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kCta; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  long long best = -1;
  for (int rep = 0; rep < 3; ++rep) {
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, idesc_of(m, n, fp8), iters, cyc, commit_each, g_src, load_bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    long long c[2]; cudaMemcpy(c, cyc, 16, cudaMemcpyDeviceToHost);
    if (best < 0 || c[0] < best) { best = c[0]; g_loaded = c[1]; }
  }
  return best;
}

int main(int argc, char** argv) {
  long long* cyc; cudaMalloc(&cyc, 16);
  const int iters = 2000;
  {
    uint8_t* srcbuf; cudaMalloc(&srcbuf, size_t(148) * 3 * (32 << 10)); cudaMemset(srcbuf, 0, size_t(148) * 3 * (32 << 10));
    g_src = srcbuf;
  }
  if (argc > 1 && argv[1][0] == 't') {
    printf("MMA pace with concurrent bulk loads into the same SM's shared memory (bf16, 148 SMs busy)\n");
    printf("  %-26s %12s | %16s %16s\n", "shape", "load bytes/blk", "cycles/K block", "loaded B/clk/SM");
    for (int lb : {0, 8 << 10, 20 << 10, 32 << 10}) {
      const long long c2 = run<2, kKindF16>(256, 64, false, iters, cyc, 148, 1, lb);
      printf("  %-26s %12d | %16.1f %16.1f\n", "cta_group::2 M=256 N=64", lb, double(c2) / iters, double(g_loaded) / c2);
      const long long c3 = run<2, kKindF16>(256, 256, false, iters, cyc, 148, 1, lb);
      printf("  %-26s %12d | %16.1f %16.1f\n", "cta_group::2 M=256 N=256", lb, double(c3) / iters, double(g_loaded) / c3);
      const long long c1 = run<1, kKindF16>(128, 64, false, iters, cyc, 148, 1, lb);
      printf("  %-26s %12d | %16.1f %16.1f\n", "cta_group::1 M=128 N=64", lb, double(c1) / iters, double(g_loaded) / c1);
    }
    return 0;
  }
  for (int fp8 = 0; fp8 < 2; ++fp8) {
    printf("%s, K per MMA = 32 bytes; cycles per MMA at steady state (4 MMAs per 128-byte K block)\n", fp8 ? "e4m3" : "bf16");
    printf("  %-28s %8s %8s %14s %22s\n", "shape", "1 SM", "148 SMs", "MAC/clk/SM", "commit per 4 MMAs");
    for (int cg = 1; cg <= 2; ++cg) {
      for (int m_cta : {64, 128}) {
        for (int n : {32, 64, 128, 256}) {
          const int m = m_cta * cg;
          long long c1, c2, c3;
          if (cg == 1) {
            c1 = fp8 ? run<1, kKindF8>(m, n, true, iters, cyc, 1) : run<1, kKindF16>(m, n, false, iters, cyc, 1);
            c2 = fp8 ? run<1, kKindF8>(m, n, true, iters, cyc, 148) : run<1, kKindF16>(m, n, false, iters, cyc, 148);
            c3 = fp8 ? run<1, kKindF8>(m, n, true, iters, cyc, 148, 1) : run<1, kKindF16>(m, n, false, iters, cyc, 148, 1);
          } else {
            c1 = fp8 ? run<2, kKindF8>(m, n, true, iters, cyc, 2) : run<2, kKindF16>(m, n, false, iters, cyc, 2);
            c2 = fp8 ? run<2, kKindF8>(m, n, true, iters, cyc, 148) : run<2, kKindF16>(m, n, false, iters, cyc, 148);
            c3 = fp8 ? run<2, kKindF8>(m, n, true, iters, cyc, 148, 1) : run<2, kKindF16>(m, n, false, iters, cyc, 148, 1);
          }
          const double per = double(c1) / (4.0 * iters), per2 = double(c2) / (4.0 * iters), per3 = double(c3) / (4.0 * iters);
          const double macs = double(m_cta) * n * (fp8 ? 32 : 16) / per2;
          char shape[64];
          snprintf(shape, sizeof shape, "cta_group::%d M=%d N=%d", cg, m, n);
          printf("  %-28s %8.1f %8.1f %14.0f %22.1f\n", shape, per, per2, macs, per3);
        }
      }
    }
  }
  return 0;
}
