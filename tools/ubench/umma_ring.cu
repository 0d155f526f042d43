// Microbenchmark: the GEMM's stage ring in miniature - a producer warp that waits on `empty[s]` and arrives on
// `full[s]` (no loads: the operands are resident zeros), and one MMA-issuing thread that consumes the stages with
// four tcgen05.mma per stage and hands them back with tcgen05.commit.  Cycles per K block by consumer loop structure.
// The tensor pipe alone needs 172 (cta_group::2 M=256 N=64) / 192 (cta_group::1 M=128 N=64) cycles per K block.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I tilelang_mesh_b200/csrc -o tools/ubench/_build/umma_ring tools/ubench/umma_ring.cu
#include <cstdio>
#include <cuda.h>
#include "ptx.cuh"
using namespace tlb200;

constexpr int kStages = 8;
enum Variant { kWaitEach, kProbeTest, kProbeTry, kNoMma, kNoMmaNoProbe, kCommitEvery2, kProbeTwoAhead, kVariants };
static const char* kNames[kVariants] = {
    "wait (try_wait spin), 4 MMAs, commit", "probe next stage (test_wait) before the MMAs", "probe next stage (try_wait) before the MMAs",
    "probe (test_wait) + commit, NO MMAs", "wait + commit, NO MMAs", "probe (test_wait), commit only every 2nd stage (releases 2)",
    "test_wait of stage i+2 issued before MMAs of stage i"};

template <int kCta, int kVar>
__global__ void __launch_bounds__(128, 1) ring(uint32_t idesc, int iters, long long* cycles) {
  extern __shared__ uint8_t raw[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t full_b[kStages], empty_b[kStages], done_b;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const uint32_t rank = kCta == 2 ? cluster_ctarank() : 0u;
  for (int i = threadIdx.x; i < (48 << 10) / 4; i += blockDim.x) asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u));
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(smem_u32(&full_b[s]), 1); mbar_init(smem_u32(&empty_b[s]), 1); }
    mbar_init(smem_u32(&done_b), 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<kCta>(smem_u32(&tptr), 512);
  fence_proxy_async_smem();
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tptr;
  const uint32_t hi = (64u) | (1u << 14) | (2u << 29);
  const uint32_t a_w = ((base >> 4) & 0x3FFFu) | (1u << 16);
  const uint32_t b_w = (((base + (16u << 10)) >> 4) & 0x3FFFu) | (1u << 16);
  const uint32_t full0 = smem_u32(&full_b[0]), empty0 = smem_u32(&empty_b[0]), done_bar = smem_u32(&done_b);
  constexpr uint16_t kMask = kCta == 2 ? 0x3 : 0x1;  // the GEMM hands a stage back to both CTAs of a pair
  constexpr int kRelease = kVar == kCommitEvery2 ? 2 : 1;  // stages handed back per commit
  if (rank == 0 && warp == 2) {
    // producer: one lane, like the GEMM's elected TMA issuer, minus the loads
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      for (int i = 0; i < iters; ++i) {
        if constexpr (kRelease == 1) mbar_wait(empty0 + 8 * s, ph ^ 1u, 0);
        else if ((s & 1) == 0) mbar_wait(empty0 + 8 * s, ph ^ 1u, 0);  // one barrier per PAIR of stages
        mbar_arrive(full0 + 8 * s);
        if (++s == kStages) { s = 0; ph ^= 1u; }
      }
    }
  } else if (rank == 0 && warp == 0) {
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      bool ready = false, ready2 = false;
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        const uint32_t f = full0 + 8 * s, e = empty0 + 8 * s;
        if constexpr (kVar == kWaitEach || kVar == kNoMmaNoProbe) mbar_wait(f, ph, 2);
        else if (!ready) mbar_wait(f, ph, 2);
        tc_fence_after();
        int s1 = s + 1; uint32_t ph1 = ph;
        if (s1 == kStages) { s1 = 0; ph1 ^= 1u; }
        if constexpr (kVar == kProbeTest || kVar == kNoMma || kVar == kCommitEvery2) ready = i + 1 < iters && mbar_test_wait(full0 + 8 * s1, ph1);
        if constexpr (kVar == kProbeTry) ready = i + 1 < iters && mbar_try_wait(full0 + 8 * s1, ph1);
        if constexpr (kVar == kProbeTwoAhead) {
          int s2 = s1 + 1; uint32_t ph2 = ph1;
          if (s2 == kStages) { s2 = 0; ph2 ^= 1u; }
          ready = ready2;
          ready2 = i + 2 < iters && mbar_test_wait(full0 + 8 * s2, ph2);
        }
        if constexpr (kVar != kNoMma && kVar != kNoMmaNoProbe)
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
        if constexpr (kRelease == 1) umma_commit<kCta>(e, kMask);
        else if (s & 1) umma_commit<kCta>(e - 8, kMask);
        s = s1; ph = ph1;
      }
      umma_commit<kCta>(done_bar, 0x1);
      mbar_wait(done_bar, 0, 0);
      if (blockIdx.x == 0) cycles[0] = clock64() - t0;
    }
  }
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) tmem_dealloc<kCta>(tmem, 512);
}

static uint32_t idesc_of(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t(n >> 3) & 0x3fu) << 17) | ((uint32_t(m >> 4) & 0x1fu) << 24);
}

template <int kCta, int kVar>
static double run(int m, int n, int iters, long long* cyc) {
  auto kern = ring<kCta, kVar>;
  cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 52 << 10);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(148); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = 52 << 10;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = kCta; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  long long best = -1;
  for (int rep = 0; rep < 3; ++rep) {
    cudaError_t e = cudaLaunchKernelEx(&cfg, kern, idesc_of(m, n), iters, cyc);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); exit(1); }
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    if (best < 0 || c < best) best = c;
  }
  return double(best) / iters;
}

template <int kVar>
static void row(int iters, long long* cyc) {
  const double c1 = run<1, kVar>(128, 64, iters, cyc), c2 = run<2, kVar>(256, 64, iters, cyc), c3 = run<2, kVar>(256, 128, iters, cyc);
  printf("  %-58s %16.1f %16.1f %16.1f\n", kNames[kVar], c1, c2, c3);
  if constexpr (kVar + 1 < kVariants) row<kVar + 1>(iters, cyc);
}

int main() {
  long long* cyc; cudaMalloc(&cyc, 8);
  printf("cycles per K block through an 8-stage full/empty ring (producer arrives, consumer issues 4 bf16 MMAs and commits)\n  %-58s %16s %16s %16s\n",
         "consumer loop", "cg1 M=128 N=64", "cg2 M=256 N=64", "cg2 M=256 N=128");
  row<0>(4000, cyc);
  return 0;
}
