// Microbenchmark: what an MMA-issue loop costs per 128-byte K block (four tcgen05.mma of a narrow shape) when the
// operands are always resident, by loop structure.  The tensor pipe alone needs 172 cycles (cta_group::2, M=256, N=64)
// or 192 (cta_group::1, M=128, N=64) per K block (tools/ubench/umma_pace.cu); everything above that is the issuing
// thread's own latency showing through, because the pipe holds next to no queue of issued MMAs.
// Every variant is its own kernel instantiation (no run-time dispatch inside the loops).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I tilelang_mesh_b200/csrc -o tools/ubench/_build/umma_loop tools/ubench/umma_loop.cu
#include <cstdio>
#include <cuda.h>
#include "ptx.cuh"
using namespace tlb200;

enum Mode { kBare, kCommit, kWait, kProbeTry, kProbeTest, kTwoBlocks, kFourBlocks, kWarpLoop, kWarpLoopVote, kHelperFlag, kModes };
static const char* kNames[kModes] = {
    "4 MMAs (one commit at the end)", "4 MMAs + commit", "try_wait spin, 4 MMAs, commit", "probe next barrier (try_wait) before MMAs",
    "probe next barrier (test_wait) before MMAs", "2 K blocks per pass (2 waits, 8 MMAs, 2 commits)", "4 K blocks per pass",
    "whole warp waits, elected lane issues", "whole warp waits on a vote, elected issues", "helper warp waits, issuer polls a shared word"};

template <int kCta, int kMode>
__global__ void __launch_bounds__(128, 1) loop_study(uint32_t idesc, int iters, long long* cycles) {
  extern __shared__ uint8_t raw[];
  __shared__ uint32_t tptr;
  __shared__ __align__(8) uint64_t done_b, open_b;
  __shared__ volatile uint32_t ready_upto;
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
  const uint32_t rank = kCta == 2 ? cluster_ctarank() : 0u;
  for (int i = threadIdx.x; i < (48 << 10) / 4; i += blockDim.x) asm volatile("st.shared.u32 [%0], %1;" ::"r"(base + 4 * i), "r"(0u));
  if (threadIdx.x == 0) { mbar_init(smem_u32(&done_b), 1); mbar_init(smem_u32(&open_b), 1); ready_upto = 0; fence_mbar_init(); }
  if (warp == 1) tmem_alloc<kCta>(smem_u32(&tptr), 512);
  fence_proxy_async_smem();
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tptr;
  const uint32_t hi = (64u) | (1u << 14) | (2u << 29);  // SBO = 1024 B, version 1, 128-byte swizzle
  const uint32_t a_w = ((base >> 4) & 0x3FFFu) | (1u << 16);
  const uint32_t b_w = (((base + (16u << 10)) >> 4) & 0x3FFFu) | (1u << 16);
  const uint32_t open_bar = smem_u32(&open_b), done_bar = smem_u32(&done_b);  // open_bar: parity 1 always passes
  if (rank == 0 && warp == 0) {
    if constexpr (kMode == kWarpLoop || kMode == kWarpLoopVote) {
      const bool elected = elect_one();
      const long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
        if constexpr (kMode == kWarpLoop) mbar_wait(open_bar, 1, 0); else mbar_wait_uniform(open_bar, 1, 0);
        tc_fence_after();
        if (elected) {
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
          umma_commit<kCta>(done_bar, 0x1);
        }
        __syncwarp();
      }
      if (elected) {
        mbar_wait(done_bar, (iters - 1) & 1, 0);
        if (blockIdx.x == 0) cycles[0] = clock64() - t0;
      }
    } else if (elect_one()) {
      const long long t0 = clock64();
      if constexpr (kMode == kBare || kMode == kCommit) {
        for (int i = 0; i < iters; ++i) {
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
          if constexpr (kMode == kCommit) umma_commit<kCta>(done_bar, 0x1);
        }
        if constexpr (kMode == kBare) { umma_commit<kCta>(done_bar, 0x1); mbar_wait(done_bar, 0, 0); }
        else mbar_wait(done_bar, (iters - 1) & 1, 0);
      } else if constexpr (kMode == kWait) {
        for (int i = 0; i < iters; ++i) {
          mbar_wait(open_bar, 1, 0);
          tc_fence_after();
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
          umma_commit<kCta>(done_bar, 0x1);
        }
        mbar_wait(done_bar, (iters - 1) & 1, 0);
      } else if constexpr (kMode == kProbeTry || kMode == kProbeTest) {
        bool ready = false;
        for (int i = 0; i < iters; ++i) {
          if (!ready) mbar_wait(open_bar, 1, 0);
          tc_fence_after();
          ready = kMode == kProbeTry ? mbar_try_wait(open_bar, 1) : mbar_test_wait(open_bar, 1);
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
          umma_commit<kCta>(done_bar, 0x1);
        }
        mbar_wait(done_bar, (iters - 1) & 1, 0);
      } else if constexpr (kMode == kTwoBlocks || kMode == kFourBlocks) {
        constexpr int g = kMode == kTwoBlocks ? 2 : 4;
        for (int i = 0; i < iters; i += g) {
#pragma unroll
          for (int j = 0; j < g; ++j) mbar_wait(open_bar, 1, 0);
          tc_fence_after();
#pragma unroll
          for (int j = 0; j < g; ++j) {
            umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, (i + j) ? 1u : 0u);
            umma_commit<kCta>(done_bar, 0x1);
          }
        }
        mbar_wait(done_bar, (iters - 1) & 1, 0);
      } else if constexpr (kMode == kHelperFlag) {
        for (int i = 0; i < iters; ++i) {
          while (ready_upto <= (uint32_t)i) {}
          tc_fence_after();
          umma_ss_x4<kCta, kKindF16>(tmem, a_w, hi, b_w, hi, 2u, 2u, idesc, i ? 1u : 0u);
          umma_commit<kCta>(done_bar, 0x1);
        }
        mbar_wait(done_bar, (iters - 1) & 1, 0);
      }
      if (blockIdx.x == 0) cycles[0] = clock64() - t0;
    }
  } else if (rank == 0 && warp == 2 && kMode == kHelperFlag) {
    // the helper: waits on each K block's barrier and publishes how many K blocks are ready
    if (elect_one()) {
      for (int i = 0; i < iters; ++i) {
        mbar_wait(open_bar, 1, 0);
        ready_upto = i + 1;
      }
    }
  }
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 1) tmem_dealloc<kCta>(tmem, 512);
}

static uint32_t idesc_of(int m, int n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t(n >> 3) & 0x3fu) << 17) | ((uint32_t(m >> 4) & 0x1fu) << 24);
}

template <int kCta, int kMode>
static double run(int m, int n, int iters, long long* cyc) {
  auto kern = loop_study<kCta, kMode>;
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

template <int kMode>
static void row(int iters, long long* cyc) {
  const double c1 = run<1, kMode>(128, 64, iters, cyc), c2 = run<2, kMode>(256, 64, iters, cyc), c3 = run<2, kMode>(256, 128, iters, cyc);
  printf("  %-52s %16.1f %16.1f %16.1f\n", kNames[kMode], c1, c2, c3);
  if constexpr (kMode + 1 < kModes) row<kMode + 1>(iters, cyc);
}

int main() {
  long long* cyc; cudaMalloc(&cyc, 8);
  printf("cycles per 128-byte K block (four bf16 MMAs), operands resident, 148 SMs busy\n  %-52s %16s %16s %16s\n", "issue loop", "cg1 M=128 N=64",
         "cg2 M=256 N=64", "cg2 M=256 N=128");
  row<0>(2000, cyc);
  return 0;
}
