// Microbenchmark: tcgen05.ld throughput per SM (bytes of TMEM read per clock) with 4 or 8 reading warps.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I tilelang_mesh_b200/csrc -o tools/ubench/_build/tmem_read_bw tools/ubench/tmem_read_bw.cu
#include <cstdio>
#include <cuda.h>
#include "ptx.cuh"
using namespace tlb200;

template <int kInFlight>
__global__ void __launch_bounds__(384, 1) k(int iters, unsigned* out, long long* cycles) {
  __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  if (warp == 0) tmem_alloc<1>(smem_u32(&tptr), 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t base = tptr + ((uint32_t)((warp & 3) * 32) << 16);
  unsigned acc = 0;
  long long t0 = clock64();
  if (warp >= 4) {
    for (int i = 0; i < iters; ++i) {
      uint32_t r[kInFlight][32];
#pragma unroll
      for (int j = 0; j < kInFlight; ++j) tmem_ld_32x32(base + (((warp >> 2) - 1) * 128 + ((i * kInFlight + j) & 3) * 32), r[j]);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < kInFlight; ++j)
#pragma unroll
        for (int q = 0; q < 32; ++q) acc ^= r[j][q];
    }
  }
  long long t1 = clock64();
  if (acc == 0x12345678u) out[0] = acc;
  if (threadIdx.x == 128 && blockIdx.x == 0) cycles[0] = t1 - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<1>(tptr, 512);
}

int main() {
  unsigned* out; long long* cyc;
  cudaMalloc(&out, 4); cudaMalloc(&cyc, 8);
  const int iters = 20000;
  for (int threads : {256, 384}) {
    for (int inflight : {1, 2, 4}) {
      for (int rep = 0; rep < 2; ++rep) {
        if (inflight == 1) k<1><<<148, threads>>>(iters, out, cyc);
        else if (inflight == 2) k<2><<<148, threads>>>(iters, out, cyc);
        else k<4><<<148, threads>>>(iters, out, cyc);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
      }
      long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      const int warps = threads / 32 - 4;
      const double bytes = (double)iters * inflight * 4096.0 * warps;
      printf("reading warps %d, x32 loads in flight %d: %.1f B/clk/SM (%lld cycles)\n", warps, inflight, bytes / c, c);
    }
  }
  return 0;
}
