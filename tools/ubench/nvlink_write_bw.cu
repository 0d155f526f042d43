// Microbenchmark (2 GPUs, one process): how fast can one GPU WRITE into a peer's memory over NVLink, by the
// path the fused GEMM + all-gather epilogue could use?  Both GPUs write to each other at the same time.
//   copy engine : cudaMemcpyPeerAsync
//   st.global   : every thread stores 16 B (coalesced 512 B per warp)
//   bulk R      : one thread per warp issues cp.async.bulk shared->global of R contiguous bytes per "row",
//                 rows 16 KiB apart in the destination (the pitch of a bf16 C row of 8192 columns)
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench/_build/nvlink_write_bw tools/ubench/nvlink_write_bw.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void st_kernel(uint4* dst, size_t n16) {
  uint4 v = make_uint4(threadIdx.x, blockIdx.x, 3, 4);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = v;
}

// Each warp owns a 4 KiB slab of shared memory and repeatedly stores it as (4096 / row_bytes) rows of row_bytes.
__global__ void bulk_kernel(char* dst, size_t total_bytes, int row_bytes, size_t pitch) {
  extern __shared__ __align__(128) char smem[];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  char* slab = smem + warp * 4096;
  for (int i = lane; i < 1024; i += 32) reinterpret_cast<uint32_t*>(slab)[i] = i;
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncwarp();
  const int rows_per_slab = 4096 / row_bytes;
  const size_t slabs = total_bytes / 4096;
  const size_t cols_per_row = pitch / row_bytes;  // slabs tile the [rows x pitch] matrix: row-block x column
  const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(slab));
  for (size_t t = blockIdx.x * (size_t)nwarps + warp; t < slabs; t += (size_t)gridDim.x * nwarps) {
    const size_t rb = t / cols_per_row, cb = t % cols_per_row;
    if (lane == 0) {
      for (int r = 0; r < rows_per_slab; ++r) {
        char* g = dst + (rb * rows_per_slab + r) * pitch + cb * row_bytes;
        asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(g), "r"(s + r * row_bytes), "r"(row_bytes) : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
    __syncwarp();
  }
  if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
}

int main() {
  int n = 0;
  CK(cudaGetDeviceCount(&n));
  if (n < 2) { printf("needs 2 GPUs\n"); return 0; }
  const size_t bytes = 512ull << 20;
  char* buf[2]; char* src[2];
  cudaStream_t st[2]; cudaEvent_t e0[2], e1[2];
  for (int d = 0; d < 2; ++d) {
    CK(cudaSetDevice(d));
    CK(cudaDeviceEnablePeerAccess(1 - d, 0));
    CK(cudaMalloc(&buf[d], bytes)); CK(cudaMalloc(&src[d], bytes));
    CK(cudaMemset(src[d], 1, bytes));
    CK(cudaStreamCreate(&st[d])); CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
    CK(cudaFuncSetAttribute(bulk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  }
  auto run = [&](const char* name, int dirs, auto launch) {
    float best = 1e9f;
    for (int rep = 0; rep < 5; ++rep) {
      for (int d = 0; d < dirs; ++d) { CK(cudaSetDevice(d)); CK(cudaDeviceSynchronize()); }
      for (int d = 0; d < dirs; ++d) { CK(cudaSetDevice(d)); CK(cudaEventRecord(e0[d], st[d])); launch(d); CK(cudaEventRecord(e1[d], st[d])); }
      float worst = 0;
      for (int d = 0; d < dirs; ++d) { CK(cudaSetDevice(d)); CK(cudaStreamSynchronize(st[d])); float ms; CK(cudaEventElapsedTime(&ms, e0[d], e1[d])); worst = ms > worst ? ms : worst; }
      if (rep > 0 && worst < best) best = worst;
    }
    printf("%-44s %s  %.3f ms  %.1f GB/s per GPU\n", name, dirs == 2 ? "both directions" : "one direction  ", best, bytes / best / 1e6);
  };
  for (int dirs = 1; dirs <= 2; ++dirs) {
    run("copy engine (cudaMemcpyPeerAsync)", dirs, [&](int d) { CK(cudaMemcpyPeerAsync(buf[1 - d], 1 - d, src[d], d, bytes, st[d])); });
    run("st.global.v4 (148x8 blocks x 256 thr)", dirs, [&](int d) { st_kernel<<<148 * 8, 256, 0, st[d]>>>(reinterpret_cast<uint4*>(buf[1 - d]), bytes / 16); });
    for (int rb : {128, 256, 512, 1024, 4096}) {
      char nm[64]; snprintf(nm, sizeof nm, "bulk store, %4d B rows at 16 KiB pitch", rb);
      run(nm, dirs, [&](int d) { bulk_kernel<<<148, 256, 8 * 4096, st[d]>>>(buf[1 - d], bytes, rb, 16384); });
    }
    run("bulk store, 4096 B rows, local memory", dirs, [&](int d) { bulk_kernel<<<148, 256, 8 * 4096, st[d]>>>(src[d], bytes, 4096, 16384); });
  }
  return 0;
}
