// Instantiations of gemm_tcgen05_kernel<2, *, 1, 1> (one translation unit per (cta_group, accumulators) pair).
#include "kernel_table.h"

namespace tlb200 {
GemmKernelFn kernel_cg2_ns1(int kind) {
  switch (kind) {
    case kKindF8: return gemm_tcgen05_kernel<2, kKindF8, 1>;
    case kKindTF32: return gemm_tcgen05_kernel<2, kKindTF32, 1>;
    case kKindI8: return gemm_tcgen05_kernel<2, kKindI8, 1>;
    default: return gemm_tcgen05_kernel<2, kKindF16, 1>;
  }
}
}  // namespace tlb200
