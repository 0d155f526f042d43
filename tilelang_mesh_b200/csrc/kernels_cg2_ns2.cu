// Instantiations of gemm_tcgen05_kernel<2, *, 2, 1> (one translation unit per (cta_group, accumulators) pair).
#include "kernel_table.h"

namespace tlb200 {
GemmKernelFn kernel_cg2_ns2(int kind) {
  switch (kind) {
    case kKindF8: return gemm_tcgen05_kernel<2, kKindF8, 2>;
    case kKindTF32: return gemm_tcgen05_kernel<2, kKindTF32, 2>;
    case kKindI8: return gemm_tcgen05_kernel<2, kKindI8, 2>;
    default: return gemm_tcgen05_kernel<2, kKindF16, 2>;
  }
}
}  // namespace tlb200
