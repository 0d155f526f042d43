// Instantiations of the A-multicast variant gemm_tcgen05_kernel<2, *, 1, 2>.
#include "kernel_table.h"

namespace tlb200 {
GemmKernelFn kernel_mc(int kind) {
  switch (kind) {
    case kKindF8: return gemm_tcgen05_kernel<2, kKindF8, 1, 2>;
    case kKindTF32: return gemm_tcgen05_kernel<2, kKindTF32, 1, 2>;
    case kKindI8: return gemm_tcgen05_kernel<2, kKindI8, 1, 2>;
    default: return gemm_tcgen05_kernel<2, kKindF16, 1, 2>;
  }
}
}  // namespace tlb200
