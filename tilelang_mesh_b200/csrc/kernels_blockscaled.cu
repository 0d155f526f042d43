// Instantiations of gemm_blockscaled_kernel<cta_group, block_n>.
#include "kernel_table.h"

namespace tlb200 {
BlockScaledFn kernel_blockscaled(int cg, int bn) {
  if (cg == 2) return bn == 128 ? gemm_blockscaled_kernel<2, 128> : gemm_blockscaled_kernel<2, 256>;
  return bn == 128 ? gemm_blockscaled_kernel<1, 128> : gemm_blockscaled_kernel<1, 256>;
}
}  // namespace tlb200
