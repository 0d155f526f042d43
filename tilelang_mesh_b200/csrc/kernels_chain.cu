// The chained-contraction kernel (A operand of the second MMA from tensor memory) in its own translation unit.
#include "kernel_table.h"

namespace tlb200 {
ChainFn kernel_chain() { return gemm_chain_kernel<0>; }
}  // namespace tlb200
