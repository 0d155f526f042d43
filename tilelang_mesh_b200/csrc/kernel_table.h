// kernel_table.h - the kernel instantiations live in their own translation units (kernels_*.cu) so that the
// library builds in parallel; the host side (tlb200.cu) reaches them through these getters.
#pragma once
#include "gemm_blockscaled_sm100.cuh"
#include "gemm_chain_sm100.cuh"
#include "gemm_sm100.cuh"

namespace tlb200 {
typedef void (*GemmKernelFn)(const GemmKernelParams);
typedef void (*BlockScaledFn)(const BlockScaledParams);
typedef void (*ChainFn)(const ChainParams);
// gemm_tcgen05_kernel<cta_group, kind, accumulators per tile, 1>
GemmKernelFn kernel_cg1_ns1(int kind);
GemmKernelFn kernel_cg1_ns2(int kind);
GemmKernelFn kernel_cg2_ns1(int kind);
GemmKernelFn kernel_cg2_ns2(int kind);
// gemm_tcgen05_kernel<2, kind, 1, 2>: pairs, A multicast across the two pairs of a 4-CTA cluster
GemmKernelFn kernel_mc(int kind);
// gemm_blockscaled_kernel<cta_group, block_n>
BlockScaledFn kernel_blockscaled(int cg, int bn);
// gemm_chain_kernel
ChainFn kernel_chain();
}  // namespace tlb200
