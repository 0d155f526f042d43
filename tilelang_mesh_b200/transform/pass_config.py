"""Pass-config keys (tilelang/transform/pass_config.py:9-152).

There is no TIR pass pipeline behind this backend; the keys exist so programs that pass
`pass_configs={...}` (the reference's sm_100 example must disable TMA lowering and warp
specialisation: examples/gemm_sm100/README.md:86-89) keep working.  They are accepted and ignored —
TMA staging and warp specialisation are always on in the hand-written kernel."""
from enum import Enum


class PassConfigKey(str, Enum):
    TL_SIMPLIFY = "tl.Simplify"
    TL_DISABLE_WARP_SPECIALIZED = "tl.disable_warp_specialized"
    TL_DISABLE_TMA_LOWER = "tl.disable_tma_lower"
    TL_DISABLE_FAST_MATH = "tl.disable_fast_math"
    TL_ENABLE_FAST_MATH = "tl.enable_fast_math"
    TL_PTXAS_REGISTER_USAGE_LEVEL = "tl.ptxas_register_usage_level"
    TL_ENABLE_PTXAS_VERBOSE_OUTPUT = "tl.enable_ptxas_verbose_output"
    TL_CONFIG_INDEX_BITWIDTH = "tl.config_index_bitwidth"
    TL_DISABLE_SAFE_MEMORY_ACCESS = "tl.disable_safe_memory_legalize"
    TL_DISABLE_VECTORIZE_256 = "tl.disable_vectorize_256"
    TL_DISABLE_DYNAMIC_TAIL_SPLIT = "tl.disable_dynamic_tail_split"
    TL_DYNAMIC_ALIGNMENT = "tl.dynamic_alignment"
    TL_ENABLE_AGGRESSIVE_SHARED_MEMORY_MERGE = "tl.enable_aggressive_shared_memory_merge"
    TL_DISABLE_SHUFFLE_ELECT = "tl.disable_shuffle_elect"
    TL_DISABLE_THREAD_STORAGE_SYNC = "tl.disable_thread_storage_sync"
    TL_FORCE_LET_INLINE = "tl.force_let_inline"
    TL_DEBUG_MERGE_SHARED_MEMORY_ALLOCATIONS = "tl.debug_merge_shared_memory_allocations"
    TL_STORAGE_REWRITE_DETECT_INPLACE = "tl.storage_rewrite_detect_inplace"
    TL_LAYOUT_VISUALIZATION_ENABLE = "tl.layout_visualization_enable"
    TIR_ENABLE_EQUIV_TERMS_IN_CSE = "tir.enable_equiv_terms_in_cse_tir"
    TIR_DISABLE_CSE = "tir.disable_cse_tir"
    TIR_SIMPLIFY = "tir.Simplify"
    TIR_DISABLE_STORAGE_REWRITE = "tir.disable_storage_rewrite"
    TIR_DISABLE_VECTORIZE = "tir.disable_vectorize"
    TIR_USE_ASYNC_COPY = "tir.use_async_copy"
    TIR_ENABLE_DEBUG = "tir.enable_debug"
    TIR_MERGE_STATIC_SMEM = "tir.merge_static_smem"
    TIR_ADD_LOWER_PASS = "tir.add_lower_pass"
    TIR_NOALIAS = "tir.noalias"
    CUDA_KERNELS_OUTPUT_DIR = "cuda.kernels_output_dir"
