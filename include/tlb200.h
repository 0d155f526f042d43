/*
 * tlb200.h — C ABI of libtlb200.so, the B200 (sm_100a) backend for TileLang's
 * dense-contraction path (T.gemm fed by T.copy global->shared staging).
 *
 * Every entry point below replaces one piece of the host C ABI that the
 * reference generates per JIT-compiled kernel, or one compile-time helper that
 * a maintainer needs at the same boundary.  Reference citations are to the
 * Tilelang-Mesh snapshot (TileLang 0.1.7 lineage):
 *
 *   generated host trio   tilelang/jit/adapter/wrapper.py:31-51
 *                         (`init`, `call`, `get_last_error`)
 *   call() body           tilelang/jit/adapter/wrapper.py:202-295
 *   TMA descriptor build  tilelang/jit/adapter/wrapper.py:76-99,
 *                         src/runtime/runtime.cc:98-110
 *   instr descriptor      src/op/tcgen5_meta.h:94-172
 *   atom selection        src/op/tcgen5_meta.h:21-92
 *   smem descriptor       src/tl_templates/cuda/common.h:481-496
 *
 * Conventions (same as the reference's generated code): plain pointers and
 * sizes only, `int` return 0 = ok / non-zero = error, message retrievable as a
 * C string.  Unlike the reference (one process-global 1024-byte buffer) the
 * error string is thread-local.  The library keeps no pointer past a call; all
 * tensors stay owned by the caller (torch in the Python adapter).
 */
#ifndef TLB200_H_
#define TLB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TLB200_ABI_VERSION 2 /* 2: tlb200_gemm_desc gained the epilogue-vector fields (vec_*) and tlb200_gemm_fused */

/* Element types.  Values follow the DLPack-free compact enum used by the
 * Python adapter; names follow tilelang/utils/tensor.py:34-51. */
typedef enum tlb200_dtype {
  TLB200_FLOAT16 = 0,      /* "float16"      */
  TLB200_BFLOAT16 = 1,     /* "bfloat16"     */
  TLB200_FLOAT8_E4M3 = 2,  /* "float8_e4m3"  (torch.float8_e4m3fn) */
  TLB200_FLOAT8_E5M2 = 3,  /* "float8_e5m2"  */
  TLB200_FLOAT32 = 4,      /* "float32"/"float": output; as operand type it is multiplied as TF32 (kind::tf32) */
  TLB200_INT8 = 5,         /* "int8" operands (kind::i8, int32 accumulate) */
  TLB200_INT32 = 6         /* "int32" output of int8 operands (exact) */
} tlb200_dtype;

/* Epilogue flags (fused into the TMEM->global drain). */
#define TLB200_EPI_NONE 0u
#define TLB200_EPI_RELU 1u /* examples/quickstart.py:42-43: C = max(C, 0) */
#define TLB200_EPI_ACCUMULATE 2u /* C = C + alpha * op(A) op(B): what the reference's split-K program does with
                                  * T.atomic_add on a caller-zeroed C (examples/gemm_splitk/
                                  * example_tilelang_gemm_splitk.py:35-40); here one deterministic read-modify-write per
                                  * element (predicated-store epilogue, or the split-K reduction).  Not with RELU,
                                  * int32 output or tlb200_gemm_bcast. */

/*
 * One dense contraction  C[M,N] = alpha * op(A) * op(B)  with fp32 accumulate
 * and a single round-to-nearest-even at the output cast — the whole-kernel
 * meaning of a TileLang GEMM program (T.copy A/B tiles -> T.gemm -> T.copy C).
 *
 *   trans_a == 0 : A is [M,K] row-major (K contiguous), lda = row stride
 *   trans_a == 1 : A is [K,M] row-major (M contiguous), lda = row stride
 *   trans_b == 1 : B is [N,K] row-major (K contiguous)      <- benchmark/matmul
 *   trans_b == 0 : B is [K,N] row-major (N contiguous)      <- examples/quickstart.py
 * Same meaning as transpose_A / transpose_B of T.gemm
 * (tilelang/language/gemm.py:105-116, K_A == K_B check at :78-81).
 * Leading dimensions are in ELEMENTS.
 *
 * The cfg_* fields are optional tuning hints (0 = choose automatically); they
 * play the role of the tile-program's block_M/N/K, num_stages and
 * T.use_swizzle(panel) knobs but never change the numerical contract.
 */
typedef struct tlb200_gemm_desc {
  int32_t abi_version; /* must be TLB200_ABI_VERSION */
  int32_t m, n, k;
  int32_t ab_dtype; /* tlb200_dtype of A and B (must be equal, as T.gemm) */
  int32_t c_dtype;  /* tlb200_dtype of C */
  int32_t trans_a, trans_b;
  int64_t lda, ldb, ldc;
  float alpha;        /* scale_a*scale_b for per-tensor-scaled fp8; 1.0f otherwise */
  uint32_t epilogue;  /* TLB200_EPI_* */
  int32_t cfg_cta_group; /* 0 auto, 1 = single-CTA tiles (128 rows), 2 = 2-CTA cluster tiles (256 rows),
                            3 = 2-CTA cluster tiles of 128 rows (64 per CTA; block_n a multiple of 64, K-contiguous A,
                            one destination; tlb200_gemm_grouped takes it only when asked by name), 4 = two pairs of 2 per cluster
                            on neighbouring column tiles, every A box fetched once and TMA-multicast to both
                            (only when all of A is proven to lie in the current device's own memory, see
                            tlb200_is_own_device_memory - an error otherwise; chosen automatically under the
                            same condition; never by tlb200_gemm_bcast / _grouped; TLB200_NO_MULTICAST=1 in the
                            environment at tlb200_init switches the plan off altogether) */
  int32_t cfg_block_n;   /* 0 auto, else a multiple of 32 in [32,256] or of 64 in (256,512] */
  int32_t cfg_stages;    /* 0 auto */
  int32_t cfg_raster;    /* 0 auto, else tile-rows per L2 raster group */
  int32_t cfg_epilogue;  /* 0 auto, 1 = direct global stores from registers, 2 = swizzled smem slab + TMA store,
                            3 = swizzled smem slab + coalesced st.global (full 128-byte lines, no TMA store) */
  int32_t reserved[3];   /* [0] = cfg_epi_warps: 0 auto, 4 or 8 epilogue warps;
                          * [1] = cfg_split_k: 0 auto, 1 off (no split-K, no tail split), n = at most n K slices,
                          *       -1 = no split-K but force the tail split (both need a workspace);
                          * [2] = cfg_l2_policy: 0 default (operands evict_last, C evict_first), 1 = no hints */
  /* ABI 2: which epilogue vectors tlb200_gemm_fused is given (ignored by every other entry point, which treat
   * them as absent).  All vectors are float32. */
  uint32_t vec_flags;    /* TLB200_VEC_* */
  int32_t reserved2[3];  /* [0] = cfg_kgroup: K blocks (pipeline stages) that share one full / empty barrier pair in
                          *       tiles of at most 128 columns - 0 auto (2; 4 for tiles of <= 32 columns), else 1, 2 or 4
                          *       (lowered when the stage ring would keep fewer than 3 groups; wider tiles always use 1;
                          *       TLB200_KGROUP in the environment at tlb200_init sets the automatic value);
                          * [1], [2] must be 0 */
} tlb200_gemm_desc;

/* Epilogue vectors ("bias / per-row scales"; the per-row form is the scales_a[m, kb] of
 * examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:34,71-73 collapsed to one K block, the per-column
 * form its scales_b at 1-wide groups; a bias has no counterpart in the reference's GEMM examples):
 *
 *   C[m, n] = cast(relu?(((alpha * acc[m, n]) * row_scale[m]) * col_scale[n] + bias))
 *
 * every operation rounded to fp32 in exactly this order (the oracle, oracle/gemm_oracle.c::oracle_gemm_fused, does
 * the same), `bias` = bias_n[n] (TLB200_VEC_BIAS_N, length N) or bias_m[m] (TLB200_VEC_BIAS_M, length M). */
#define TLB200_VEC_ROW_SCALE 1u /* row_scale[M] */
#define TLB200_VEC_COL_SCALE 2u /* col_scale[N] */
#define TLB200_VEC_BIAS_N 4u    /* bias[N]: one value per output column (a linear layer's bias) */
#define TLB200_VEC_BIAS_M 8u    /* bias[M]: one value per output row; not together with TLB200_VEC_BIAS_N */
typedef struct tlb200_epilogue_vectors {
  const float* row_scale; /* device pointers, 4-byte aligned; NULL where the flag is not set */
  const float* col_scale;
  const float* bias;
} tlb200_epilogue_vectors;

/* The configuration the library resolved for a descriptor (for reports/tests). */
typedef struct tlb200_gemm_plan {
  int32_t cta_group, block_m, block_n, block_k, stages;
  int32_t tiles_m, tiles_n, grid, threads, smem_bytes;
  int32_t raster, epilogue, tmem_cols;
  uint32_t instr_desc;
  int32_t umma_n, epi_warps; /* N of one tcgen05.mma (block_n / accumulators per tile); epilogue warps */
  int32_t split_k;           /* K slices per tile (1 = no split-K) */
  int32_t multicast; /* 2 when A boxes are multicast across two pairs (cluster of 4 CTAs), else 1 */
  int64_t workspace_bytes;   /* device scratch tlb200_gemm_ex can use for this plan (split-K, tail split, re-pitched operands) */
} tlb200_gemm_plan;

/* Replaces the generated `int init()` (wrapper.py:40-44): binds the CUDA
 * driver entry points, raises the dynamic-smem limit of every kernel on
 * `device` and caches its SM count.  Idempotent, may be called per device. */
int tlb200_init(int device);

/* Replaces the generated `int call(A, B, C, ..., cudaStream_t)`
 * (wrapper.py:46-51, 202-295): encodes the TMA descriptors on the stack and
 * launches ONE persistent kernel on `stream` (a cudaStream_t; 0 = default) of
 * the CURRENT device.  A, B, C are device pointers.  Asynchronous. */
int tlb200_gemm(const tlb200_gemm_desc* desc, const void* a, const void* b, void* c,
                void* stream);

/* tlb200_gemm with a caller-owned device workspace (scratch memory; 256-byte
 * aligned) of at least tlb200_gemm_workspace_bytes(desc) bytes.  Three things
 * use it:
 *  - split-K (the reference's examples/gemm_splitk idea): when the output has
 *    too few tiles to fill the GPU, K is cut into slices accumulated in fp32 in
 *    the workspace and summed in slice order by a second kernel;
 *  - the tail split (stream-K's remedy for wave quantisation, cf. the reference's
 *    examples/gemm_streamk): when the last wave of whole tiles would occupy
 *    only a fraction of the GPU, just those tiles are cut along K over all
 *    clusters, their fp32 partial tiles go to the workspace and a small second
 *    kernel sums them in slice order.  Optional: a workspace that lacks this
 *    part runs the ordinary schedule.  cfg_split_k = 1 switches it off, -1
 *    forces it whenever the shape allows it;
 *  - operands TMA cannot address (row pitch or base not a multiple of 16 bytes,
 *    e.g. K = 763 fp16 - the reference's test_pad_* cases, which it serves
 *    with SIMT copies, src/op/copy.cc:823-855) are first re-pitched into it.
 * workspace == NULL behaves exactly like tlb200_gemm (which rejects such
 * operands).  The workspace may be reused by the next call on the same
 * stream; calls on different streams need different workspaces.
 * tlb200_gemm_workspace_bytes assumes 16-byte aligned base addresses. */
int64_t tlb200_gemm_workspace_bytes(const tlb200_gemm_desc* desc);
int tlb200_gemm_ex(const tlb200_gemm_desc* desc, const void* a, const void* b, void* c,
                   void* workspace, int64_t workspace_bytes, void* stream);

/* tlb200_gemm_ex with epilogue vectors fused into the TMEM -> global drain (desc->vec_flags says which of
 * `vec`'s pointers are read).  Works with every plan of tlb200_gemm_ex, split-K and the tail split included (their
 * reduction kernels apply the vectors).  Not offered for int32 output or TLB200_EPI_ACCUMULATE.
 * `workspace` may be NULL as in tlb200_gemm_ex. */
int tlb200_gemm_fused(const tlb200_gemm_desc* desc, const void* a, const void* b, void* c,
                      const tlb200_epilogue_vectors* vec, void* workspace, int64_t workspace_bytes, void* stream);

/* Grouped contraction (the reference's examples/grouped_gemm: MoE expert
 * GEMMs): rows of A[sum_rows, K] and C[sum_rows, N] are cut into n_groups
 * consecutive row groups (group_rows[g] rows each, a HOST array, 0 allowed,
 * n_groups <= 256) and group g is multiplied by its own B block:
 *   transpose_B = 1 : B is [n_groups, N, K]      transpose_B = 0 : B is [n_groups, K, N]
 * (blocks contiguous, ldb = row pitch).  desc.m = sum of group_rows.  One
 * persistent launch walks the row tiles of all groups; a tile never spans two
 * groups. */
int tlb200_gemm_grouped(const tlb200_gemm_desc* desc, const void* a, const void* b, void* c,
                        const int32_t* group_rows, int n_groups, void* stream);

/* tlb200_gemm whose C tiles are stored into n_c matrices at once (n_c <= 8):
 * c_list[0] is the local C, the others are the same window of C in PEER GPUs'
 * memory (addresses mapped into this process, e.g. CUDA IPC / symmetric
 * memory).  Every epilogue slab is written by TMA to all destinations, so an
 * N-sharded contraction materialises the full C on every GPU over NVLink while
 * the tensor pipe keeps working - the GEMM and its all-gather are ONE kernel
 * (the reference has no multi-GPU path; this serves north_star's optional
 * N-sharded variant).  All destinations share ldc, shape and dtype.  The
 * caller synchronises the GPUs afterwards (stream completion on every rank +
 * a cross-rank barrier) before reading. */
int tlb200_gemm_bcast(const tlb200_gemm_desc* desc, const void* a, const void* b, void* const* c_list,
                      int n_c, void* stream);

/* fp8 contraction with fine-grained scales ("2x accumulation"):
 *
 *   C[m, n] = sum_kb scales_a[m, kb] * scales_b[n / 128, kb]
 *                    * sum_{k in [128 kb, 128 kb + 128)} A[m, k] * B[n, k]
 *
 * Replaces the tile program of examples/deepseek_deepgemm/
 * example_deepgemm_fp8_2xAcc.py:12-83 (`tl_gemm`): A[M, K] and B[N, K] fp8,
 * K-contiguous; scales_a one fp32 per (row, 128-wide K block), shape
 * (M, ceil(K/128)) (:34); scales_b one fp32 per (128-row block of B, K block),
 * shape (ceil(N/128), ceil(K/128)) (:36); the partial sum of every K block is
 * multiplied by scales_a * scales_b (the product is rounded to fp32 first, as
 * the program's Scale_C_shared, :71-73) and added to an fp32 accumulator
 * (:77-78); one rounding to c_dtype at the end (bfloat16 / float16 / float32;
 * the reference asserts bfloat16 or float32, :24-27).  Scale tensors are
 * addressed through element strides so both the row-major layout of the
 * example and DeepGEMM's M-contiguous layout work.  group_k and group_n must
 * be 128 (the reference's group_size, :29).  cfg_* = 0 selects the default
 * plan (2-CTA pairs, 256-wide tiles). */
typedef struct tlb200_blockscaled_desc {
  int32_t abi_version; /* must be TLB200_ABI_VERSION */
  int32_t m, n, k;
  int32_t ab_dtype; /* TLB200_FLOAT8_E4M3 or TLB200_FLOAT8_E5M2 */
  int32_t c_dtype;  /* TLB200_BFLOAT16, TLB200_FLOAT16 or TLB200_FLOAT32 */
  int32_t group_n, group_k; /* 128, 128 */
  int64_t lda, ldb, ldc;    /* row pitches in elements */
  int64_t sa_stride_m, sa_stride_k; /* scales_a[m, kb] = scales_a[m*sa_stride_m + kb*sa_stride_k] */
  int64_t sb_stride_n, sb_stride_k; /* scales_b[nb, kb] likewise */
  int32_t cfg_cta_group; /* 0 auto, 1, 2 */
  int32_t cfg_block_n;   /* 0 auto, 128, 256 */
  int32_t cfg_stages;    /* 0 auto */
  int32_t cfg_raster;    /* 0 auto */
  int32_t reserved[4];
} tlb200_blockscaled_desc;

int tlb200_gemm_blockscaled(const tlb200_blockscaled_desc* desc, const void* a, const void* b, void* c,
                            const float* scales_a, const float* scales_b, void* stream);

/* Two contractions back to back, the intermediate kept in tensor memory (the "TS" operand mode:
 * the A operand of the second tcgen05.mma comes from TMEM).  The reference declares the instruction wrapper
 * `tl::tcgen05mma_ts` (src/tl_templates/cuda/instruction/tcgen05mma.h:26-133) and the emitter hook
 * (tilelang/intrinsics/tcgen05_macro_generator.py:147-148) but leaves `tcgen5mma_gemm_ts` a TODO
 * (src/tl_templates/cuda/gemm_sm100.h:419-424); this is that operand mode as a whole-problem call:
 *
 *   P[M, N1] = cast_ab(relu?(alpha1 * A[M, K] . B1[N1, K]^T))     fp32 accumulate, ONE rounding to the operand type
 *   D[M, N2] = cast_c (alpha2 * P[M, N1] . B2[N2, N1]^T)          fp32 accumulate, ONE rounding
 *
 * i.e. exactly what two tlb200_gemm calls with an intermediate of the operand type compute (and what the oracle
 * checks), without the intermediate ever leaving the SM: attention's S = Q K^T -> O = P V step without the softmax,
 * or a low-rank product x A^T B^T.  A, B1, B2 are K-contiguous float16 or bfloat16 (B2 is [N2, N1]); N1 is a
 * multiple of 64 and at most 256 (the intermediate of a 128-row tile must fit TMEM beside its fp32 source and one
 * output accumulator); 16-byte aligned bases and row pitches; C float16 / bfloat16 / float32, any ldc. */
typedef struct tlb200_chain_desc {
  int32_t abi_version; /* must be TLB200_ABI_VERSION */
  int32_t m, k, n1, n2;
  int32_t ab_dtype;    /* TLB200_FLOAT16 or TLB200_BFLOAT16 (A, B1, B2 and the intermediate) */
  int32_t c_dtype;     /* TLB200_FLOAT16, TLB200_BFLOAT16 or TLB200_FLOAT32 */
  int32_t relu_mid;    /* 1: relu on the intermediate before it is rounded */
  int64_t lda, ldb1, ldb2, ldc; /* row pitches in elements */
  float alpha1, alpha2;
  int32_t cfg_stages;  /* 0 auto */
  int32_t reserved[3]; /* must be 0 */
} tlb200_chain_desc;

int tlb200_gemm_chain(const tlb200_chain_desc* desc, const void* a, const void* b1, const void* b2, void* c, void* stream);

/* No counterpart in the reference: an allocation-free form of tlb200_gemm_ex.
 *
 * Who needs it: callers that are handed tensor pointers and a stream and nothing
 * else, and are not allowed to obtain device memory themselves.  The typical
 * case is a wrapper that must keep the reference's generated signature
 * `call(A, B, C, stream)` (seam 2): that signature has no
 * workspace argument, and the host framework above it owns every byte of
 * device memory through its own pool, so a cudaMalloc from inside the call is
 * either forbidden by policy or would bypass the pool's accounting.  Latency-
 * critical serving loops that freeze their memory layout after start-up are in
 * the same position.
 *
 * What it does: the split-K partial sums go into a scratch array that is part
 * of the library image itself (TLB200_STATIC_SCRATCH_BYTES, reserved on every
 * GPU that loads the library, for the lifetime of the process).  The call makes
 * no cudaMalloc / cudaFree and uses no allocating library routine.  Problems
 * whose plan needs no split-K run exactly as tlb200_gemm does.  A plan whose
 * partial sums do not fit the static scratch is REJECTED with an error that
 * names both sizes (set cfg_split_k lower, or use tlb200_gemm_ex with your own
 * workspace); nothing is silently truncated.
 *
 * Rules of use:
 *  - The first call on each device looks up the device address of the scratch
 *    array (a one-time set-up step on the then-current device).  Make that first
 *    call during start-up, before any stream capture begins.
 *  - All calls on one device share the one array and it keeps whatever the
 *    previous call left in it.  Nothing in it has to start at zero: every slice
 *    overwrites its whole partial tile before the reduction reads it.  Two calls
 *    that can overlap on the GPU (different streams, no ordering between them)
 *    race on the array - order them, or give one of them tlb200_gemm_ex. */
#define TLB200_STATIC_SCRATCH_BYTES (32u << 20)
int tlb200_gemm_static_scratch(const tlb200_gemm_desc* desc, const void* a, const void* b, void* c,
                               void* stream);

/* Same contract through HOST buffers (pinned memory for full speed): stages B,
 * then A in row panels, to the device; runs tlb200_gemm per panel on `stream`
 * while the next panel is copied in and the previous panel's C is copied out
 * (two internal copy streams created by tlb200_init), then synchronises.
 * Operands whose host row pitch TMA cannot address (not a multiple of 16
 * bytes) are re-pitched by the copy itself.
 * `workspace` is a caller-owned device buffer of at least
 * tlb200_gemm_host_workspace_bytes(desc) bytes (the library never allocates).
 * This is the host-buffer form of the reference's adapter call path
 * (tilelang/jit/adapter/cython/cython_wrapper.pyx:164-287 with CPU tensors moved
 * by the caller) and is what bench.py times as `e2e`. */
int64_t tlb200_gemm_host_workspace_bytes(const tlb200_gemm_desc* desc);
int tlb200_gemm_host(const tlb200_gemm_desc* desc, const void* host_a, const void* host_b,
                     void* host_c, void* workspace, int64_t workspace_bytes, void* stream);

/* Resolve the plan without launching (pure host arithmetic; no GPU needed): what the reference reports through
 * its tile-program constants and `JITKernel.get_kernel_source()` (tilelang/jit/kernel.py:420-440).  Nothing is
 * known here about where A lives, so `multicast` is 2 only when the descriptor asks for it (cfg_cta_group = 4). */
int tlb200_gemm_plan_query(const tlb200_gemm_desc* desc, int sm_count, tlb200_gemm_plan* out);

/* The plan a tlb200_gemm / tlb200_gemm_ex call with exactly these pointers would run on the current device (no
 * launch).  Same role as above; it exists because the A-multicast plan (`multicast` = 2) depends on the pointer:
 * TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
 * memory hung the GPU, so that plan is taken only when the whole extent of A (or of its re-pitched copy in the
 * workspace) is proven to be one cudaMalloc allocation of the current device - see tlb200_is_own_device_memory. */
int tlb200_gemm_plan_query_call(const tlb200_gemm_desc* desc, const void* a, const void* b, const void* c,
                                const void* workspace, int64_t workspace_bytes, tlb200_gemm_plan* out);

/* No counterpart in the reference (it has no multi-GPU path).  1 when [ptr, ptr + bytes) is proven to be ordinary
 * memory of the CURRENT device allocated by this process with cudaMalloc / cuMemAlloc and lying inside one
 * allocation; 0 for everything else - host, managed, stream-ordered-pool, virtual-memory-mapped (cuMemMap:
 * symmetric memory, expandable segments), imported (IPC) and peer memory, ranges that cross an allocation
 * boundary, and any query error.  This is the condition under which the library uses TMA multicast loads. */
int tlb200_is_own_device_memory(const void* ptr, int64_t bytes);

/* Replaces `const char* get_last_error()` (wrapper.py:36-38). Thread-local. */
const char* tlb200_last_error(void);

/* ABI version of the loaded library. */
int tlb200_abi_version(void);

/* Number of kernels this library has launched in the calling process (for the
 * `gpu_launches` figure bench.py reports). */
int64_t tlb200_launch_count(void);

/* Restates GetTCGEN5InstrDesc (src/op/tcgen5_meta.h:94-172): the 32-bit
 * tcgen05.mma instruction descriptor.  Pure arithmetic; no GPU needed.
 * Returns 0 on unsupported arguments (and sets the error string). */
uint32_t tlb200_instr_desc(int atom_m, int atom_n, int atom_k, int ab_dtype, int c_dtype,
                           int a_is_k_major, int b_is_k_major, int scale_in_a, int scale_in_b);

/* Restates GetTCGEN5MMAMeta (src/op/tcgen5_meta.h:21-92): atom the REFERENCE
 * would pick for a (M,N,K) tile.  out5 = {atom_m, atom_n, atom_k, ws, 2cta}.
 * Returns 1 if supported, 0 if not. */
int tlb200_tcgen5_meta(int m, int n, int k, int ab_dtype, int c_dtype, int32_t* out5);

/* Restates initialize_tcgen05_descriptor (src/tl_templates/cuda/common.h:481-496):
 * 64-bit shared-memory matrix descriptor from a shared-window byte address. */
uint64_t tlb200_smem_desc(uint32_t smem_addr, int leading_byte_offset_16, int stride_byte_offset_16,
                          int base_offset, int leading_is_absolute, int swizzle_mode);

#ifdef __cplusplus
}
#endif
#endif /* TLB200_H_ */
