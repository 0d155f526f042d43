// gemm_sm100.cuh — the persistent, warp-specialised tcgen05 GEMM kernel.
//
// One kernel covers the whole TileLang GEMM program  (T.copy A,B tiles -> T.gemm -> T.copy C):
//
//   warps 0,2,3 producers (K blocks round-robin): TMA bulk-tensor loads of the A / B K-blocks into 128B-swizzled smem,
//                      multi-stage full/empty mbarrier ring          (replaces T.copy g->s:
//                      src/op/copy.cc:1399-1688, tl::tma_load copy_sm90.h:17-268,
//                      and the cp.async path the reference is forced onto, copy.h:16-78)
//   warp 1  MMA      : one lane issues tcgen05.mma (cta_group::1 or ::2) into a
//                      double-buffered TMEM accumulator; tcgen05.commit frees smem stages
//                      and publishes finished accumulators     (replaces tl::tcgen5mma_gemm_ss
//                      gemm_sm100.h:342-436 and tcgen05_macro_generator.py:140-349)
//   warp 2  also     : tcgen05.alloc / dealloc                  (replaces tcgen_05.h:13-27)
//   warps 4-7 epilogue: tcgen05.ld -> alpha/relu -> cast -> swizzled smem -> TMA store (or
//                      predicated 16 B global stores)           (replaces copy.cc:1191-1377
//                      tmem->fragment, copy.cc:920-958 fragment->shared->global)
//
// Tiles: BLOCK_M = 128 rows per CTA (256 per 2-CTA pair; 64 per CTA / 128 per pair with half_m), BLOCK_N <= 256
// columns per accumulator (two accumulators for 512-wide tiles), BLOCK_K = 128 bytes of K per stage (64 x 16-bit or
// 128 x 8-bit elements) = one 128B swizzle atom, consumed by 4 MMAs of K = 32 bytes each.  A static persistent
// scheduler walks the tile grid in L2-friendly row groups (replaces threadblock_swizzle.h:7-41).
//
// Pipeline protocol = the reference's warp-specialised ring: parity (k / stages) & 1, producer waits the
// inverted parity, expect_tx = bytes of the stage - with `kgroup` consecutive stages behind one full / empty barrier
// pair for narrow tiles, whose K block costs less than a barrier hand-shake does (profiles/r2_notes.md section 12).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>

#include "ptx.cuh"

namespace tlb200 {

constexpr int kBlockM = 128;          // rows of C per CTA (TMEM lanes)
constexpr int kStageRowBytes = 128;   // K bytes per stage row = one 128B swizzle atom
constexpr int kUmmaKBytes = 32;       // K bytes consumed by one tcgen05.mma
constexpr int kMaxThreads = 384;      // 4 control warps + up to 8 epilogue warps
constexpr int kEpiWarp0 = 4;          // first epilogue warp (must be a multiple of 4)
constexpr int kEpiBufBytes = 32 * 128;     // one staging slab: 32 rows x 128 B
constexpr int kEpiSlabs = 8;               // staging slabs per CTA by default: 4 warps x 2 or 8 warps x 1 (32 KiB); 16 = 8 warps x 2
constexpr int kMaxStages = 32;  // small stages (short problems, narrow tiles) want many: cold operands are latency-bound by the bytes in flight
constexpr int kMaxGroups = 256;  // grouped GEMM: row groups of A/C, each with its own B (an expert count of current MoE layers)
constexpr int kMaxPeerDst = 7;  // fused all-gather: a C tile is stored into up to 1 + 7 matrices
constexpr int kABytesPerStage = kBlockM * kStageRowBytes;  // 16 KiB

enum OutKind : int { kOutF16 = 0, kOutBF16 = 1, kOutE4M3 = 2, kOutE5M2 = 3, kOutF32 = 4, kOutI32 = 6 };

struct alignas(64) GemmKernelParams {
  CUtensorMap tm_a;  // operand A boxes
  CUtensorMap tm_b;  // operand B boxes
  CUtensorMap tm_c;  // C store boxes (32 rows x 128 B), valid iff epi_tma
  CUtensorMap tm_c_peer[kMaxPeerDst];  // additional destinations of every C store (peer GPUs' copies of C)
  void* c_peer[kMaxPeerDst];
  int n_peer;        // number of additional destinations (0 = single-GPU store)
  // grouped GEMM (examples/grouped_gemm): rows [group_row[g], group_row[g+1]) of A and C use B block g.
  // tiles_m then counts the row tiles of all groups; group_tile[g] is the first row tile of group g.
  // Tail split: the tiles of the last, partial wave are cut into K slices so that the whole machine works on them
  // (a stream-K style fix of wave quantisation).  Work items [0, tail_begin) are whole tiles; item tail_begin +
  // s * tail_tiles + j is K slice s of tile tail_begin + j and stores raw fp32 partial sums as tile (s * tail_tiles + j)
  // of the workspace matrix [tail_split * tail_tiles * block_m rows, block_n columns]; a reduction kernel finishes them.
  CUtensorMap tm_ws;
  void* ws;
  int tail_begin, tail_tiles, tail_split, tail_kb;  // tail_tiles == 0: off
  int n_groups;      // 0 = ordinary GEMM
  int b_group_rows;  // rows of one group's B in the 2-D view TMA sees (N if K-major, K if N-major)
  void* c;
  long long ldc;  // elements
  int m, n;
  int num_k_blocks;
  int split_k;        // K slices per tile (1 = none).  Slice s covers K-blocks [s*kb_per_split, ...)
  int kb_per_split;
  int split_row_pitch; // split-K partial sums: slice s stores its tile rows at row + s*split_row_pitch of the workspace
  int tiles_m, tiles_n;  // in units of (kBlockM * cta_group) rows, block_n columns
  int raster;            // tile rows per raster group
  int block_n;           // tile width = n_sub * umma_n
  int umma_n;            // N of one tcgen05.mma (<= 256)
  int n_sub;             // accumulators per tile (1, or 2 for 512-wide tiles sharing the A stage)
  int acc_stages;        // 2 = double-buffered accumulator (block_n <= 256), 1 = the tile fills TMEM
  int epi_warps;         // 4 or 8
  int epi_slabs;         // staging slabs per CTA: 8, or 16 (two per warp with 8 epilogue warps)
  int stages;
  int k_elems_per_block;  // 128 / sizeof(ab)
  int a_mn_major, b_mn_major;
  int a_boxes, b_boxes;            // TMA issues per stage for A / B
  int a_box_bytes, b_box_bytes;    // bytes each issue writes
  int b_rows_per_cta;              // umma_n / cta_group: rows of B one CTA stages per accumulator
  int b_sub_bytes;                 // bytes of one accumulator's B rows in a stage
  uint32_t idesc;
  uint32_t desc_hi_a, desc_hi_b;   // upper 32 bits of the smem matrix descriptors (SBO, version, swizzle)
  uint32_t desc_lbo_a, desc_lbo_b; // (LBO >> 4) << 16: the leading-byte-offset field of the lower word
  uint32_t kadv_a, kadv_b;         // (bytes >> 4) to advance desc.lo per MMA along K
  int out_kind;
  int acc_int;  // accumulators hold int32 (int8 operands)
  int epi_tma;
  int epi_lsu;  // with epi_tma: slabs leave shared memory through coalesced st.global (no TMA store)
  int relu;
  int accumulate;  // C = C_in + alpha * acc (predicated-store epilogue only)
  int l2_policy;
  float alpha;
  // epilogue vectors (all fp32, null = absent): v = ((alpha * acc) * row_scale[m]) * col_scale[n] + bias
  const float* row_scale;
  const float* col_scale;
  const float* bias;
  int bias_per_row;  // bias is indexed by the output row (else by the column)
  // ceil(2^64 / d) for the divisors of the work-item decomposition (0 when d <= 1): tiles_mn, raster * tiles_n_c,
  // raster, tiles_m % raster and tail_tiles.  Divisions by run-time values (~60 dependent instructions each) sat
  // between the set-up barrier and the first TMA issue of every CTA.
  unsigned long long magic_tiles_mn, magic_group, magic_raster, magic_last_rows, magic_tail_tiles;
  int a_stage_bytes; // shared-memory bytes between consecutive A stages: 16 KB (128 rows); 8 KB with half_m; the rows actually
                     // fetched, in whole 8-row swizzle atoms, for a problem shorter than one tile (the MMA still reads 128
                     // rows - what lies behind the fetched ones feeds accumulator rows that are never stored)
  int half_m;        // pair tiles of 128 rows: 64 rows per CTA (cta_group::2 MMAs with M = 128).  An accumulator then is
                     // block_n / 2 TMEM columns over all 128 lanes: lanes [0, 64) hold rows 0-63 of the tile columns
                     // [0, block_n / 2), lanes [64, 128) the same rows of columns [block_n / 2, block_n)
  int kgroup;        // K blocks per pipeline barrier (1, 2 or 4): `full` / `empty` [g] cover stages [g*kgroup, (g+1)*kgroup)
  long long* trace;  // development builds (-DTLB200_TRACE): per-tile clock64 stamps of CTA 0; null otherwise
  int debug_mode;    // development builds: 1 = issue no MMAs (stages are handed straight back), 2 = issue no TMA loads
                     // (the stage barrier is just arrived on), 4 / 8 = one / two producer warps instead of three
  // Last, so that the scalars every CTA reads before its first TMA issue sit in a few adjacent constant-cache lines.
  int group_row[kMaxGroups + 1];
  int group_tile[kMaxGroups + 1];
};

#ifdef TLB200_TRACE
// slot layout per tile of CTA 0 (16 slots): 0 MMA warp holds both accumulators, 1 first K block landed, 2 last commit
// issued, 3/4 epilogue warp 0 sees accumulator 0 full / has drained it, 5/6 same for accumulator 1, 7-10 the same four
// for epilogue warp 4 (the other column half), 11 producer warp 0 issues its first K block of the tile
#define TLB_TRACE(seq, slot) do { if (p.trace && blockIdx.x == 0 && lane == 0) p.trace[(seq) * 16 + (slot)] = clock64(); } while (0)
#else
#define TLB_TRACE(seq, slot) do { } while (0)
#endif

// ---------------------------------------------------------------- output conversion helpers
template <int kOut>
struct OutTraits;
template <>
struct OutTraits<kOutF16> { static constexpr int kBytes = 2; };
template <>
struct OutTraits<kOutBF16> { static constexpr int kBytes = 2; };
template <>
struct OutTraits<kOutE4M3> { static constexpr int kBytes = 1; };
template <>
struct OutTraits<kOutE5M2> { static constexpr int kBytes = 1; };
template <>
struct OutTraits<kOutF32> { static constexpr int kBytes = 4; };
template <>
struct OutTraits<kOutI32> { static constexpr int kBytes = 4; };

// Convert 16/sizeof(out) consecutive fp32 values to one 16-byte packet, round-to-nearest-even
// (reference epilogue: __float22bfloat162_rn / __float22half2_rn in generated code; fp8 through
// cute::float_e4m3_t(float), common.h:505-523 = cvt.rn.satfinite).
template <int kOut>
__device__ __forceinline__ uint4 pack16(const float* v) {
  uint4 r;
  if constexpr (kOut == kOutBF16) {
    __nv_bfloat162 p0 = __float22bfloat162_rn(make_float2(v[0], v[1]));
    __nv_bfloat162 p1 = __float22bfloat162_rn(make_float2(v[2], v[3]));
    __nv_bfloat162 p2 = __float22bfloat162_rn(make_float2(v[4], v[5]));
    __nv_bfloat162 p3 = __float22bfloat162_rn(make_float2(v[6], v[7]));
    r.x = *reinterpret_cast<uint32_t*>(&p0);
    r.y = *reinterpret_cast<uint32_t*>(&p1);
    r.z = *reinterpret_cast<uint32_t*>(&p2);
    r.w = *reinterpret_cast<uint32_t*>(&p3);
  } else if constexpr (kOut == kOutF16) {
    __half2 p0 = __float22half2_rn(make_float2(v[0], v[1]));
    __half2 p1 = __float22half2_rn(make_float2(v[2], v[3]));
    __half2 p2 = __float22half2_rn(make_float2(v[4], v[5]));
    __half2 p3 = __float22half2_rn(make_float2(v[6], v[7]));
    r.x = *reinterpret_cast<uint32_t*>(&p0);
    r.y = *reinterpret_cast<uint32_t*>(&p1);
    r.z = *reinterpret_cast<uint32_t*>(&p2);
    r.w = *reinterpret_cast<uint32_t*>(&p3);
  } else if constexpr (kOut == kOutF32 || kOut == kOutI32) {  // I32: v[] carries the raw integer bits
    r.x = __float_as_uint(v[0]);
    r.y = __float_as_uint(v[1]);
    r.z = __float_as_uint(v[2]);
    r.w = __float_as_uint(v[3]);
  } else {
    constexpr __nv_fp8_interpretation_t kI = (kOut == kOutE4M3) ? __NV_E4M3 : __NV_E5M2;
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t lo = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * i + 0], v[4 * i + 1]), __NV_SATFINITE, kI);
      uint32_t hi = __nv_cvt_float2_to_fp8x2(make_float2(v[4 * i + 2], v[4 * i + 3]), __NV_SATFINITE, kI);
      w[i] = (lo & 0xffffu) | (hi << 16);
    }
    r.x = w[0]; r.y = w[1]; r.z = w[2]; r.w = w[3];
  }
  return r;
}

template <int kOut>
__device__ __forceinline__ void store_scalar(void* c, long long idx, float v) {
  if constexpr (kOut == kOutBF16) {
    reinterpret_cast<__nv_bfloat16*>(c)[idx] = __float2bfloat16_rn(v);
  } else if constexpr (kOut == kOutF16) {
    reinterpret_cast<__half*>(c)[idx] = __float2half_rn(v);
  } else if constexpr (kOut == kOutF32 || kOut == kOutI32) {
    reinterpret_cast<float*>(c)[idx] = v;
  } else {
    constexpr __nv_fp8_interpretation_t kI = (kOut == kOutE4M3) ? __NV_E4M3 : __NV_E5M2;
    reinterpret_cast<uint8_t*>(c)[idx] = (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, kI);
  }
}

template <int kOut>
__device__ __forceinline__ float load_out(const void* c, long long idx) {
  if constexpr (kOut == kOutBF16) return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(c)[idx]);
  else if constexpr (kOut == kOutF16) return __half2float(reinterpret_cast<const __half*>(c)[idx]);
  else if constexpr (kOut == kOutF32 || kOut == kOutI32) return reinterpret_cast<const float*>(c)[idx];
  else {
    constexpr __nv_fp8_interpretation_t kI = (kOut == kOutE4M3) ? __NV_E4M3 : __NV_E5M2;
    return __half2float(__half(__nv_cvt_fp8_to_halfraw(reinterpret_cast<const uint8_t*>(c)[idx], kI)));
  }
}

// Where and how one tile's accumulator is written: the final C (cast, alpha, relu, optional += C), or - for the K
// slices of a split tail tile - raw fp32 partial sums into the caller's workspace.
struct EpiOut {
  const CUtensorMap* tm;  // TMA-store map (valid iff tma)
  void* c;
  long long ldc;
  int n;                  // columns of the destination matrix (stores are clipped to it)
  float alpha;
  int relu, accumulate, tma, n_peer;
  int lsu;                // staged slabs leave shared memory through coalesced st.global instead of a TMA store
  int evict_first;        // C is write-once: keep it from displacing the operands in L2
  const float* row_scale; // epilogue vectors of the final C (null = absent; never set for raw partial sums)
  const float* col_scale;
  const float* bias;
  int bias_per_row;
  int trace_seq;          // development builds: tile sequence number for TLB_TRACE, -1 = do not trace this call
};
#ifdef TLB200_TRACE
#define TLB_TRACE_EPI(slot) do { if (o.trace_seq >= 0 && cs == cbeg + kColsPerSlab) TLB_TRACE(o.trace_seq, slot); } while (0)
#else
#define TLB_TRACE_EPI(slot) do { } while (0)
#endif

// Epilogue vectors on 32 consecutive columns [gcol, gcol + 32) of global row `grow` (after alpha, before relu and the
// cast): v = (v * row_scale[grow]) * col_scale[col] + bias, each step rounded to fp32.  Columns / rows outside the
// matrix read nothing (their values are never stored).  Warp-uniform branches: a GEMM without vectors pays nothing.
__device__ __forceinline__ void epi_apply_vectors(float (&v)[32], const EpiOut& o, int grow, int row_limit, int gcol) {
  if (o.row_scale != nullptr) {
    const float rs = grow < row_limit ? __ldg(o.row_scale + grow) : 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] * rs;
  }
  if (o.col_scale != nullptr) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = v[i] * (gcol + i < o.n ? __ldg(o.col_scale + gcol + i) : 0.f);
  }
  if (o.bias != nullptr) {
    if (o.bias_per_row) {
      const float bv = grow < row_limit ? __ldg(o.bias + grow) : 0.f;
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = v[i] + bv;
    } else {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = v[i] + (gcol + i < o.n ? __ldg(o.bias + gcol + i) : 0.f);
    }
  }
}

// 32 accumulator elements (one tcgen05.ld) -> the values to be cast: alpha, relu.  Kept to ONE instruction per
// element on the ordinary path (an FMUL): two epilogue warps share a scheduler and every instruction here is
// issued 32-64 times per slab, so a per-element select between integer and float accumulators or a per-element
// relu test made the drain of a 128 x 256 accumulator issue-bound (4.5k cycles, profiles/r2_notes.md).  kAccInt is
// a property of the kernel instantiation (int8 operands); relu is a warp-uniform branch around a second pass.
// int32 accumulators are converted to float unless the output is int32 too: then the raw integer bits are carried.
template <int kOut, bool kAccInt>
__device__ __forceinline__ void epi_values(const uint32_t (&raw)[32], float (&v)[32], float alpha, bool relu, const EpiOut* vec = nullptr,
                                           int grow = 0, int row_limit = 0, int gcol = 0) {
  if constexpr (kOut == kOutI32) {
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(raw[i]);
    if (relu) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = __int_as_float(max(__float_as_int(v[i]), 0));
    }
  } else {
#pragma unroll
    for (int i = 0; i < 32; ++i)
      v[i] = (kAccInt ? static_cast<float>(static_cast<int>(raw[i])) : __uint_as_float(raw[i])) * alpha;
    if (vec != nullptr && (vec->row_scale != nullptr || vec->col_scale != nullptr || vec->bias != nullptr))
      epi_apply_vectors(v, *vec, grow, row_limit, gcol);
    if (relu) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = fmaxf(v[i], 0.f);
    }
  }
}

// ---------------------------------------------------------------- epilogue for one tile
// What one epilogue warp needs to drain its share of an accumulator (always inlined: a context passed by reference).
struct EpiCtx {
  const GemmKernelParams& p;
  const EpiOut& o;
  uint32_t lane_taddr;  // TMEM address of this warp's lane quarter, column 0 of the accumulator
  int q, lane, row0, row_limit, row_off, col0;  // q: which 32 rows of the CTA's part of the tile
  uint32_t stage_smem;
  int nslabs;
};

// registers -> swizzled slab -> global (TMA store, or coalesced st.global)
template <int kOut>
__device__ __forceinline__ void epi_stage(const EpiCtx& e, int cs, const uint4 (&w)[8], int& slab_iter) {
  constexpr int kOB = OutTraits<kOut>::kBytes;
  const EpiOut& o = e.o;
  const int lane = e.lane;
  const uint32_t slab = e.stage_smem + (e.nslabs == 2 ? (slab_iter & 1) : 0) * kEpiBufBytes;
  const int first_row = e.row0 + e.q * 32;
  if (!o.lsu) {  // the TMA stores that last read this slab must have drained it
    // (bulk groups are per thread: every lane that issues stores waits for its own)
    if (lane <= o.n_peer) {
      if (e.nslabs == 2) tma_store_wait_read<1>(); else tma_store_wait_read<0>();
    }
    __syncwarp();
  }
#pragma unroll
  for (int chunk = 0; chunk < 8; ++chunk) {  // 16 B chunk index in the 128 B row
    const uint32_t addr = slab + lane * 128 + ((chunk ^ (lane & 7)) << 4);
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[chunk].x), "r"(w[chunk].y),
                 "r"(w[chunk].z), "r"(w[chunk].w)
                 : "memory");
  }
  if (o.lsu) {
    // Coalesced read-back: every instruction moves 4 rows x 128 B (8 lanes x 16 B per row), so each row segment
    // reaches L2 as one full 128-byte line, and no asynchronous agent reads the slab.  Measured equal to the TMA
    // store on one GPU (profiles/r2_notes.md); kept as cfg_epilogue = 3.
    __syncwarp();
    const int rsub = lane >> 3, chunk = lane & 7;
    const long long row_bytes = o.ldc * kOB;
    const long long col_byte = static_cast<long long>(e.col0 + cs) * kOB + chunk * 16;
    const bool col_ok = col_byte < static_cast<long long>(o.n) * kOB;  // the row length is a multiple of 16 bytes
    char* gptr = reinterpret_cast<char*>(o.c) + static_cast<long long>(e.row_off + first_row + rsub) * row_bytes + col_byte;
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int row = j * 4 + rsub;
      const uint32_t addr = slab + row * 128 + ((chunk ^ (row & 7)) << 4);
      asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "r"(addr) : "memory");
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (col_ok && first_row + j * 4 + rsub < e.row_limit)
        asm volatile("st.global.cs.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(gptr + static_cast<long long>(j) * 4 * row_bytes),
                     "r"(v[j].x), "r"(v[j].y), "r"(v[j].z), "r"(v[j].w) : "memory");
    }
    // Fused all-gather through the load/store unit: the same registers go to every peer GPU's C as full 128-byte lines
    // (st.global reaches 701 GB/s per GPU over NVLink, bulk stores of 128-byte rows 669: profiles/nvlink_write_bw_r2.txt)
    // and the TMA unit stays free for the operand loads.  Destinations are visited in an order rotated per CTA and slab.
    for (int gi = 0; gi < o.n_peer; ++gi) {
      const int g = (gi + static_cast<int>(blockIdx.x) + slab_iter) % o.n_peer;
      char* pptr = gptr - reinterpret_cast<char*>(o.c) + reinterpret_cast<char*>(e.p.c_peer[g]);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (col_ok && first_row + j * 4 + rsub < e.row_limit)
          asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};" ::"l"(pptr + static_cast<long long>(j) * 4 * row_bytes),
                       "r"(v[j].x), "r"(v[j].y), "r"(v[j].z), "r"(v[j].w) : "memory");
      }
    }
    __syncwarp();  // the slab may be rewritten once every lane has its data in registers
    ++slab_iter;
    return;
  }
  fence_proxy_async_smem();  // generic-proxy smem writes -> visible to the TMA engine
  __syncwarp();
  if (o.n_peer == 0) {
    if (lane == 0) {
      if (first_row < e.row_limit) {
        // C is write-once: evict_first keeps the operands in L2
        if (o.evict_first) tma_store_2d_hint(o.tm, slab, e.col0 + cs, e.row_off + first_row, policy_evict_first());
        else tma_store_2d(o.tm, slab, e.col0 + cs, e.row_off + first_row);
      }
      tma_store_commit();
    }
  } else if (lane <= o.n_peer) {
    // Fused all-gather: the same slab goes into the local C and straight into every peer GPU's C over NVLink.
    // One lane per destination issues (and later waits for) its own store, so the up-to-8 stores of a slab enter
    // the TMA queue together instead of one after another from a single thread; the lane -> destination map is
    // rotated per CTA and slab so that the CTAs of a wave do not all address the same peer at the same moment.
    if (first_row < e.row_limit) {
      const int dst = (lane + static_cast<int>(blockIdx.x) + slab_iter) % (o.n_peer + 1);  // n_peer = the local C
      if (dst == o.n_peer) tma_store_2d_hint(o.tm, slab, e.col0 + cs, e.row_off + first_row, policy_evict_first());
      else tma_store_2d(&e.p.tm_c_peer[dst], slab, e.col0 + cs, e.row_off + first_row);
    }
    tma_store_commit();
  }
  ++slab_iter;
}

// Executed by the epilogue warps.  The warp owns TMEM lanes [32q, 32q+32), tile rows [32rq, 32rq+32) (rq == q except
// for pair tiles of 128 rows) and the tile columns [cbeg, cend).  tmem_acc = TMEM address of tile column 0 in those lanes.
//
// The accumulator goes back to the MMA warp the moment this warp's LAST tcgen05.ld of it has completed: what
// follows for that last slab (cast, wait for the TMA engine to have read the staging slab, st.shared, proxy fence,
// TMA store) works from registers while the tensor pipe is already on the next tile.  Before this the accumulator
// was held until the last store had been issued - 4.5k cycles per 128 x 256 accumulator, all of them exposed for a
// tile that fills TMEM (profiles/r2_notes.md).
template <int kOut, bool kAccInt>
__device__ __forceinline__ void epilogue_tile(const GemmKernelParams& p, const EpiOut& o, uint32_t tmem_acc, int q, int rq,
                                              int lane, int row0, int row_limit, int row_off, int col0, int cbeg, int cend,
                                              uint32_t stage_smem, int nslabs, int& slab_iter, uint32_t tempty_bar,
                                              bool tempty_remote) {
  constexpr int kOB = OutTraits<kOut>::kBytes;
  constexpr int kColsPerSlab = 128 / kOB;   // fp32 columns that fill one 128 B staging row
  constexpr int kPacketsPerLd = 32 * kOB / 16;  // 16 B packets produced by one x32 load
  constexpr int kValsPerPacket = 16 / kOB;
  const int grow = row0 + rq * 32 + lane;  // global row of this thread (rq: which 32 rows; q: which 32 TMEM lanes)
  const uint32_t lane_taddr = tmem_acc + (static_cast<uint32_t>(q * 32) << 16);
  const float alpha = o.alpha;
  const bool relu = o.relu != 0;
  // All tcgen05.ld of this warp have completed (wait::ld): hand the accumulator back.
  auto release_accumulator = [&]() {
    tc_fence_before();
    __syncwarp();
    if (lane == 0 && tempty_bar != 0) {  // 0: no later tile waits for this accumulator (see the call site)
      if (tempty_remote) mbar_arrive_cluster(tempty_bar);
      else mbar_arrive(tempty_bar);
    }
  };

  // A TMA store writes the warp's whole 32-row slab: use it only when all 32 rows belong to this tile's
  // row range (always true for an ordinary GEMM, where the tensor map clips at M; a grouped GEMM's last
  // row tile of a group falls back to predicated stores).
  const bool slab_rows_ok = p.n_groups == 0 || row0 + rq * 32 + 32 <= row_limit;
  if (o.tma && slab_rows_ok) {
    const EpiCtx e{p, o, lane_taddr, rq, lane, row0, row_limit, row_off, col0, stage_smem, nslabs};
    // slabs of this warp that hold at least one column of C (warp-uniform)
    int nsl = 0;
    for (int cs = cbeg; cs < cend && col0 + cs < o.n; cs += kColsPerSlab) ++nsl;
    if (nsl == 0) release_accumulator();
    constexpr int kLdsPerSlab = kColsPerSlab / 32;
    constexpr int kPair = kLdsPerSlab >= 2 ? 2 : 1;  // x32 loads in flight
    for (int i = 0; i < nsl; ++i) {
      const int cs = cbeg + i * kColsPerSlab;
      uint4 w[8];
#pragma unroll
      for (int l0 = 0; l0 < kLdsPerSlab; l0 += kPair) {
        uint32_t r[kPair][32];
        tmem_ld_32x32(lane_taddr + cs + l0 * 32, r[0]);
        if constexpr (kPair == 2) tmem_ld_32x32(lane_taddr + cs + (l0 + 1) * 32, r[1]);
        tmem_ld_wait();
        if (l0 + kPair >= kLdsPerSlab && i + 1 == nsl) release_accumulator();  // that was the last read
#pragma unroll
        for (int h = 0; h < kPair; ++h) {
          float v[32];
          epi_values<kOut, kAccInt>(r[h], v, alpha, relu, &o, grow, row_limit, col0 + cs + (l0 + h) * 32);
#pragma unroll
          for (int pk = 0; pk < kPacketsPerLd; ++pk) w[(l0 + h) * kPacketsPerLd + pk] = pack16<kOut>(&v[pk * kValsPerPacket]);
        }
      }
      epi_stage<kOut>(e, cs, w, slab_iter);
    }
  } else {
    const bool row_ok = grow < row_limit;
    const bool vec_ok = ((o.ldc * kOB) % 16 == 0) && ((reinterpret_cast<uintptr_t>(o.c) % 16) == 0);
    bool released = false;
    for (int cs = cbeg; cs < cend; cs += 32) {
      if (col0 + cs >= o.n) break;
      uint32_t r[32];
      tmem_ld_32x32(lane_taddr + cs, r);
      tmem_ld_wait();
      if (cs + 32 >= cend || col0 + cs + 32 >= o.n) {  // that was the last read of this accumulator
        release_accumulator();
        released = true;
      }
      float v[32];
      epi_values<kOut, kAccInt>(r, v, alpha, relu, &o, grow, row_limit, col0 + cs);
      if (row_ok) {
        const int gcol = col0 + cs;
        const long long base = static_cast<long long>(grow + row_off) * o.ldc + gcol;
        if (o.accumulate) {  // C += ...: every element is owned by exactly one thread of one CTA, no atomics needed
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (gcol + i < o.n) v[i] += load_out<kOut>(o.c, base + i);
        }
        for (int g = -1; g < o.n_peer; ++g) {  // g == -1: the local C; others: peer GPUs' copies
          void* cdst = g < 0 ? o.c : p.c_peer[g];
          if (vec_ok && gcol + 32 <= o.n && (gcol * kOB) % 16 == 0 && (reinterpret_cast<uintptr_t>(cdst) % 16) == 0) {
            uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<char*>(cdst) + base * kOB);
#pragma unroll
            for (int pk = 0; pk < kPacketsPerLd; ++pk) dst[pk] = pack16<kOut>(&v[pk * kValsPerPacket]);
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (gcol + i < o.n) store_scalar<kOut>(cdst, base + i, v[i]);
          }
        }
      }
      __syncwarp();  // tcgen05.ld is .sync.aligned: re-converge after the predicated stores
    }
    if (!released) release_accumulator();
  }
}

// ---------------------------------------------------------------- the kernel
// Control flow of every role is kept WARP-UNIFORM (role chosen from a shuffled warp index, loop
// bounds from kernel parameters) and only the issuing instructions are predicated on one elected
// lane.  tcgen05.mma / TMA take their operands from uniform registers; inside a lane-divergent
// region the compiler cannot prove uniformity and wraps every such instruction in an
// ELECT / R2UR.BROADCAST / BRA.U.ANY "waterfall" loop, which made the first version of this kernel
// issue-bound (~870 cycles per K-block instead of 512; profiles/r1_notes.md).
//
// kMc = 2 (with kCta = 2): clusters of FOUR CTAs = two pairs working on two neighbouring column tiles of the same
// row tile.  Both pairs need the same A rows, so each CTA fetches only HALF of its A box and TMA multicasts it to
// the CTA of the other pair that holds the same rows (A traffic per CTA halves: what cuBLAS's 2x2-cluster kernels
// do for small tiles, where L2->SM bandwidth, not the tensor pipe, is the bound).  A stage may be refilled only
// when BOTH pairs have consumed it: every pair leader's tcgen05.commit arrives on empty[s] of all four CTAs.
template <int kCta, int kKind, int kNSub, int kMc = 1>
__global__ void __launch_bounds__(kMaxThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ GemmKernelParams p) {
  static_assert(kMc == 1 || (kCta == 2 && kNSub == 1), "A multicast is implemented for pairs with one accumulator per tile");
  constexpr int kCluster = kCta * kMc;
  extern __shared__ uint8_t smem_raw[];
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const uint32_t cluster_rank = (kCluster > 1) ? cluster_ctarank() : 0u;
  const uint32_t cta_rank = cluster_rank & (kCta - 1);      // rank inside the pair
  const uint32_t mc_idx = cluster_rank / kCta;              // which of the cluster's column tiles this pair owns
  const uint32_t leader_rank = cluster_rank - cta_rank;     // cluster rank of this pair's leader
  const bool leader = cta_rank == 0;
  if (warp == 0) TLB_TRACE(0, 13);  // kernel entry

  // ---- shared memory carve-up (1024 B aligned for the 128B swizzle pattern)
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const bool half_m = kCta == 2 && kNSub == 1 && kMc == 1 && p.half_m != 0;
  const int rows_cta = half_m ? kBlockM / 2 : kBlockM;                       // tile rows of one CTA
  const uint32_t a_stage_bytes = static_cast<uint32_t>(p.a_stage_bytes);
  const int acc_w = half_m ? p.block_n / 2 : p.block_n;                      // TMEM columns of one accumulator
  const int kg = (kNSub == 2 || p.kgroup < 2) ? 1 : p.kgroup;  // K blocks (stages) per full / empty barrier pair
  const int groups = stages / kg;            // ring slots; stages beyond groups * kg stay unused
  constexpr int n_sub = kNSub;  // accumulators per tile (compile-time: a predicated-off tcgen05.mma is not free)
  const int acc_stages = p.acc_stages;
  const uint32_t b_sub_bytes = static_cast<uint32_t>(p.b_sub_bytes);
  const uint32_t b_stage_bytes = b_sub_bytes * n_sub;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_a + stages * a_stage_bytes;
  const uint32_t smem_epi = smem_b + stages * b_stage_bytes;
  const uint32_t smem_bar = smem_epi + p.epi_slabs * kEpiBufBytes;
  const uint32_t bar_full = smem_bar;                    // [stages]
  const uint32_t bar_empty = bar_full + 8 * stages;      // [stages]
  const uint32_t bar_tfull = bar_empty + 8 * stages;     // [2]
  const uint32_t bar_tempty = bar_tfull + 16;            // [2]
  const uint32_t bar_tmem = bar_tempty + 16;             // clusters: this CTA's tensor memory is allocated (1 arrive)
  const uint32_t bar_tmem_pair = bar_tmem + 8;           // clusters, pair leader: both CTAs' tensor memory is (kCta arrives)
  const uint32_t smem_tptr = bar_tmem_pair + 8;          // u32 TMEM base address
  // bytes that land in one stage: the A boxes actually issued (a short problem, M < 128, fetches only its own rows;
  // with A multicast each CTA issues half a box and receives the other half) plus the B rows, per CTA of the pair
  const uint32_t stage_tx_bytes = (static_cast<uint32_t>(p.a_boxes * p.a_box_bytes) * kMc + b_stage_bytes) * kCta;
  // Kernel parameters are cold in the constant cache at every launch, and the producers' way to their first TMA issue
  // reads a dozen of them one after another (a miss each time it enters a new line: ~1.2k cycles between the set-up
  // barrier and the first load).  One word of every line of the scalar block is read here, all at once, while the
  // set-up is still going on; the sum goes to a spare shared-memory word so that the loads cannot be dropped.
  if (lane == 0) {
    const uint32_t h = static_cast<uint32_t>(p.tail_begin) ^ static_cast<uint32_t>(p.n_groups) ^ static_cast<uint32_t>(p.m) ^
                       static_cast<uint32_t>(p.raster) ^ static_cast<uint32_t>(p.stages) ^ static_cast<uint32_t>(p.a_boxes) ^ p.idesc ^
                       p.kadv_a ^ static_cast<uint32_t>(p.out_kind) ^ static_cast<uint32_t>(p.l2_policy) ^
                       static_cast<uint32_t>(p.bias_per_row) ^ static_cast<uint32_t>(p.magic_tiles_mn) ^
                       static_cast<uint32_t>(p.magic_tail_tiles) ^ static_cast<uint32_t>(p.kgroup);
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(smem_tptr + 8 + 0 * warp), "r"(h) : "memory");
  }
  const uint32_t acc_cols = static_cast<uint32_t>(acc_w) * acc_stages;
  const uint32_t tmem_cols = acc_cols <= 32u ? 32u : acc_cols <= 64u ? 64u : acc_cols <= 128u ? 128u
                             : acc_cols <= 256u ? 256u : 512u;

  const int tiles_n_c = (p.tiles_n + kMc - 1) / kMc;  // column tiles per cluster step (kMc neighbours at a time)
  const int tiles_mn = p.tiles_m * tiles_n_c;
  const bool tail_mode = p.tail_tiles > 0;
  const int num_tiles = tail_mode ? p.tail_begin + p.tail_tiles * p.tail_split : tiles_mn * p.split_k;  // work items
  const int cluster_id = blockIdx.x / kCluster;
  const int num_clusters = gridDim.x / kCluster;
  const int nkb = p.num_k_blocks;

  // Work item -> (row tile, column tile); for a grouped GEMM also the group, the first row of the tile
  // and the end of the group's rows (the tile's row limit).
  // x / d through the host's reciprocal ceil(2^64 / d): the high half of x * magic is exact for every 32-bit x and
  // d >= 2 (magic 0 stands for d == 1).  No run-time division is left on the way to the first TMA issue.
  auto fdiv = [](int x, unsigned long long magic) -> int {
    return magic ? static_cast<int>(__umul64hi(static_cast<unsigned long long>(static_cast<unsigned>(x)), magic)) : x;
  };
  auto tile_coord = [&](int w, int& tm, int& tn, int& grp, int& row_base, int& row_limit) {
    int t;
    if (!tail_mode) t = w - fdiv(w, p.magic_tiles_mn) * tiles_mn;
    else t = w < p.tail_begin ? w : p.tail_begin + (w - p.tail_begin) - fdiv(w - p.tail_begin, p.magic_tail_tiles) * p.tail_tiles;
    const int group = p.raster * tiles_n_c;
    const int g = fdiv(t, p.magic_group);
    const int first = g * p.raster;
    const int rows = min(p.raster, p.tiles_m - first);
    const int within = t - g * group;
    const int wq = fdiv(within, rows == p.raster ? p.magic_raster : p.magic_last_rows);
    tm = first + within - wq * rows;
    tn = wq * kMc + (int)mc_idx;  // may equal tiles_n for the last odd tile: that pair computes zeros
    if (p.n_groups == 0) {
      grp = 0;
      row_base = tm * kCta * rows_cta;
      row_limit = p.m;
    } else {
      int lo = 0, hi = p.n_groups - 1;  // the last group whose first tile is <= tm (empty groups share a first tile)
      while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (p.group_tile[mid] <= tm) lo = mid; else hi = mid - 1;
      }
      const int gi = lo;
      grp = gi;
      row_base = p.group_row[gi] + (tm - p.group_tile[gi]) * kCta * rows_cta;
      row_limit = p.group_row[gi + 1];
    }
  };
  // K-block range [kb0, kb1) of work item w
  auto k_range = [&](int w, int& kb0, int& kb1) {
    if (tail_mode) {
      kb0 = w < p.tail_begin ? 0 : fdiv(w - p.tail_begin, p.magic_tail_tiles) * p.tail_kb;
      kb1 = w < p.tail_begin ? nkb : min(nkb, kb0 + p.tail_kb);
    } else {
      kb0 = fdiv(w, p.magic_tiles_mn) * p.kb_per_split;
      kb1 = min(nkb, kb0 + p.kb_per_split);
    }
  };

  // One pipeline stage: issue the A and B boxes of K block `kb` into stage `s`; their bytes are counted on `full`
  // (armed by the caller for all the K blocks of a group).
  // K-major operand: box = [rows x 128 B of K], coords {k, row}; MN-major: boxes = [K-block rows x 128 B of M/N],
  // coords {row + j*span, k}.  Called by one elected lane.
  const bool kmajor_single_box = p.a_boxes == 1 && p.b_boxes == 1 && !p.a_mn_major && !p.b_mn_major;
  auto issue_stage = [&](int s, uint32_t full, int kb, int row_a, int row_b, int kb_off, uint64_t pol) {
    const int kspan = p.k_elems_per_block;
    const int a_rows_per_box = p.a_box_bytes / kStageRowBytes;
    const int b_rows_per_box = p.b_box_bytes / kStageRowBytes;
#ifdef TLB200_TRACE_KB
    const bool tri = p.trace && blockIdx.x == 0 && kb < 64;  // CTA 0's producers, first 64 K blocks, step by step
#define TLB_TRI(j) do { if (tri) p.trace[(64 + kb) * 16 + (j)] = clock64(); } while (0)
#else
#define TLB_TRI(j) do { } while (0)
#endif
    TLB_TRI(1);
    const int k0 = kb * kspan;
    const uint32_t dst_a = smem_a + s * a_stage_bytes;
    const uint32_t dst_b = smem_b + s * b_stage_bytes;
    if (kmajor_single_box) {
      // Both operands K-contiguous, one box each (the benchmark layout): straight-line issue.  The general loops below
      // (several boxes per operand, either major-ness chosen per box) compile to ~150 dependent instructions per K
      // block - 840 cycles per issue, which capped every tile narrower than 256 columns at ~330 cycles per K block
      // (profiles/r2_notes.md); this path is ~20.
      if constexpr (kMc == 2) {
        tma_load_2d_pair_mc(dst_a + mc_idx * (kABytesPerStage / 2), &p.tm_a, full, k0, row_a + (int)mc_idx * (kBlockM / 2),
                            static_cast<uint16_t>(0x5u << cta_rank), pol);
      } else if constexpr (kCta == 2) tma_load_2d_pair(dst_a, &p.tm_a, full, k0, row_a, pol);
      else tma_load_2d(dst_a, &p.tm_a, full, k0, row_a, pol);
      TLB_TRI(2);
#pragma unroll
      for (int h = 0; h < n_sub; ++h) {
        if constexpr (kCta == 2) tma_load_2d_pair(dst_b + h * b_sub_bytes, &p.tm_b, full, k0, row_b + h * p.umma_n, pol);
        else tma_load_2d(dst_b + h * b_sub_bytes, &p.tm_b, full, k0, row_b + h * p.umma_n, pol);
      }
      TLB_TRI(3);
      return;
    }
    for (int j = 0; j < p.a_boxes; ++j) {
      const int c0 = p.a_mn_major ? row_a + j * kspan : k0;
      const int c1 = p.a_mn_major ? k0 : row_a + j * a_rows_per_box;
      if constexpr (kMc == 2) {
        // K-major A only (host guarantees it): this CTA fetches rows [64*mc_idx, +64) of its 128-row box
        // and multicasts them to itself and to the CTA of the other pair with the same pair rank.
        tma_load_2d_pair_mc(dst_a + mc_idx * (kABytesPerStage / 2), &p.tm_a, full, c0, c1 + (int)mc_idx * (kBlockM / 2),
                            static_cast<uint16_t>(0x5u << cta_rank), pol);
      } else if constexpr (kCta == 2) tma_load_2d_pair(dst_a + j * p.a_box_bytes, &p.tm_a, full, c0, c1, pol);
      else tma_load_2d(dst_a + j * p.a_box_bytes, &p.tm_a, full, c0, c1, pol);
    }
    // accumulator h of the tile covers tile columns [h*umma_n, (h+1)*umma_n)
#pragma unroll
    for (int h = 0; h < n_sub; ++h) {
      const int rb = row_b + h * p.umma_n;
      for (int j = 0; j < p.b_boxes; ++j) {
        const int c0 = p.b_mn_major ? rb + j * kspan : k0;
        const int c1 = p.b_mn_major ? k0 + kb_off : rb + j * b_rows_per_box;
        const uint32_t dst = dst_b + h * b_sub_bytes + j * p.b_box_bytes;
        if constexpr (kCta == 2) tma_load_2d_pair(dst, &p.tm_b, full, c0, c1, pol);
        else tma_load_2d(dst, &p.tm_b, full, c0, c1, pol);
      }
    }
  };
  // Operand coordinates of work item t for this CTA.
  auto operand_rows = [&](int t, int& row_a, int& row_b, int& kb_off) {
    int tm, tn, grp, row_base, row_limit;
    tile_coord(t, tm, tn, grp, row_base, row_limit);
    row_a = row_base + (int)cta_rank * rows_cta;
    // group g's B is block g of the 2-D view: rows [g*b_group_rows, ...) (K-major) or K offset (N-major)
    const int b_goff = grp * p.b_group_rows;
    row_b = tn * p.block_n + (int)cta_rank * p.b_rows_per_cta + (p.b_mn_major ? 0 : b_goff);
    kb_off = p.b_mn_major ? b_goff : 0;
  };
  // Single-CTA kernels: warp 0 (a TMA producer) does not wait at the block-wide set-up barrier.  It initialises the
  // stage barriers itself, ARRIVES on the set-up barrier and goes straight to its K blocks, while warp 2 is still
  // allocating tensor memory; the other warps synchronise as before (producers 2 and 3 join the round-robin after the
  // barrier).  Short problems are a set-up, one or two dozen K blocks and an epilogue, and ~2.2k of their ~12k cycles
  // went by between kernel entry and the first TMA issue (profiles/r2_notes.md).  With clusters the peer's loads signal
  // this CTA's barriers, which must not happen before the cluster-wide barrier, so pairs keep the ordinary order.
  const bool early_producer = kCluster == 1 && warp == 0;

  // The ring's barrier pairs are initialised one per lane (at most 32 groups; only `groups` of the `stages` slots are
  // used): a single thread looping over up to 64 mbarrier.init sat at the head of the producer's way to its first load.
  static_assert(kMaxStages <= 32, "one lane per stage-group barrier pair");
  if (warp == 0) {
    if constexpr (kCluster == 1) {
      if (lane < groups) {
        mbar_init(bar_full + 8 * lane, 1);
        mbar_init(bar_empty + 8 * lane, 1);
      }
      fence_mbar_init();
    }
    if (elect_one()) {
      tma_prefetch_desc(&p.tm_a);
      tma_prefetch_desc(&p.tm_b);
      if (p.epi_tma) {
        tma_prefetch_desc(&p.tm_c);
        for (int g = 0; g < p.n_peer; ++g) tma_prefetch_desc(&p.tm_c_peer[g]);
      }
    }
  } else if (warp == 1) {
    if constexpr (kCluster > 1) {
      if (lane < groups) {
        mbar_init(bar_full + 8 * lane, 1);    // producer's arrive.expect_tx (leader CTA only for a pair)
        mbar_init(bar_empty + 8 * lane, kMc);  // one tcgen05.commit per pair that reads the stages
      }
    }
    if (elect_one()) {
      for (int a = 0; a < 2; ++a) {
        mbar_init(bar_tfull + 8 * a, 1);                      // one tcgen05.commit
        mbar_init(bar_tempty + 8 * a, p.epi_warps * kCta);    // one arrive per epilogue warp (per CTA)
      }
      if constexpr (kCluster > 1) {
        mbar_init(bar_tmem, 1);
        mbar_init(bar_tmem_pair, kCta);
      }
    }
    fence_mbar_init();
  } else if (warp == 2) {
    if constexpr (kCluster == 1) tmem_alloc<kCta>(smem_tptr, tmem_cols);
  }
  __syncwarp();
  tc_fence_before();
  if constexpr (kCluster > 1) {
    // Clusters: the cluster-wide barrier only has to publish the mbarriers (the peer's loads and commits signal this CTA's),
    // so the tensor-memory allocation (~270 cycles and a second warp's worth of latency) is taken out of its way: the
    // producers start ~1k cycles earlier, the allocation runs next to their first loads and is announced through two
    // mbarriers of its own - to this CTA's MMA and epilogue warps, and to the pair's leader, whose MMAs write both CTAs'.
    // The arrive is relaxed: fence.mbarrier_init.release.cluster (above, by the initialising threads) is what publishes
    // the barriers; a release arrive adds MEMBAR.ALL.GPU + ERRBAR to every thread's way through the set-up.
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  } else {
    // named barrier 1 over the whole block: warp 0 only arrives (it needs neither the TMEM address nor the accumulator
    // barriers), everybody else waits for all arrivals
    if (early_producer) asm volatile("barrier.cta.arrive 1, %0;" ::"r"(static_cast<uint32_t>(blockDim.x)) : "memory");
    else asm volatile("barrier.cta.sync 1, %0;" ::"r"(static_cast<uint32_t>(blockDim.x)) : "memory");
  }
  tc_fence_after();
  uint32_t tmem_base = 0;
  if constexpr (kCluster > 1) {
    if (warp == 2) {
      tmem_alloc<kCta>(smem_tptr, tmem_cols);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(bar_tmem);
        mbar_arrive_cluster(mapa(bar_tmem_pair, leader_rank));
      }
      __syncwarp();
    }
    if (warp == 1 || warp == 2 || warp >= kEpiWarp0) {  // producers 0 and 3 need no tensor memory
      if (warp == 1 && leader) mbar_wait(bar_tmem_pair, 0, 5);  // the MMAs of a pair write both CTAs' accumulators
      mbar_wait(bar_tmem, 0, 5);
      tc_fence_after();
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_tptr) : "memory");
    }
  } else {
    if (!early_producer) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_tptr) : "memory");
  }
  // Programmatic dependent launch: everything above touched only this CTA's shared / tensor memory, so it may
  // overlap the tail of the previous kernel in the stream; the next kernel may start its own set-up now.  The warps
  // that read or write global memory wait for the previous grid first (producers before the first TMA load,
  // epilogue warps before the first store).
  griddep_launch_dependents();
  if (warp == 0) TLB_TRACE(0, 14);  // set-up done

  if (warp != 1 && warp < kEpiWarp0) {
    // ===================================================== TMA producers (one elected lane of each issues)
    // Three producer warps (0, 2, 3) take the K blocks round-robin, each filling whole stages.  One warp's issue
    // path (try_wait on `empty`, arrive.expect_tx, two TMA issues) costs ~280 ns per K block, which starved every
    // tile narrower than 256 columns (282 ns per block whatever the tile, stage count or cluster shape;
    // profiles/r1_notes.md); three interleaved streams bring the issue cost under the MMA time of a 64-wide tile.
    // Parity waits only tell adjacent phases apart: a producer that skips P groups between its visits must not
    // be able to see a group's `empty` barrier two phases behind, which holds for P <= groups.
    // A barrier pair covers `kg` consecutive stages (GemmKernelParams::kgroup): one wait, one expect_tx and, on the
    // other side, one wait and one commit per kg K blocks - the hand-shake, not the loads or the MMAs, is what a
    // narrow tile's K block costs (see the MMA issuer below).
#ifdef TLB200_TRACE
    const int kProducers = (p.debug_mode & 4) ? 1 : (p.debug_mode & 8) ? 2 : groups < 3 ? groups : 3;
#else
    const int kProducers = groups < 3 ? groups : 3;
#endif
    const int pidx = warp == 0 ? 0 : warp - 1;  // warps 0, 2, 3 -> 0, 1, 2 (a warp with pidx >= kProducers idles)
    int turn = 0;                               // whose K block the next one is
    const bool elected = elect_one();
    // Operands are re-read by other tiles of the raster group: keep them in L2 (evict_last measured +1.5 % bf16,
    // +3 % fp8 sustained over evict_normal; profiles/r1_notes.md).  l2_policy 1 restores evict_normal.
    if (warp == 0) TLB_TRACE(200, 0);
    const uint64_t pol = p.l2_policy == 1 ? policy_evict_normal() : policy_evict_last();
    int gs = 0;  // ring slot: a GROUP of kg consecutive stages behind one full / empty barrier pair
    uint32_t ph = 0;
    if (warp == 0) TLB_TRACE(200, 1);
    griddep_wait();
    if (warp == 0) TLB_TRACE(200, 2);
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int row_a, row_b, kb_off;
      operand_rows(t, row_a, row_b, kb_off);
      if (warp == 0 && t == cluster_id) TLB_TRACE(200, 3);
      int kb0, kb1;
      k_range(t, kb0, kb1);
      if (warp == 0) TLB_TRACE((t - cluster_id) / num_clusters, 11);
      for (int kb = kb0; kb < kb1; kb += kg) {
        const bool mine = turn == pidx;
        if (++turn == kProducers) turn = 0;
#ifdef TLB200_TRACE_KB
        if (mine && elected && kb < 64 && p.trace && blockIdx.x == 0 && t == cluster_id) p.trace[(64 + kb) * 16 + 4] = clock64();
#endif
        if (mine) mbar_wait(bar_empty + 8 * gs, ph ^ 1u, 0);
        if (mine && elected) {
          const int nk = kb1 - kb < kg ? kb1 - kb : kg;
          const uint32_t full = bar_full + 8 * gs;
#ifdef TLB200_TRACE_KB
          if (kb < 64 && p.trace && blockIdx.x == 0 && t == cluster_id) p.trace[(64 + kb) * 16 + 0] = clock64();
#endif
#ifdef TLB200_TRACE
          if (p.debug_mode & 2) { if (leader) mbar_arrive(full); } else
#endif
          {
            if (leader) mbar_arrive_expect_tx(full, stage_tx_bytes * nk);
            for (int j = 0; j < nk; ++j) issue_stage(gs * kg + j, full, kb + j, row_a, row_b, kb_off, pol);
          }
        }
        __syncwarp();
#ifdef TLB200_TRACE
        if (warp == 0 && t == cluster_id && kb - kb0 == 6 * kg) TLB_TRACE(0, 12);  // warp 0 has issued its third group
#endif
        if (++gs == groups) { gs = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer (leader CTA; one elected lane)
    if (leader) {
      const bool elected = elect_one();
      const uint32_t idesc = p.idesc;
      const uint64_t hi_a = static_cast<uint64_t>(p.desc_hi_a) << 32 | p.desc_lbo_a;
      const uint64_t hi_b = static_cast<uint64_t>(p.desc_hi_b) << 32 | p.desc_lbo_b;
      const uint32_t kadv_a = p.kadv_a, kadv_b = p.kadv_b;
      const uint32_t a_lo0 = (smem_a & 0x3FFFFu) >> 4, b_lo0 = (smem_b & 0x3FFFFu) >> 4;
      const uint32_t a_step = a_stage_bytes >> 4, b_step = b_stage_bytes >> 4, b_sub = b_sub_bytes >> 4;
      const uint32_t umma_n = p.umma_n;
      int s = 0;
      uint32_t ph = 0;
      int as = 0;
      uint32_t aph = 0;
      if constexpr (kNSub == 1) {
        // Narrow tiles are bound by THIS loop, not by the tensor pipe or the loads.  tools/ubench/umma_pace.cu: a
        // 256x64 MMA runs every 43 cycles when issued back to back (172 per K block), but the pipe holds next to no
        // queue - whatever the issuing thread does between two K blocks is added to them, and a barrier wait alone is
        // 100-200 cycles (the loop that waited, then issued, ran at ~320 per K block with operands already resident).
        // So: ONE elected thread runs the whole loop (ptxas then keeps descriptors and barrier addresses in uniform
        // registers: no R2UR in front of every MMA), the barrier of the NEXT K block is probed before this block's
        // MMAs are issued and only waited on if that probe failed, the stage state advances by additions, and the four
        // MMAs of a K block are one asm block.  Even so one pass (probe, commit, bookkeeping, with the producers'
        // hand-shake on the other side) is ~250-320 cycles (tools/ubench/umma_ring.cu), above the 172 of a 64-wide K
        // block, so narrow tiles put `kgroup` K blocks behind each barrier pair: one pass then issues 4 * kgroup MMAs.
        static_assert(kStageRowBytes / kUmmaKBytes == 4, "umma_ss_x4 issues four MMAs per K block");
        const uint16_t empty_mask = kMc == 2 ? 0xF : 0x3;
        if (elected) {
          const uint32_t a_w0 = static_cast<uint32_t>(hi_a) | a_lo0, b_w0 = static_cast<uint32_t>(hi_b) | b_lo0;
          const uint32_t a_hi = static_cast<uint32_t>(hi_a >> 32), b_hi = static_cast<uint32_t>(hi_b >> 32);
          const uint32_t a_gstep = a_step * kg, b_gstep = b_step * kg;
          uint32_t a_w = a_w0, b_w = b_w0, bar_e = bar_empty, bar_f = bar_full;  // of the coming group
          int gs = 0;
          bool ready = false;  // the probe of the coming group's barrier succeeded
          for (int t = cluster_id; t < num_tiles; t += num_clusters) {
            mbar_wait(bar_tempty + 8 * as, aph ^ 1u, 1);  // epilogue has drained this accumulator
            tc_fence_after();
            TLB_TRACE((t - cluster_id) / num_clusters, 0);
            const uint32_t d_tmem = tmem_base + as * acc_w;
            int kb0, kb1;
            k_range(t, kb0, kb1);
            const bool more_tiles = t + num_clusters < num_tiles;
            for (int kb = kb0; kb < kb1; kb += kg) {
              if (!ready) mbar_wait(bar_f, ph, 2);  // TMA bytes of this group's stages have landed
              tc_fence_after();
              if (kb == kb0) TLB_TRACE((t - cluster_id) / num_clusters, 1);
#ifdef TLB200_TRACE_KB
              if (p.trace && blockIdx.x == 0 && t == cluster_id && kb < 64) p.trace[(64 + kb) * 16 + 8] = clock64();  // this group's operands seen
#endif
              uint32_t a_cur = a_w, b_cur = b_w;
              const uint32_t e_cur = bar_e;
              a_w += a_gstep; b_w += b_gstep; bar_e += 8; bar_f += 8;
              if (++gs == groups) { gs = 0; ph ^= 1u; a_w = a_w0; b_w = b_w0; bar_e = bar_empty; bar_f = bar_full; }
              // non-blocking probe (try_wait may suspend the thread, which would hold back the MMAs below)
              ready = (kb + kg < kb1 || more_tiles) && mbar_test_wait(bar_f, ph);
              const int nk = kb1 - kb < kg ? kb1 - kb : kg;
#ifdef TLB200_TRACE
              if (!(p.debug_mode & 1))
#endif
              for (int j = 0; j < nk; ++j) {
                umma_ss_x4<kCta, kKind>(d_tmem, a_cur, a_hi, b_cur, b_hi, kadv_a, kadv_b, idesc, (kb == kb0 && j == 0) ? 0u : 1u);
                a_cur += a_step; b_cur += b_step;
              }
              umma_commit<kCta>(e_cur, empty_mask);  // the group's stages are reusable once these MMAs retire
            }
            umma_commit<kCta>(bar_tfull + 8 * as, static_cast<uint16_t>(0x3u << leader_rank));  // accumulator complete -> epilogue
            TLB_TRACE((t - cluster_id) / num_clusters, 2);
            if (++as == acc_stages) { as = 0; aph ^= 1u; }
          }
        }
        __syncwarp();
      } else {
        // Two accumulators per tile sharing every A stage (they fill TMEM, so there is no second buffer: what the
        // epilogue cannot hide is exposed).  tmem_full/empty[h] belong to accumulator h.  The K blocks are issued
        // interleaved (both accumulators consume a stage, then it is released) except for the LAST `tail` blocks, all
        // of which are resident at once: there accumulator 0 runs through all of them first and is published, then
        // accumulator 1 - so the epilogue drains accumulator 0 while the tensor pipe still works on accumulator 1
        // (2048 cycles for four stages) instead of after it.
        auto mma_half = [&](int stage, uint32_t d_tmem, uint32_t b_extra, bool first_kb) {
          const uint32_t a_lo = a_lo0 + stage * a_step;
          const uint32_t b_lo = b_lo0 + stage * b_step + b_extra;
#pragma unroll
          for (int k = 0; k < kStageRowBytes / kUmmaKBytes; ++k) {
            const uint64_t da = hi_a | ((a_lo + k * kadv_a) & 0x3FFFu);
            const uint64_t db = hi_b | ((b_lo + k * kadv_b) & 0x3FFFu);
            umma_ss<kCta, kKind>(d_tmem, da, db, idesc, (first_kb && k == 0) ? 0u : 1u);
          }
        };
        for (int t = cluster_id; t < num_tiles; t += num_clusters) {
          mbar_wait(bar_tempty, aph ^ 1u, 1);
          mbar_wait(bar_tempty + 8, aph ^ 1u, 4);
          tc_fence_after();
          TLB_TRACE((t - cluster_id) / num_clusters, 0);
          int kb0, kb1;
          k_range(t, kb0, kb1);
          const int tail = min(stages, kb1 - kb0);
          for (int kb = kb0; kb < kb1 - tail; ++kb) {
            mbar_wait(bar_full + 8 * s, ph, 2);
            tc_fence_after();
            if (kb == kb0) TLB_TRACE((t - cluster_id) / num_clusters, 1);
            if (elected) {
              const uint32_t a_lo = a_lo0 + s * a_step;
              const uint32_t b_lo = b_lo0 + s * b_step;
#pragma unroll
              for (int k = 0; k < kStageRowBytes / kUmmaKBytes; ++k) {
                const uint64_t da = hi_a | ((a_lo + k * kadv_a) & 0x3FFFu);
                const uint32_t acc = (kb != kb0 || k != 0) ? 1u : 0u;
                umma_ss<kCta, kKind>(tmem_base, da, hi_b | ((b_lo + k * kadv_b) & 0x3FFFu), idesc, acc);
                umma_ss<kCta, kKind>(tmem_base + umma_n, da, hi_b | ((b_lo + b_sub + k * kadv_b) & 0x3FFFu), idesc, acc);
              }
              umma_commit<kCta>(bar_empty + 8 * s);
            }
            __syncwarp();
            if (++s == stages) { s = 0; ph ^= 1u; }
          }
          // tail, accumulator 0: every remaining stage (no stage is released yet)
          {
            int s1 = s;
            uint32_t ph1 = ph;
            for (int kb = kb1 - tail; kb < kb1; ++kb) {
              mbar_wait(bar_full + 8 * s1, ph1, 2);
              tc_fence_after();
              if (kb == kb0) TLB_TRACE((t - cluster_id) / num_clusters, 1);
              if (elected) mma_half(s1, tmem_base, 0u, kb == kb0);
              __syncwarp();
              if (++s1 == stages) { s1 = 0; ph1 ^= 1u; }
            }
            if (elected) umma_commit<kCta>(bar_tfull);
            __syncwarp();
          }
          // tail, accumulator 1: the same stages again (their barriers have been observed), releasing each
          for (int kb = kb1 - tail; kb < kb1; ++kb) {
            if (elected) {
              mma_half(s, tmem_base + umma_n, b_sub, kb == kb0);
              umma_commit<kCta>(bar_empty + 8 * s);
              if (kb + 1 == kb1) umma_commit<kCta>(bar_tfull + 8);
            }
            __syncwarp();
            if (++s == stages) { s = 0; ph ^= 1u; }
          }
          TLB_TRACE((t - cluster_id) / num_clusters, 2);
          aph ^= 1u;
        }
      }
    }
  } else if (warp >= kEpiWarp0) {
    // ===================================================== epilogue warps
    const int ew = warp - kEpiWarp0;
    const int q = ew & 3;            // == warp % 4: the TMEM lane quarter this warp may read
    const int chalf = ew >> 2;       // with 8 epilogue warps: which half of the tile's columns
    const int nslabs = p.epi_slabs / p.epi_warps;  // 2 or 1 staging slabs per warp
    const uint32_t my_stage = smem_epi + ew * nslabs * kEpiBufBytes;
    const uint32_t tempty_leader = (kCta == 2) ? mapa(bar_tempty, leader_rank) : bar_tempty;
    // Column range of this warp inside ONE accumulator (all warps drain accumulator 0, then 1).
    const int acc_width = (kNSub == 2) ? p.umma_n : p.block_n;
    // Pair tiles of 128 rows (half_m): this warp's lane quarter q holds tile rows [32 * (q & 1), +32) of the column half
    // q >> 1, at TMEM columns [0, block_n / 2) - so the warp's columns start at col_half and its TMEM address is
    // shifted back by as much (the epilogue adds the tile column to it).
    const int rq = half_m ? (q & 1) : q;                           // which 32 rows of the CTA's part of the tile
    const int col_half = half_m ? (q >> 1) * (p.block_n / 2) : 0;  // first tile column this warp can reach
    const int ccols = (half_m ? acc_width / 2 : acc_width) / (p.epi_warps >> 2);
    int as = 0;
    uint32_t aph = 0;
    int slab_iter = 0;
    griddep_wait();
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int tm, tn, grp, row_base, row_limit;
      tile_coord(t, tm, tn, grp, row_base, row_limit);
      const bool tail_item = tail_mode && t >= p.tail_begin;
      const bool rows_in_range = row_base + (int)cta_rank * rows_cta < row_limit;
      // a tail slice writes tile (t - tail_begin) of the workspace matrix, whole (the reduction clips to M x N)
      const int row0 = (tail_item ? (t - p.tail_begin) * kCta * rows_cta : row_base) + (int)cta_rank * rows_cta;
      const int col0 = tail_item ? 0 : tn * p.block_n;
      const int row_off = (tail_mode ? 0 : fdiv(t, p.magic_tiles_mn)) * p.split_row_pitch;
      if (tail_item) row_limit = 0x7fffffff;
      // partial sums of a tail slice go to the workspace raw; everything else is the final C
      EpiOut o{tail_item ? &p.tm_ws : &p.tm_c, tail_item ? p.ws : p.c, tail_item ? (long long)p.block_n : p.ldc,
                     tail_item ? p.block_n : p.n, tail_item ? 1.0f : p.alpha, tail_item ? 0 : p.relu,
                     tail_item ? 0 : p.accumulate, tail_item ? 1 : p.epi_tma, tail_item ? 0 : p.n_peer,
                     tail_item ? 0 : p.epi_lsu, (!tail_item && p.l2_policy != 1) ? 1 : 0,
                     tail_item ? nullptr : p.row_scale, tail_item ? nullptr : p.col_scale, tail_item ? nullptr : p.bias,
                     p.bias_per_row, -1};
      const int out_kind = tail_item ? static_cast<int>(kOutF32) : p.out_kind;
#pragma unroll
      for (int h = 0; h < kNSub; ++h) {
        const int bi = (kNSub == 2) ? h : as;  // barrier / accumulator slot
        mbar_wait(bar_tfull + 8 * bi, aph, 3);
        tc_fence_after();
        if ((ew & 3) == 0) TLB_TRACE((t - cluster_id) / num_clusters, 3 + 2 * h + (ew >> 2) * 4);
        const uint32_t acc = ((kNSub == 2) ? tmem_base : tmem_base + as * acc_w) - col_half;  // column 0 of the tile
#ifdef TLB200_TRACE
        o.trace_seq = (ew == 0 && h == 0) ? (t - cluster_id) / num_clusters : -1;
#endif
        const int cbeg = h * acc_width + col_half + chalf * ccols, cend = cbeg + ccols;
        // Nothing to read (a fully out-of-range half-tile, or an output type this instantiation cannot produce - the
        // host rejects those): hand the accumulator straight back.  Otherwise epilogue_tile does so itself, as soon as
        // its last tcgen05.ld has completed.
        // The hand-back of an accumulator that no later tile of this cluster will use is left out: nobody would wait
        // for it, and as an arrive on the LEADER's barrier from the other CTA of a pair it would be the one remote
        // operation of this kernel that is never observed by its target - which lets the teardown below do without a
        // cluster-scope release (it compiles to MEMBAR.ALL.GPU and waits for every thread's stores to be acknowledged).
        const bool acc_reused = t + (kNSub == 2 ? 1 : acc_stages) * num_clusters < num_tiles;
        auto release_unread = [&]() {
          tc_fence_before();
          __syncwarp();
          if (lane == 0 && acc_reused) {
            if constexpr (kCta == 2) mbar_arrive_cluster(tempty_leader + 8 * bi);
            else mbar_arrive(bar_tempty + 8 * bi);
          }
        };
        if (rows_in_range) {  // warp-uniform
          constexpr bool kAccInt = kKind == kKindI8;  // int8 operands accumulate in int32
#define TLB_EPI(K) epilogue_tile<K, kAccInt>(p, o, acc, q, rq, lane, row0, row_limit, row_off, col0, cbeg, cend, my_stage, nslabs, slab_iter, \
                                             acc_reused ? (kCta == 2 ? tempty_leader : bar_tempty) + 8 * bi : 0u, kCta == 2)
          switch (out_kind) {
            case kOutBF16: TLB_EPI(kOutBF16); break;
            case kOutF16: TLB_EPI(kOutF16); break;
            case kOutF32: TLB_EPI(kOutF32); break;
            case kOutI32:
              if constexpr (kAccInt) TLB_EPI(kOutI32); else release_unread();  // int32 output exists for int8 operands only
              break;
            case kOutE4M3:
              if constexpr (!kAccInt) TLB_EPI(kOutE4M3); else release_unread();
              break;
            default:
              if constexpr (!kAccInt) TLB_EPI(kOutE5M2); else release_unread();
              break;
          }
#undef TLB_EPI
        } else {
          release_unread();
        }
        if ((ew & 3) == 0) TLB_TRACE((t - cluster_id) / num_clusters, 4 + 2 * h + (ew >> 2) * 4);
      }
      if constexpr (kNSub == 2) { aph ^= 1u; } else if (++as == acc_stages) { as = 0; aph ^= 1u; }
    }
    // The staging slabs must have been read before the CTA's shared memory goes away; the writes themselves are
    // complete when the grid is.  Stores into peer GPUs' memory are waited for in full.
    if (p.epi_tma && (!p.epi_lsu || tail_mode) && lane <= p.n_peer) {
      if (p.n_peer > 0) tma_store_wait_all<0>(); else tma_store_wait_read<0>();
    }
  }

  // ---- teardown: nobody may still be using TMEM / peer smem when it is released
  __syncwarp();  // re-converge the single-lane regions before the .aligned barriers
  if (warp == kEpiWarp0) TLB_TRACE(0, 15);  // first epilogue warp done with every tile
  tc_fence_before();
  // Execution order is all this barrier has to give: every remote operation on a CTA's shared memory (the pair's TMA
  // loads crediting the leader's `full` barriers, the MMA thread's multicast commits, the hand-backs of accumulators
  // that a later tile uses) has been OBSERVED by a thread of that CTA through an mbarrier wait before that thread can
  // get here, and the staging slabs have been read by the stores (wait_group.read above).  A release arrive would
  // add MEMBAR.ALL.GPU + ERRBAR per thread - the CTA then holds its SM until its global stores are acknowledged,
  // ~0.6 us that the next kernel's set-up (programmatic dependent launch) can use.
  if constexpr (kCluster > 1) {
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  } else __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc<kCta>(tmem_base, tmem_cols);
}

}  // namespace tlb200
