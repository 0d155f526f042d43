// gemm_chain_sm100.cuh - two contractions back to back with the intermediate kept in TENSOR memory:
//
//     P[M, N1] = cast_ab(alpha1 * A[M, K] . B1[N1, K]^T)        (fp32 accumulate, one rounding to the operand type)
//     D[M, N2] = cast_c (alpha2 * P[M, N1] . B2[N2, N1]^T)      (fp32 accumulate, one rounding)
//
// The second contraction reads its A operand from TMEM (the "TS" form of tcgen05.mma: A from tensor memory, B from a
// shared-memory descriptor).  This is the operand mode the reference declares but never lowers: `tl::tcgen05mma_ts`
// exists as a PTX wrapper (src/tl_templates/cuda/instruction/tcgen05mma.h:26-133) while `tcgen5mma_gemm_ts` is a TODO
// (src/tl_templates/cuda/gemm_sm100.h:419-424).  It is the inner step of attention (S = Q K^T,
// O = P V) and of low-rank products x A^T B^T: the [128 x N1] intermediate never goes to shared or global memory.
//
// One CTA owns a 128-row tile: it computes S once (N1 <= 256 columns of fp32 in TMEM), its epilogue warps read S,
// scale, round to the operand type and write the packed result back to TMEM with tcgen05.st (two K-consecutive
// 16-bit elements per 32-bit column, row = lane), then the MMA warp walks the column tiles of D (128 wide), each a
// chain of tcgen05.mma TS instructions over N1, while the epilogue warps store the previous one.
//
//   TMEM columns, N1 > 128:   [0, 256) S (fp32)   [256, 384) P (packed 16-bit, N1 / 2 columns)   [384, 512) O
//                 N1 <= 128:  [0, 128) S          [128, 192) P        [256, 384) O buffer 0   [384, 512) O buffer 1
//   (O is one 128-column accumulator per D tile; with two buffers the MMAs of tile j+1 run while tile j is stored)
//   warp 0: TMA producer    warp 1: MMA issuer    warp 2: TMEM alloc    warps 4-7: conversion + epilogue
#pragma once
#include "gemm_sm100.cuh"

namespace tlb200 {

constexpr int kChainThreads = 256;
constexpr int kChainStageBytes = kABytesPerStage + 256 * kStageRowBytes;  // A box + up to 256 rows of B1 / 128 of B2
constexpr int kChainOWidth = 128;                                          // columns of one D tile

struct alignas(64) ChainParams {
  CUtensorMap tm_a;   // A  [M,  K ] boxes of 128 rows x 128 B
  CUtensorMap tm_b1;  // B1 [N1, K ] boxes of n1 rows x 128 B
  CUtensorMap tm_b2;  // B2 [N2, N1] boxes of 128 rows x 128 B
  void* c;
  long long ldc;
  int m, n1, n2;
  int kb1;        // K blocks of the first contraction (128 B of K each)
  int kb2;        // K blocks of the second = n1 / 64
  int tiles_m, tiles_n2;
  int stages;
  uint32_t idesc1, idesc2;
  uint32_t desc_hi;
  int out_kind;   // kOutF16 / kOutBF16 / kOutF32
  int mid_bf16;   // the intermediate is rounded to bfloat16 (else float16)
  int relu_mid;   // relu on the intermediate before it is rounded
  float alpha1, alpha2;
};

// D[tmem] (+)= A[tmem] * B[smem desc]   (kind::f16, one CTA)
__device__ __forceinline__ void umma_ts_f16(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b),
               "r"(idesc), "r"(accumulate)
               : "memory");
}

// registers -> TMEM: thread l of the warp writes 16 consecutive 32-bit columns of lane (32 * (warp % 4) + l)
__device__ __forceinline__ void tmem_st_32x16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]), "r"(r[10]),
      "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// (a template only so that the definition can live in this header: it is instantiated in kernels_chain.cu alone)
template <int kMidTypeAgnostic = 0>
__global__ void __launch_bounds__(kChainThreads, 1) gemm_chain_kernel(const __grid_constant__ ChainParams p) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int stages = p.stages;
  const uint32_t smem_a = smem_base;                                    // [stages] A boxes (first contraction only)
  const uint32_t smem_b = smem_a + stages * kABytesPerStage;            // [stages] B1 or B2 boxes
  const uint32_t smem_bar = smem_b + stages * (256 * kStageRowBytes);
  const uint32_t bar_full = smem_bar;                  // [stages]
  const uint32_t bar_empty = bar_full + 8 * stages;    // [stages]
  const uint32_t bar_sfull = bar_empty + 8 * stages;   // S complete (MMA -> converters)
  const uint32_t bar_pfull = bar_sfull + 8;            // P written (4 converter warps -> MMA)
  const uint32_t bar_ofull = bar_pfull + 8;            // [2] O tile complete (MMA -> epilogue)
  const uint32_t bar_oempty = bar_ofull + 16;          // [2] O tile drained (4 epilogue warps -> MMA)
  const uint32_t smem_tptr = bar_oempty + 16;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tm_a);
      tma_prefetch_desc(&p.tm_b1);
      tma_prefetch_desc(&p.tm_b2);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(bar_full + 8 * s, 1);
        mbar_init(bar_empty + 8 * s, 1);
      }
      mbar_init(bar_sfull, 1);
      mbar_init(bar_pfull, 4);
      for (int o = 0; o < 2; ++o) {
        mbar_init(bar_ofull + 8 * o, 1);
        mbar_init(bar_oempty + 8 * o, 4);
      }
      fence_mbar_init();
    }
  } else if (warp == 2) {
    tmem_alloc<1>(smem_tptr, 512);
  }
  __syncwarp();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_tptr) : "memory");
  const int o_bufs = p.n1 <= 128 ? 2 : 1;
  const uint32_t tmem_s = tmem_base, tmem_p = tmem_base + (o_bufs == 2 ? 128 : 256);
  const uint32_t tmem_o0 = tmem_base + (o_bufs == 2 ? 256 : 384);  // buffer b at tmem_o0 + 128 * b
  const uint32_t b1_bytes = static_cast<uint32_t>(p.n1) * kStageRowBytes;
  const uint32_t b2_bytes = kChainOWidth * kStageRowBytes;

  if (warp == 0) {
    // ===================================================== TMA producer
    const bool elected = elect_one();
    const uint64_t pol = policy_evict_last();
    int s = 0;
    uint32_t ph = 0;
    for (int tm = blockIdx.x; tm < p.tiles_m; tm += gridDim.x) {
      for (int kb = 0; kb < p.kb1; ++kb) {  // first contraction: A and B1 K blocks
        mbar_wait(bar_empty + 8 * s, ph ^ 1u, 0);
        if (elected) {
          const uint32_t full = bar_full + 8 * s;
          mbar_arrive_expect_tx(full, kABytesPerStage + b1_bytes);
          tma_load_2d(smem_a + s * kABytesPerStage, &p.tm_a, full, kb * 64, tm * kBlockM, pol);
          tma_load_2d(smem_b + s * (256 * kStageRowBytes), &p.tm_b1, full, kb * 64, 0, pol);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
      for (int tn = 0; tn < p.tiles_n2; ++tn) {  // second contraction: B2 only (A is P, in TMEM)
        for (int kk = 0; kk < p.kb2; ++kk) {
          mbar_wait(bar_empty + 8 * s, ph ^ 1u, 0);
          if (elected) {
            const uint32_t full = bar_full + 8 * s;
            mbar_arrive_expect_tx(full, b2_bytes);
            tma_load_2d(smem_b + s * (256 * kStageRowBytes), &p.tm_b2, full, kk * 64, tn * kChainOWidth, pol);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer
    const bool elected = elect_one();
    const uint64_t hi = static_cast<uint64_t>(p.desc_hi) << 32 | (1u << 16);
    const uint32_t a_lo0 = (smem_a & 0x3FFFFu) >> 4, b_lo0 = (smem_b & 0x3FFFFu) >> 4;
    int s = 0;
    uint32_t ph = 0, tile_ph = 0, o_ph = 0;
    int ob = 0;
    for (int tm = blockIdx.x; tm < p.tiles_m; tm += gridDim.x) {
      // S = A . B1^T: both operands from shared memory.  (The previous row tile's last O tile completed before its
      // epilogue released it, so every TS instruction that read P or S's columns has retired.)
      for (int kb = 0; kb < p.kb1; ++kb) {
        mbar_wait(bar_full + 8 * s, ph, 2);
        tc_fence_after();
        if (elected) {
          const uint32_t a_lo = a_lo0 + s * (kABytesPerStage >> 4);
          const uint32_t b_lo = b_lo0 + s * ((256 * kStageRowBytes) >> 4);
#pragma unroll
          for (int k = 0; k < kStageRowBytes / kUmmaKBytes; ++k)
            umma_ss<1, kKindF16>(tmem_s, hi | ((a_lo + k * (kUmmaKBytes >> 4)) & 0x3FFFu), hi | ((b_lo + k * (kUmmaKBytes >> 4)) & 0x3FFFu),
                                 p.idesc1, (kb == 0 && k == 0) ? 0u : 1u);
          umma_commit<1>(bar_empty + 8 * s);
          if (kb + 1 == p.kb1) umma_commit<1>(bar_sfull);
        }
        __syncwarp();
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
      mbar_wait(bar_pfull, tile_ph, 5);  // the converters have written P
      tc_fence_after();
      for (int tn = 0; tn < p.tiles_n2; ++tn) {
        mbar_wait(bar_oempty + 8 * ob, o_ph ^ 1u, 1);
        tc_fence_after();
        const uint32_t tmem_o = tmem_o0 + 128 * ob;
        for (int kk = 0; kk < p.kb2; ++kk) {
          mbar_wait(bar_full + 8 * s, ph, 2);
          tc_fence_after();
          if (elected) {
            const uint32_t b_lo = b_lo0 + s * ((256 * kStageRowBytes) >> 4);
#pragma unroll
            for (int k = 0; k < 4; ++k)  // 16 K elements of P = 8 packed columns per instruction
              umma_ts_f16(tmem_o, tmem_p + kk * 32 + k * 8, hi | ((b_lo + k * (kUmmaKBytes >> 4)) & 0x3FFFu), p.idesc2,
                          (kk == 0 && k == 0) ? 0u : 1u);
            umma_commit<1>(bar_empty + 8 * s);
            if (kk + 1 == p.kb2) umma_commit<1>(bar_ofull + 8 * ob);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
        if (++ob == o_bufs) { ob = 0; o_ph ^= 1u; }
      }
      tile_ph ^= 1u;
    }
  } else if (warp >= kEpiWarp0) {
    // ===================================================== conversion of the intermediate, then the epilogue of D
    const int q = warp & 3;
    const uint32_t lane_off = static_cast<uint32_t>(q * 32) << 16;
    uint32_t tile_ph = 0, o_ph = 0;
    int ob = 0;
    const int ob_bytes = p.out_kind == kOutF32 ? 4 : 2;
    const bool vec_ok = (reinterpret_cast<uintptr_t>(p.c) % 16 == 0) && ((p.ldc * ob_bytes) % 16 == 0);
    for (int tm = blockIdx.x; tm < p.tiles_m; tm += gridDim.x) {
      mbar_wait(bar_sfull, tile_ph, 3);
      tc_fence_after();
      for (int cs = 0; cs < p.n1; cs += 32) {
        uint32_t r[32];
        tmem_ld_32x32(tmem_s + lane_off + cs, r);
        tmem_ld_wait();
        uint32_t w[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float x0 = __uint_as_float(r[2 * i]) * p.alpha1, x1 = __uint_as_float(r[2 * i + 1]) * p.alpha1;
          if (p.relu_mid) { x0 = fmaxf(x0, 0.f); x1 = fmaxf(x1, 0.f); }
          if (p.mid_bf16) {
            __nv_bfloat162 h = __float22bfloat162_rn(make_float2(x0, x1));  // .x = lower K index = low half
            w[i] = *reinterpret_cast<uint32_t*>(&h);
          } else {
            __half2 h = __float22half2_rn(make_float2(x0, x1));
            w[i] = *reinterpret_cast<uint32_t*>(&h);
          }
        }
        tmem_st_32x16(tmem_p + lane_off + cs / 2, w);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_pfull);
      tile_ph ^= 1u;
      const int grow = tm * kBlockM + q * 32 + lane;
      for (int tn = 0; tn < p.tiles_n2; ++tn) {
        mbar_wait(bar_ofull + 8 * ob, o_ph, 3);
        tc_fence_after();
        const uint32_t tmem_o = tmem_o0 + 128 * ob;
        for (int cs = 0; cs < kChainOWidth; cs += 32) {
          const int gcol = tn * kChainOWidth + cs;
          if (gcol >= p.n2) break;  // warp-uniform
          uint32_t r[32];
          tmem_ld_32x32(tmem_o + lane_off + cs, r);
          tmem_ld_wait();
          if (cs + 32 >= kChainOWidth || gcol + 32 >= p.n2) {  // last read of this accumulator: hand it back
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_oempty + 8 * ob);
          }
          if (grow < p.m) {
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha2;
            const long long base = static_cast<long long>(grow) * p.ldc + gcol;
            if (vec_ok && gcol + 32 <= p.n2) {  // this thread's 32 columns as 16-byte packets (64 or 128 contiguous bytes)
              if (p.out_kind == kOutBF16) {
                uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.c) + base);
#pragma unroll
                for (int pk = 0; pk < 4; ++pk) dst[pk] = pack16<kOutBF16>(&v[pk * 8]);
              } else if (p.out_kind == kOutF16) {
                uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__half*>(p.c) + base);
#pragma unroll
                for (int pk = 0; pk < 4; ++pk) dst[pk] = pack16<kOutF16>(&v[pk * 8]);
              } else {
                uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<float*>(p.c) + base);
#pragma unroll
                for (int pk = 0; pk < 8; ++pk) dst[pk] = pack16<kOutF32>(&v[pk * 4]);
              }
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                if (gcol + i < p.n2) {
                  if (p.out_kind == kOutBF16) store_scalar<kOutBF16>(p.c, base + i, v[i]);
                  else if (p.out_kind == kOutF16) store_scalar<kOutF16>(p.c, base + i, v[i]);
                  else store_scalar<kOutF32>(p.c, base + i, v[i]);
                }
              }
            }
          }
          __syncwarp();
        }
        if (++ob == o_bufs) { ob = 0; o_ph ^= 1u; }
      }
    }
  }

  __syncwarp();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc<1>(tmem_base, 512);
}

}  // namespace tlb200
