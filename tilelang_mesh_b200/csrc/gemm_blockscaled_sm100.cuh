// gemm_blockscaled_sm100.cuh — fp8 contraction with fine-grained scales ("2x accumulation").
//
//   C[m, n] = sum_kb  scales_a[m, kb] * scales_b[n / 128, kb] * ( sum_{k in K-block kb} A[m, k] * B[n, k] )
//
// This is SURVEY §8(f) N1: the reference writes it as a tile program whose K loop promotes the
// tensor-core partial sum of every 128-wide K block into a second fp32 accumulator
// (examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:43-83:
//     T.gemm(A_shared, B_shared, C_local, transpose_B=True)
//     C_local_accum[i, j] += C_local[i, j] * Scale_C_shared[i];  T.clear(C_local)).
//
// B200 form: the same persistent warp-specialised ring as gemm_sm100.cuh, except that the MMA warp
// CLEARS the TMEM accumulator at every K block (one K block = 128 fp8 elements = one 128B-swizzled
// stage = 4 tcgen05.mma of K = 32) and publishes each partial through a ring of TMEM buffers that
// fills all 512 columns (2 buffers of 256 columns, or 4 of 128); the eight promotion warps pull each
// partial out with tcgen05.ld (32 columns at a time, the next load in flight while the current
// chunk is folded) and add it into register accumulators with packed fp32 fmas (fma.rn.f32x2:
// two IEEE fmas per instruction, bit-identical to fmaf).  scale = scales_a * scales_b is rounded
// once, as the reference computes Scale_C_shared.  The accumulators are cast and stored once per
// tile.  The promotion warpgroups raise their register budget with setmaxnreg (128 accumulators +
// 64 staging registers per thread for BN = 256); the control warpgroup lowers its own.
//
// What bounds it (profiles/r1_notes.md, "block-scaled"): not TMEM bandwidth (tcgen05.ld sustains
// ~900 B/clk/SM with 8 warps, tools/ubench/tmem_read_width.cu; a K block's partial is 128 KiB =
// ~145 clk) and not the tensor pipe (512 clk per K block at BN = 256, 75 % busy), but the promotion
// warps themselves: 2 per scheduler, each K block costs them 64 FFMA2 + 4 loads + waits on a chain
// that has only two TMEM buffers of slack.  Measured 2.4-2.5 PFLOP/s at 8192^3 = 0.75 x the plain
// fp8 kernel of this library on the same box.
#pragma once
#include "gemm_sm100.cuh"

namespace tlb200 {

constexpr int kBsEpiWarps = 8;
constexpr int kBsThreads = (kEpiWarp0 + kBsEpiWarps) * 32;  // 384

struct alignas(64) BlockScaledParams {
  CUtensorMap tm_a;  // A[M, K] fp8, box 128 rows x 128 B
  CUtensorMap tm_b;  // B[N, K] fp8, box (BN / cta_group) rows x 128 B
  CUtensorMap tm_c;  // C store boxes (32 rows x 128 B), valid iff epi_tma
  void* c;
  long long ldc;
  const float* sa;   // scales_a[m, kb] at sa[m * sa_sm + kb * sa_sk]
  const float* sb;   // scales_b[nb, kb] at sb[nb * sb_sn + kb * sb_sk]
  long long sa_sm, sa_sk, sb_sn, sb_sk;
  int m, n;
  int num_k_blocks;
  int tiles_m, tiles_n;
  int raster;
  int stages;
  int b_rows_per_cta;
  uint32_t idesc;
  uint32_t desc_hi;  // upper word of the K-major 128B-swizzle smem descriptor
  int out_kind;
  int epi_tma;
  int l2_policy;
};


// Two independent IEEE fp32 fmas in one instruction (fma.rn.f32x2, sm_100): {d0, d1} += {x0, x1} * {s, s}.
// Same results as two fmaf; half the issue slots of the promotion loop.
__device__ __forceinline__ void fma2(float& d0, float& d1, uint32_t x0, uint32_t x1, uint64_t s2) {
  uint64_t d, x;
  asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(d0), "f"(d1));
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "r"(x0), "r"(x1));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(x), "l"(s2));
  asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

// Store this thread's kCols consecutive accumulators (tile row = TMEM lane) of one tile.
template <int kOut, int kCols>
__device__ __forceinline__ void bs_store_tile(const BlockScaledParams& p, const float (&acc)[kCols], int q, int lane,
                                              int row0, int col_beg, uint32_t slab) {
  constexpr int kOB = OutTraits<kOut>::kBytes;
  constexpr int kColsPerSlab = 128 / kOB;
  constexpr int kValsPerPacket = 16 / kOB;
  constexpr int kPacketsPerSlab = 8;
  const int grow = row0 + q * 32 + lane;
  if (p.epi_tma) {
#pragma unroll
    for (int cs = 0; cs < kCols; cs += kColsPerSlab) {
      if (col_beg + cs < p.n) {  // warp-uniform
        uint4 w[kPacketsPerSlab];
#pragma unroll
        for (int pk = 0; pk < kPacketsPerSlab; ++pk) w[pk] = pack16<kOut>(&acc[cs + pk * kValsPerPacket]);
        if (lane == 0) tma_store_wait_read<0>();  // the previous store has finished reading the slab
        __syncwarp();
#pragma unroll
        for (int pk = 0; pk < kPacketsPerSlab; ++pk) {
          const uint32_t addr = slab + lane * 128 + ((pk ^ (lane & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[pk].x), "r"(w[pk].y), "r"(w[pk].z),
                       "r"(w[pk].w)
                       : "memory");
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          if (row0 + q * 32 < p.m) {
            if (p.l2_policy != 1) tma_store_2d_hint(&p.tm_c, slab, col_beg + cs, row0 + q * 32, policy_evict_first());
            else tma_store_2d(&p.tm_c, slab, col_beg + cs, row0 + q * 32);
          }
          tma_store_commit();
        }
      }
    }
  } else {
    if (grow < p.m) {
      const long long base = static_cast<long long>(grow) * p.ldc + col_beg;
#pragma unroll
      for (int i = 0; i < kCols; ++i)
        if (col_beg + i < p.n) store_scalar<kOut>(p.c, base + i, acc[i]);
    }
    __syncwarp();
  }
}

// kCta: CTAs per tile (2 = a cluster pair shares every B stage, tile = 256 x kBN); kBN: tile width (128 or 256).
template <int kCta, int kBN>
__global__ void __launch_bounds__(kBsThreads, 1)
gemm_blockscaled_kernel(const __grid_constant__ BlockScaledParams p) {
  extern __shared__ uint8_t smem_raw[];
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = static_cast<int>(lane_id());
  const uint32_t cta_rank = (kCta == 2) ? cluster_ctarank() : 0u;
  const bool leader = cta_rank == 0;

  constexpr int kBRows = kBN / kCta;                       // rows of B one CTA stages per K block
  constexpr uint32_t kBBytesPerStage = kBRows * kStageRowBytes;
  constexpr uint32_t kStageBytes = kABytesPerStage + kBBytesPerStage;
  constexpr int kNumBuf = 512 / kBN;                       // partial-sum buffers: the ring fills TMEM (2 or 4)
  constexpr uint32_t kTmemCols = 512;
  constexpr int kCols = kBN / 2;                           // accumulator columns per epilogue thread
  const int stages = p.stages;
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smem_a = smem_base;
  const uint32_t smem_b = smem_a + stages * kABytesPerStage;
  const uint32_t smem_epi = smem_b + stages * kBBytesPerStage;
  const uint32_t smem_bar = smem_epi + kBsEpiWarps * kEpiBufBytes;
  const uint32_t bar_full = smem_bar;
  const uint32_t bar_empty = bar_full + 8 * stages;
  const uint32_t bar_tfull = bar_empty + 8 * stages;   // [kNumBuf]
  const uint32_t bar_tempty = bar_tfull + 32;          // [kNumBuf]
  const uint32_t smem_tptr = bar_tempty + 32;

  if (warp == 0) {
    if (elect_one()) {
      tma_prefetch_desc(&p.tm_a);
      tma_prefetch_desc(&p.tm_b);
      if (p.epi_tma) tma_prefetch_desc(&p.tm_c);
    }
  } else if (warp == 1) {
    if (elect_one()) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(bar_full + 8 * s, 1);
        mbar_init(bar_empty + 8 * s, 1);
      }
      for (int a = 0; a < kNumBuf; ++a) {
        mbar_init(bar_tfull + 8 * a, 1);
        mbar_init(bar_tempty + 8 * a, kBsEpiWarps * kCta);
      }
      fence_mbar_init();
    }
  } else if (warp == 2) {
    tmem_alloc<kCta>(smem_tptr, kTmemCols);
  }
  __syncwarp();
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem_base) : "r"(smem_tptr) : "memory");

  const int num_tiles = p.tiles_m * p.tiles_n;
  const int cluster_id = blockIdx.x / kCta;
  const int num_clusters = gridDim.x / kCta;
  const int nkb = p.num_k_blocks;
  auto tile_coord = [&](int t, int& tm, int& tn) {
    const int group = p.raster * p.tiles_n;
    const int g = t / group;
    const int first = g * p.raster;
    const int rows = min(p.raster, p.tiles_m - first);
    const int within = t - g * group;
    tm = first + within % rows;
    tn = within / rows;
  };

  if (warp < kEpiWarp0) {
    // ===================================================== control warpgroup: give registers away
    setmaxnreg_dec<56>();
    if (warp != 1) {
      // ---- TMA producers: warps 0, 2, 3 take the K blocks round-robin (at most `stages` of them; gemm_sm100.cuh)
      const int kProducers = stages < 3 ? stages : 3;
      const int pidx = warp == 0 ? 0 : warp - 1;
      int turn = 0;
      const bool elected = elect_one();
      const uint64_t pol = p.l2_policy == 1 ? policy_evict_normal() : policy_evict_last();
      int s = 0;
      uint32_t ph = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        int tm, tn;
        tile_coord(t, tm, tn);
        const int row_a = (tm * kCta + (int)cta_rank) * kBlockM;
        const int row_b = tn * kBN + (int)cta_rank * kBRows;
        for (int kb = 0; kb < nkb; ++kb) {
          const bool mine = turn == pidx;
          if (++turn == kProducers) turn = 0;
          if (mine) mbar_wait(bar_empty + 8 * s, ph ^ 1u, 0);
          if (mine && elected) {
            const uint32_t full = bar_full + 8 * s;
            if (leader) mbar_arrive_expect_tx(full, kStageBytes * kCta);
            const int k0 = kb * kStageRowBytes;  // fp8: one element per byte
            if constexpr (kCta == 2) {
              tma_load_2d_pair(smem_a + s * kABytesPerStage, &p.tm_a, full, k0, row_a, pol);
              tma_load_2d_pair(smem_b + s * kBBytesPerStage, &p.tm_b, full, k0, row_b, pol);
            } else {
              tma_load_2d(smem_a + s * kABytesPerStage, &p.tm_a, full, k0, row_a, pol);
              tma_load_2d(smem_b + s * kBBytesPerStage, &p.tm_b, full, k0, row_b, pol);
            }
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    } else if (leader) {
      // ---- MMA issuer (warp 1): every K block starts a fresh partial sum in TMEM buffer `as`
      const bool elected = elect_one();
      const uint64_t hi = static_cast<uint64_t>(p.desc_hi) << 32 | (1u << 16);
      const uint32_t a_lo0 = (smem_a & 0x3FFFFu) >> 4, b_lo0 = (smem_b & 0x3FFFFu) >> 4;
      int s = 0, as = 0;
      uint32_t ph = 0, aph = 0;
      for (int t = cluster_id; t < num_tiles; t += num_clusters) {
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait(bar_tempty + 8 * as, aph ^ 1u, 1);  // the promotion has read this buffer
          mbar_wait(bar_full + 8 * s, ph, 2);
          tc_fence_after();
          if (elected) {
            const uint32_t a_lo = a_lo0 + s * (kABytesPerStage >> 4);
            const uint32_t b_lo = b_lo0 + s * (kBBytesPerStage >> 4);
#pragma unroll
            for (int k = 0; k < kStageRowBytes / kUmmaKBytes; ++k) {
              const uint64_t da = hi | ((a_lo + k * (kUmmaKBytes >> 4)) & 0x3FFFu);
              const uint64_t db = hi | ((b_lo + k * (kUmmaKBytes >> 4)) & 0x3FFFu);
              umma_ss<kCta, kKindF8>(tmem_base + as * kBN, da, db, p.idesc, k == 0 ? 0u : 1u);
            }
            umma_commit<kCta>(bar_empty + 8 * s);
            umma_commit<kCta>(bar_tfull + 8 * as);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1u; }
          if (++as == kNumBuf) { as = 0; aph ^= 1u; }
        }
      }
    }
  } else {
    // ===================================================== promotion + epilogue warps
    setmaxnreg_inc<224>();
    const int ew = warp - kEpiWarp0;
    const int q = ew & 3;        // TMEM lane quarter
    const int chalf = ew >> 2;   // column half of the tile
    const uint32_t my_slab = smem_epi + ew * kEpiBufBytes;
    const uint32_t tempty_leader = (kCta == 2) ? mapa(bar_tempty, 0) : bar_tempty;
    // Loop-invariant addresses are pinned in registers (an opaque mov): left alone, ptxas re-derives the shared-
    // memory window base (S2UR SR_CgaCtaId + a ULEA chain) in every K block, ~40 % of the loop's instructions.
    auto pin = [](uint32_t v) { uint32_t r; asm volatile("mov.b32 %0, %1;" : "=r"(r) : "r"(v)); return r; };
    const uint32_t lane_taddr = pin(tmem_base + (static_cast<uint32_t>(q * 32) << 16) + chalf * kCols);
    const uint32_t tfull0 = pin(bar_tfull);
    const uint32_t tempty0 = pin(tempty_leader);
    constexpr int kChunks = kCols / 32;
    static_assert(kChunks % 2 == 0, "the two staging buffers alternate across K blocks");
    int as = 0;
    uint32_t aph = 0;
    for (int t = cluster_id; t < num_tiles; t += num_clusters) {
      int tm, tn;
      tile_coord(t, tm, tn);
      const int row0 = (tm * kCta + (int)cta_rank) * kBlockM;
      const int col_beg = tn * kBN + chalf * kCols;
      const int grow = row0 + q * 32 + lane;
      // scale pointers of this thread's row / this warp's 128-column block (rows / columns outside the
      // problem contribute zero: their operands were zero-filled by TMA, the scale just has to be finite)
      const bool row_ok = grow < p.m, col_ok = col_beg < p.n;
      const float* sa_ptr = p.sa + static_cast<long long>(row_ok ? grow : 0) * p.sa_sm;
      const float* sb_ptr = p.sb + static_cast<long long>(col_ok ? col_beg / 128 : 0) * p.sb_sn;
      float acc[kCols];
#pragma unroll
      for (int i = 0; i < kCols; ++i) acc[i] = 0.f;
      // The two factors of the next K block's scale are fetched a whole K block ahead and multiplied only when
      // needed: a multiply placed right behind the loads stalled every iteration for the full global-load latency
      // (20 % of the promotion warps' samples, profiles/r1_notes.md).
      float sa_next = __ldg(sa_ptr), sb_next = __ldg(sb_ptr);
      // One continuous pipeline over (K block, 32-column chunk): while chunk c is folded into the accumulators the
      // load of the next chunk - possibly the first chunk of the NEXT K block's buffer - is already in flight.
      uint32_t r[2][32];
      bool next_started = false;
      mbar_wait(tfull0 + 8 * as, aph, 3);
      tc_fence_after();
      tmem_ld_32x32(lane_taddr + as * kBN, r[0]);
      for (int kb = 0; kb < nkb; ++kb) {
        const float scale = sa_next * sb_next;
        uint64_t scale2;  // {scale, scale} for the packed fma
        asm("mov.b64 %0, {%1, %1};" : "=l"(scale2) : "f"(scale));
        const bool more = kb + 1 < nkb;
        if (more) {
          sa_ptr += p.sa_sk;
          sb_ptr += p.sb_sk;
          sa_next = __ldg(sa_ptr);
          sb_next = __ldg(sb_ptr);
        }
        const uint32_t taddr = lane_taddr + as * kBN;
#pragma unroll
        for (int c = 0; c < kChunks; ++c) {
          tmem_ld_wait();
          if (c + 1 < kChunks) {
            tmem_ld_32x32(taddr + (c + 1) * 32, r[(c + 1) & 1]);
          } else {
            // the whole partial is in registers: hand the TMEM buffer back, then start on the next K block's
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if constexpr (kCta == 2) mbar_arrive_cluster(tempty0 + 8 * as);
              else mbar_arrive(tempty0 + 8 * as);
            }
            if (++as == kNumBuf) { as = 0; aph ^= 1u; }
            // Start on the next K block's partial right away if it is already complete; otherwise fold the last
            // chunk first and wait afterwards (waiting first idled the FFMAs: measured 2366 -> 2084 TFLOP/s).
            next_started = false;
            if (more) {
              next_started = __shfl_sync(0xffffffffu, mbar_test_wait(tfull0 + 8 * as, aph) ? 1 : 0, 0) != 0;
              if (next_started) {
                tc_fence_after();
                tmem_ld_32x32(lane_taddr + as * kBN, r[0]);
              }
            }
          }
#pragma unroll
          for (int i = 0; i < 32; i += 2)
            fma2(acc[c * 32 + i], acc[c * 32 + i + 1], r[c & 1][i], r[c & 1][i + 1], scale2);
        }
        if (more && !next_started) {
          mbar_wait(tfull0 + 8 * as, aph, 3);
          tc_fence_after();
          tmem_ld_32x32(lane_taddr + as * kBN, r[0]);
        }
      }
      if (row0 < p.m) {
        switch (p.out_kind) {
          case kOutBF16: bs_store_tile<kOutBF16, kCols>(p, acc, q, lane, row0, col_beg, my_slab); break;
          case kOutF16: bs_store_tile<kOutF16, kCols>(p, acc, q, lane, row0, col_beg, my_slab); break;
          default: bs_store_tile<kOutF32, kCols>(p, acc, q, lane, row0, col_beg, my_slab); break;
        }
      }
    }
    if (p.epi_tma && lane == 0) tma_store_wait_all<0>();
  }

  __syncwarp();
  tc_fence_before();
  if constexpr (kCta == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  if (warp == 2) tmem_dealloc<kCta>(tmem_base, kTmemCols);
}

}  // namespace tlb200
