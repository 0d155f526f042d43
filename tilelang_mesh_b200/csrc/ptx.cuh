// ptx.cuh — thin inline-PTX wrappers for the sm_100a features the GEMM kernel uses:
// mbarrier, TMA (cp.async.bulk.tensor), tcgen05 (alloc / mma / commit / ld / fences) and
// cluster helpers.  Hand-written; no CUTLASS/CuTe.
//
// Reference counterparts (replaced, not ported):
//   mbarrier / fences      src/tl_templates/cuda/barrier.h:7-162
//   TMA load / store       src/tl_templates/cuda/copy_sm90.h:17-268
//   tcgen05 alloc/commit   src/tl_templates/cuda/tcgen_05.h:13-72
//   tcgen05.mma            src/tl_templates/cuda/instruction/tcgen05mma.h:14-337
//   tcgen05.ld             src/tl_templates/cuda/tcgen_05_ld.h, copy_sm100.h:100-147
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdio>

namespace tlb200 {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ uint32_t lane_id() {
  uint32_t l;
  asm volatile("mov.u32 %0, %%laneid;" : "=r"(l));
  return l;
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// True on exactly one lane of a fully converged warp.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint64_t globaltimer_ns() {
  uint64_t t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// ----------------------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Arrive on a barrier addressed in the shared::cluster window (possibly another CTA's).  Relaxed: the barriers this
// is used for hand back TENSOR memory whose reads have already completed (tcgen05.wait::ld) and are ordered by
// tcgen05.fence::before_thread_sync; no global or shared memory is published.  The .release.cluster form compiles to
// MEMBAR.ALL.GPU + ERRBAR, which also waits for every TMA store the thread has in flight: ~1.8k cycles per
// accumulator hand-back in the epilogue (profiles/r2_notes.md).
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.relaxed.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar)
               : "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t smem_addr, uint32_t cta_rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(cta_rank));
  return r;
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

// Non-blocking probe of a phase parity (try_wait may suspend the warp for a while before answering).
__device__ __forceinline__ bool mbar_test_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}

#ifndef TLB200_WATCHDOG_NS
#define TLB200_WATCHDOG_NS 4000000000ull  // 4 s: a pipeline that waits this long is deadlocked
#endif

// Spin on a phase parity.  try_wait suspends in hardware for a bounded time, so the loop
// body runs rarely; the watchdog converts a protocol deadlock into a trap (launch error)
// instead of a hung GPU.  `tag` identifies the waiting role in the diagnostic.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity, int tag) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!mbar_try_wait(bar, parity)) {
    if ((++spins & 0x3ffu) == 0) {  // the timer is read only after ~1000 failed polls: %globaltimer is slow
      const uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > TLB200_WATCHDOG_NS) {
        printf("tlb200 watchdog: block %d thread %d tag %d bar 0x%x parity %u\n", (int)blockIdx.x,
               (int)threadIdx.x, tag, bar, parity);
        __trap();
      }
    }
  }
}

// The same wait for a CONVERGED warp, with a warp-uniform exit: every lane polls and the loop ends on a vote.  After
// mbar_wait() ptxas treats what follows as possibly divergent and keeps it out of the uniform datapath (every
// tcgen05.mma operand then goes through an R2UR); the vote's result is uniform by construction, so the code after
// this wait stays in uniform registers.  All 32 lanes must call it.
__device__ __forceinline__ void mbar_wait_uniform(uint32_t bar, uint32_t parity, int tag) {
  uint64_t t0 = 0;
  uint32_t spins = 0;
  while (!__all_sync(0xffffffffu, mbar_try_wait(bar, parity))) {
    if ((++spins & 0x3ffu) == 0) {
      const uint64_t now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > TLB200_WATCHDOG_NS) {
        printf("tlb200 watchdog: block %d thread %d tag %d bar 0x%x parity %u\n", (int)blockIdx.x,
               (int)threadIdx.x, tag, bar, parity);
        __trap();
      }
    }
  }
}

// ----------------------------------------------------------------------------- fences
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::
                   : "memory");
}

// ----------------------------------------------------------------------------- register budget per warpgroup
// setmaxnreg moves registers between the warpgroups (4 consecutive warps) of a CTA: the control warpgroup gives
// most of its share away, the epilogue warpgroups take it.  Every warp of a warpgroup must execute the same one.
template <int kRegs>
__device__ __forceinline__ void setmaxnreg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegs));
}
template <int kRegs>
__device__ __forceinline__ void setmaxnreg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegs));
}

// ----------------------------------------------------------------------------- programmatic dependent launch
// `griddep_wait` blocks until the grids this launch depends on have completed and their memory is visible (a no-op
// for a launch without the programmatic-serialization attribute); `griddep_launch_dependents` lets the next launch
// of the stream begin its own set-up while this grid is still running.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ----------------------------------------------------------------------------- TMA
__device__ __forceinline__ void tma_prefetch_desc(const void* desc) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(desc)) : "memory");
}
// 2-D tiled load global -> shared, completion signalled on `bar` (this CTA's).
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* desc, uint32_t bar, int c0,
                                            int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// Same, issued by either CTA of a cta_group::2 pair; the transaction bytes are credited
// to the barrier at the same offset in the EVEN (leader) CTA of the pair: clearing bit 24
// of a shared::cluster address selects the pair's leader.
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const void* desc, uint32_t bar,
                                                 int c0, int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// Pair variant that also MULTICASTS the box: it lands at the same smem offset in every CTA of `cta_mask`
// (bit i = cluster rank i) and each destination's bytes are credited to the barrier at the same offset in
// that destination's pair leader.  Used to share one A box between the two pairs of a 4-CTA cluster.
__device__ __forceinline__ void tma_load_2d_pair_mc(uint32_t dst, const void* desc, uint32_t bar, int c0, int c1,
                                                    uint16_t cta_mask, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      ".multicast::cluster.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5, %6;" ::"r"(dst),
      "l"(reinterpret_cast<uint64_t>(desc)), "r"(bar & 0xFEFFFFFFu), "r"(c0), "r"(c1), "h"(cta_mask), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const void* desc, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(desc)),
               "r"(src), "r"(c0), "r"(c1)
               : "memory");
}
// Same with an L2 eviction-priority hint (C is written once and never re-read by this kernel).
__device__ __forceinline__ void tma_store_2d_hint(const void* desc, uint32_t src, int c0, int c1, uint64_t policy) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group.L2::cache_hint [%0, {%2, %3}], [%1], %4;" ::"l"(
                   reinterpret_cast<uint64_t>(desc)),
               "r"(src), "r"(c0), "r"(c1), "l"(policy)
               : "memory");
}
__device__ __forceinline__ void tma_store_commit() {
  asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// L2 eviction policies for TMA loads (createpolicy encodings).
__device__ __forceinline__ uint64_t policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}

// ----------------------------------------------------------------------------- TMEM
template <int kCta>
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t ncols) {
  if constexpr (kCta == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(dst_smem),
                 "r"(ncols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}
template <int kCta>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  if constexpr (kCta == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
  } else {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols)
                 : "memory");
  }
}

// ----------------------------------------------------------------------------- tcgen05.mma
// D[tmem] (+)= A[smem desc] * B[smem desc]; `accumulate` = 0 overwrites D.
// kKind selects the instruction kind by operand type: fp16/bf16, e4m3/e5m2, fp32 read as tf32, int8.
enum MmaKind : int { kKindF16 = 0, kKindF8 = 1, kKindTF32 = 2, kKindI8 = 3 };

#define TLB200_UMMA(GROUP, KIND)                                                                   \
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"                                  \
               "tcgen05.mma.cta_group::" GROUP ".kind::" KIND " [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), \
               "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)                                \
               : "memory")

template <int kCta, int kKind>
__device__ __forceinline__ void umma_ss(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b,
                                        uint32_t idesc, uint32_t accumulate) {
  if constexpr (kCta == 1) {
    if constexpr (kKind == kKindF8) TLB200_UMMA("1", "f8f6f4");
    else if constexpr (kKind == kKindTF32) TLB200_UMMA("1", "tf32");
    else if constexpr (kKind == kKindI8) TLB200_UMMA("1", "i8");
    else TLB200_UMMA("1", "f16");
  } else {
    if constexpr (kKind == kKindF8) TLB200_UMMA("2", "f8f6f4");
    else if constexpr (kKind == kKindTF32) TLB200_UMMA("2", "tf32");
    else if constexpr (kKind == kKindI8) TLB200_UMMA("2", "i8");
    else TLB200_UMMA("2", "f16");
  }
}
#undef TLB200_UMMA

// The four MMAs of one 128-byte K block as ONE asm statement: the descriptors' low words advance by `kadv` (16-byte
// units) per MMA, their high words and the instruction descriptor stay.  Written as a block so that nothing is scheduled
// between the four issues and the descriptor arithmetic is four additions (the issue loop of a narrow tile is bound by
// this thread's own instruction latency, not by the tensor pipe: profiles/r2_notes.md section 12).  The low word must
// not carry out of its 14-bit address field: a K block never leaves its stage, and shared memory ends below 2^18.
#define TLB200_UMMA4(GROUP, KIND)                                                                                   \
  asm volatile(                                                                                                     \
      "{\n\t.reg .pred p, t;\n\t.reg .b64 da, db;\n\t.reg .b32 aw, bw;\n\t"                                          \
      "setp.ne.b32 p, %8, 0;\n\tsetp.eq.b32 t, %7, %7;\n\t"                                                         \
      "mov.b64 da, {%1, %2};\n\tmov.b64 db, {%3, %4};\n\t"                                                          \
      "tcgen05.mma.cta_group::" GROUP ".kind::" KIND " [%0], da, db, %7, p;\n\t"                                     \
      "add.u32 aw, %1, %5;\n\tadd.u32 bw, %3, %6;\n\tmov.b64 da, {aw, %2};\n\tmov.b64 db, {bw, %4};\n\t"             \
      "tcgen05.mma.cta_group::" GROUP ".kind::" KIND " [%0], da, db, %7, t;\n\t"                                     \
      "add.u32 aw, aw, %5;\n\tadd.u32 bw, bw, %6;\n\tmov.b64 da, {aw, %2};\n\tmov.b64 db, {bw, %4};\n\t"             \
      "tcgen05.mma.cta_group::" GROUP ".kind::" KIND " [%0], da, db, %7, t;\n\t"                                     \
      "add.u32 aw, aw, %5;\n\tadd.u32 bw, bw, %6;\n\tmov.b64 da, {aw, %2};\n\tmov.b64 db, {bw, %4};\n\t"             \
      "tcgen05.mma.cta_group::" GROUP ".kind::" KIND " [%0], da, db, %7, t;\n\t}" ::"r"(tmem_d),                     \
      "r"(a_w), "r"(a_hi), "r"(b_w), "r"(b_hi), "r"(kadv_a), "r"(kadv_b), "r"(idesc), "r"(accumulate)               \
      : "memory")
template <int kCta, int kKind>
__device__ __forceinline__ void umma_ss_x4(uint32_t tmem_d, uint32_t a_w, uint32_t a_hi, uint32_t b_w, uint32_t b_hi,
                                           uint32_t kadv_a, uint32_t kadv_b, uint32_t idesc, uint32_t accumulate) {
  if constexpr (kCta == 1) {
    if constexpr (kKind == kKindF8) TLB200_UMMA4("1", "f8f6f4");
    else if constexpr (kKind == kKindTF32) TLB200_UMMA4("1", "tf32");
    else if constexpr (kKind == kKindI8) TLB200_UMMA4("1", "i8");
    else TLB200_UMMA4("1", "f16");
  } else {
    if constexpr (kKind == kKindF8) TLB200_UMMA4("2", "f8f6f4");
    else if constexpr (kKind == kKindTF32) TLB200_UMMA4("2", "tf32");
    else if constexpr (kKind == kKindI8) TLB200_UMMA4("2", "i8");
    else TLB200_UMMA4("2", "f16");
  }
}
#undef TLB200_UMMA4

// Make `bar` observe completion of every tcgen05.mma this thread issued so far.
// (implies tcgen05.fence::before_thread_sync).  For a pair the arrive is multicast to the
// same barrier offset in both CTAs.
// `cta_mask`: cluster ranks whose barrier (same offset) receives the arrive; default = both CTAs of a pair
// that sits at cluster ranks 0 and 1.
template <int kCta>
__device__ __forceinline__ void umma_commit(uint32_t bar, uint16_t cta_mask = 0x3) {
  if constexpr (kCta == 1) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     bar)
                 : "memory");
  } else {
    asm volatile(
        "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64"
        " [%0], %1;" ::"r"(bar),
        "h"(cta_mask)
        : "memory");
  }
}

// ----------------------------------------------------------------------------- tcgen05.ld
// 32 lanes x 32 consecutive fp32 columns: thread `l` of the warp receives row
// (32*(warp%4) + l), columns [col, col+32) of the accumulator.
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]),
        "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]),
        "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() {
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

}  // namespace tlb200
