// tlb200.cu — host side of libtlb200.so: the C ABI declared in include/tlb200.h.
//
// Mirrors the host wrapper the reference generates per kernel
// (tilelang/jit/adapter/wrapper.py:31-51 trio, :76-99 TMA descriptor encode, :202-295 call):
// validate -> plan -> encode CUtensorMaps on the stack -> one launch on the caller's stream.
// No allocation, no retained pointers, thread-local error text.
#include "tlb200.h"

#include <array>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "kernel_table.h"

namespace {

using namespace tlb200;

thread_local char g_err[1024] = {0};
std::atomic<long long> g_launches{0};

int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return -1;
}

// ---- driver entry point (bound at init; the library does not link libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn g_encode = nullptr;
// Pointer queries behind own_device_extent() below (bound at init like the encoder).
typedef CUresult (*PtrAttrFn)(void*, CUpointer_attribute, CUdeviceptr);
typedef CUresult (*AddrRangeFn)(CUdeviceptr*, size_t*, CUdeviceptr);
typedef CUresult (*RetainHandleFn)(CUmemGenericAllocationHandle*, void*);
typedef CUresult (*ReleaseHandleFn)(CUmemGenericAllocationHandle);
PtrAttrFn g_ptr_attr = nullptr;
AddrRangeFn g_addr_range = nullptr;
RetainHandleFn g_retain_handle = nullptr;
ReleaseHandleFn g_release_handle = nullptr;
bool g_no_multicast = false;  // TLB200_NO_MULTICAST=1 in the environment: the A-multicast plan is never used
int g_kgroup = 0;             // TLB200_KGROUP=1|2|4 in the environment: K blocks per stage barrier of narrow tiles (0 = automatic)
std::mutex g_init_mu;
constexpr int kMaxDevices = 64;
std::atomic<int> g_sm_count[kMaxDevices];   // zero-initialised (static storage)
std::atomic<bool> g_dev_ready[kMaxDevices];
// Host-buffer path: per device one copy-in and one copy-out stream plus an event ring, created at
// init (streams/events are driver objects, not device memory; the library still allocates none).
constexpr int kMaxPanels = 32;  // arrivals per call: two operands x up to 16 panels
struct HostPipe {
  cudaStream_t in = nullptr, out = nullptr;
  cudaEvent_t fork = nullptr, loaded[kMaxPanels] = {}, done[kMaxPanels] = {}, drained = nullptr;
};
HostPipe g_pipe[kMaxDevices];

constexpr int kSmemLimit = 232448;  // 227 KiB opt-in per block on sm_100
constexpr int smem_fixed(int epi_slabs) { return 1024 /*alignment slack*/ + epi_slabs * kEpiBufBytes + (2 * kMaxStages + 6) * 8 + 16; }

int dtype_bytes(int dt) {
  switch (dt) {
    case TLB200_FLOAT16:
    case TLB200_BFLOAT16: return 2;
    case TLB200_FLOAT8_E4M3:
    case TLB200_FLOAT8_E5M2: return 1;
    case TLB200_INT8: return 1;
    case TLB200_FLOAT32:
    case TLB200_INT32: return 4;
    default: return 0;
  }
}
bool is_ab_dtype(int dt) {
  return dt == TLB200_FLOAT16 || dt == TLB200_BFLOAT16 || dt == TLB200_FLOAT8_E4M3 ||
         dt == TLB200_FLOAT8_E5M2 || dt == TLB200_FLOAT32 || dt == TLB200_INT8;
}
// tcgen05.mma kind of an operand type (fp32 operands are read as tf32)
int mma_kind_of(int ab) {
  switch (ab) {
    case TLB200_FLOAT8_E4M3:
    case TLB200_FLOAT8_E5M2: return kKindF8;
    case TLB200_FLOAT32: return kKindTF32;
    case TLB200_INT8: return kKindI8;
    default: return kKindF16;
  }
}

uint32_t instr_desc_impl(int atom_m, int atom_n, int atom_k, int ab, int c, bool a_k_major,
                         bool b_k_major, int sa, int sb) {
  // Field map: src/op/tcgen5_meta.h:142-171.
  uint32_t fmt;
  switch (ab) {
    case TLB200_FLOAT16: fmt = 0; break;
    case TLB200_BFLOAT16: fmt = 1; break;
    case TLB200_FLOAT8_E4M3: fmt = 0; break;
    case TLB200_FLOAT8_E5M2: fmt = 1; break;
    case TLB200_FLOAT32: fmt = 2; break;  // kind::tf32 (the reference's emitter never encodes it; PTX ISA value)
    case TLB200_INT8: fmt = 1; break;     // kind::i8: 0 = unsigned, 1 = signed
    default: return 0;
  }
  uint32_t cfmt;
  if (c == TLB200_FLOAT16) cfmt = 0;
  else if (c == TLB200_FLOAT32) cfmt = 1;
  else if (c == TLB200_INT32) cfmt = 2;
  else return 0;
  if ((ab == TLB200_INT8) != (c == TLB200_INT32)) return 0;
  if (atom_m % 16 || atom_n % 8 || (atom_k != 8 && atom_k != 16 && atom_k != 32)) return 0;
  if ((sa != 1 && sa != -1) || (sb != 1 && sb != -1)) return 0;
  uint32_t d = 0;
  d |= (cfmt & 3u) << 4;
  d |= (fmt & 7u) << 7;
  d |= (fmt & 7u) << 10;
  d |= (sa == -1 ? 1u : 0u) << 13;
  d |= (sb == -1 ? 1u : 0u) << 14;
  d |= (a_k_major ? 0u : 1u) << 15;
  d |= (b_k_major ? 0u : 1u) << 16;
  d |= (static_cast<uint32_t>(atom_n >> 3) & 0x3fu) << 17;
  d |= (static_cast<uint32_t>(atom_m >> 4) & 0x1fu) << 24;
  return d;
}

uint64_t smem_desc_impl(uint32_t addr, int lbo16, int sbo16, int base_offset, int lbo_abs, int swz) {
  // Bit layout: src/tl_templates/cuda/common.h:481-496 (Tcgen05SMemDescriptor).
  uint64_t d = 0;
  d |= static_cast<uint64_t>((addr >> 4) & 0x3fffu);
  d |= static_cast<uint64_t>(lbo16 & 0x3fff) << 16;
  d |= static_cast<uint64_t>(sbo16 & 0x3fff) << 32;
  d |= 1ull << 46;  // version_ = 1
  d |= static_cast<uint64_t>(base_offset & 7) << 49;
  d |= static_cast<uint64_t>(lbo_abs ? 1 : 0) << 52;
  d |= static_cast<uint64_t>(swz & 7) << 61;
  return d;
}

struct Resolved {
  tlb200_gemm_plan plan;
  int es, ob;             // operand / output element bytes
  int span;               // elements per 128 B of an operand row
  bool a_mn, b_mn;
  int n_sub, umma_n, acc_stages, epi_warps, epi_slabs;
  int mc;                 // 2 = clusters of two pairs sharing every A box by TMA multicast, else 1
  int half;               // 1 = pair tiles of 128 rows (64 per CTA, cta_group::2 MMAs with M = 128)
  int a_stage_bytes;      // shared-memory bytes per A stage (GemmKernelParams::a_stage_bytes)
  bool short_m;           // the problem is shorter than one single-CTA tile: A boxes hold only its rows
  int b_rows_per_cta;
  int num_k_blocks;
  bool epi_tma, epi_lsu;
  // split-K: the main kernel writes fp32 partial tiles into a workspace, a second kernel reduces
  int split_k, kb_per_split, m_pad;
  long long ws_ld, ws_bytes;  // ws_bytes = aligned operand copies + split-K partial sums
  // operands TMA cannot address (row pitch or base not a multiple of 16 bytes) are first copied into the
  // workspace with a 128-byte-multiple pitch (the reference falls back to SIMT copies: src/op/copy.cc:823-855)
  bool copy_a, copy_b, tr_a, tr_b;
  long long lda_c, ldb_c, a_copy_bytes, b_copy_bytes, split_bytes;
  // tail split (optional: used only when the caller's workspace also holds tail_bytes beyond ws_bytes): the
  // tiles of the last, partial wave are cut into K slices; see GemmKernelParams::tail_*
  int tail_begin, tail_tiles, tail_split, tail_kb;
  long long tail_bytes;
};

// Resident 4-CTA clusters (one CTA per SM) on a GPU with `sms` SMs: measured by tlb200_init with
// cudaOccupancyMaxActiveClusters (33 on a 148-SM B200: GPC sizes are not multiples of 4); estimate otherwise.
int g_mc_clusters_sms = 0, g_mc_clusters = 0;
// Cycle-model efficiencies of the A-multicast plans (tools/sweep.py); 0 = never chosen automatically.
constexpr double kEffMc256 = 1.05, kEffMc128 = 0.78, kEffMc64 = 0.41;
constexpr bool kAutoMc = true;  // multicast plans are chosen automatically
int mc_clusters_of(int sms) { return (g_mc_clusters > 0 && sms == g_mc_clusters_sms) ? g_mc_clusters : sms * 33 / 148; }

int resolve(const tlb200_gemm_desc* d, int sm_count, const void* c_ptr, Resolved* r, bool allow_split = false,
            const void* a_ptr = nullptr, const void* b_ptr = nullptr, bool allow_mc = false, bool allow_tail = true,
            bool allow_half = true) {
  if (!d) return fail("tlb200: null descriptor");
  if (d->abi_version != TLB200_ABI_VERSION)
    return fail("tlb200: descriptor abi_version %d, library speaks %d", d->abi_version,
                TLB200_ABI_VERSION);
  if (d->m < 0 || d->n < 0 || d->k < 0) return fail("tlb200: negative extent (m=%d n=%d k=%d)", d->m, d->n, d->k);
  if (d->reserved2[1] || d->reserved2[2]) return fail("tlb200: descriptor fields reserved2[1..2] must be zero");
  if (d->reserved2[0] != 0 && d->reserved2[0] != 1 && d->reserved2[0] != 2 && d->reserved2[0] != 4)
    return fail("tlb200: cfg_kgroup must be 0 (automatic), 1, 2 or 4 (got %d)", d->reserved2[0]);
  if (d->vec_flags & ~(TLB200_VEC_ROW_SCALE | TLB200_VEC_COL_SCALE | TLB200_VEC_BIAS_N | TLB200_VEC_BIAS_M))
    return fail("tlb200: unknown bits in vec_flags (0x%x)", d->vec_flags);
  if (!is_ab_dtype(d->ab_dtype))
    return fail("tlb200: unsupported operand dtype %d (float16, bfloat16, float8_e4m3, float8_e5m2)", d->ab_dtype);
  if (dtype_bytes(d->c_dtype) == 0 || d->c_dtype == TLB200_INT8) return fail("tlb200: unsupported output dtype %d", d->c_dtype);
  if (d->c_dtype == TLB200_INT32 && d->ab_dtype != TLB200_INT8) return fail("tlb200: int32 output needs int8 operands");
  if (d->ab_dtype == TLB200_INT8 && (d->c_dtype == TLB200_FLOAT8_E4M3 || d->c_dtype == TLB200_FLOAT8_E5M2))
    return fail("tlb200: int8 operands produce int32, float32, float16 or bfloat16 output (not fp8)");
  if (d->c_dtype == TLB200_INT32 && d->alpha != 1.0f) return fail("tlb200: int32 output cannot be scaled (alpha must be 1)");
  r->es = dtype_bytes(d->ab_dtype);
  r->ob = dtype_bytes(d->c_dtype);
  r->span = 128 / r->es;
  r->a_mn = d->trans_a != 0;
  r->b_mn = d->trans_b == 0;
  const long long a_cols = r->a_mn ? d->m : d->k, b_cols = r->b_mn ? d->n : d->k;
  if (d->lda < a_cols) return fail("tlb200: lda=%lld < %lld", (long long)d->lda, a_cols);
  if (d->ldb < b_cols) return fail("tlb200: ldb=%lld < %lld", (long long)d->ldb, b_cols);
  if (d->ldc < d->n) return fail("tlb200: ldc=%lld < n=%d", (long long)d->ldc, d->n);
  // TMA needs 16-byte row pitches and bases (same constraint as the reference's bulk copy,
  // src/op/copy.cc:1489-1499: strides in bytes, % 16, < 2^40).  Operands that violate it are staged
  // through the workspace; without one they are rejected.
  const bool live = d->k > 0 && d->m > 0 && d->n > 0;
  // fp32 (tf32) operands whose M/N dimension is contiguous would need the 32-byte-atom swizzle variant in
  // shared memory; they are transposed into K-major form through the workspace instead.
  r->tr_a = live && d->ab_dtype == TLB200_FLOAT32 && r->a_mn;
  r->tr_b = live && d->ab_dtype == TLB200_FLOAT32 && r->b_mn;
  r->copy_a = r->tr_a || (live && ((d->lda * r->es) % 16 || reinterpret_cast<uintptr_t>(a_ptr) % 16));
  r->copy_b = r->tr_b || (live && ((d->ldb * r->es) % 16 || reinterpret_cast<uintptr_t>(b_ptr) % 16));
  if ((r->copy_a || r->copy_b) && !allow_split) {
    if (r->tr_a || r->tr_b)
      return fail("tlb200: float32 (tf32) operands with a contiguous M/N dimension are transposed through a workspace; "
                  "call tlb200_gemm_ex with one");
    return fail("tlb200: operand row pitch and base must be a multiple of 16 bytes for TMA (lda=%lld ldb=%lld, %d-byte elements); "
                "call tlb200_gemm_ex with a workspace and the operands are re-pitched first",
                (long long)d->lda, (long long)d->ldb, r->es);
  }
  if (r->tr_a) r->a_mn = false;  // after the transposing copy the kernel sees K-major operands
  if (r->tr_b) r->b_mn = false;
  const long long a_rows_c = r->a_mn ? d->k : d->m, a_cols_c = r->a_mn ? d->m : d->k;  // shape of the staged copy
  const long long b_rows_c = r->b_mn ? d->k : d->n, b_cols_c = r->b_mn ? d->n : d->k;
  r->lda_c = r->copy_a ? (a_cols_c * r->es + 127) / 128 * 128 / r->es : d->lda;
  r->ldb_c = r->copy_b ? (b_cols_c * r->es + 127) / 128 * 128 / r->es : d->ldb;
  r->a_copy_bytes = r->copy_a ? (a_rows_c * r->lda_c * r->es + 255) / 256 * 256 : 0;
  r->b_copy_bytes = r->copy_b ? (b_rows_c * r->ldb_c * r->es + 255) / 256 * 256 : 0;

  tlb200_gemm_plan& p = r->plan;
  memset(&p, 0, sizeof(p));
  if (d->cfg_cta_group < 0 || d->cfg_cta_group > 4)
    return fail("tlb200: cfg_cta_group must be 0, 1, 2, 3 or 4 (got %d)", d->cfg_cta_group);
  if (d->reserved[1] < -1 || d->reserved[1] > 64)
    return fail("tlb200: cfg_split_k must be in [0,64], or -1 to force the tail split (got %d)", d->reserved[1]);
  if (d->reserved[1] > 1 && d->ab_dtype == TLB200_INT8) return fail("tlb200: split-K is not offered for int8 (partial sums are kept in fp32)");
  // A tile is n_sub accumulators of umma_n (<= 256) columns each.  The epilogue drains TMEM 32
  // columns at a time, so umma_n is a whole multiple of 32; an N-major B is staged in whole 128 B
  // spans per CTA.  Tiles wider than 256 columns (two accumulators sharing one A stage: 25 % less
  // L2->SM traffic) fill the 512 TMEM columns, so their accumulator is single-buffered.
  // cfg_cta_group = 4: pairs (cta_group::2 MMAs) in clusters of two pairs that share A by multicast
  // cfg_cta_group = 3: pairs whose tile is 128 rows (64 per CTA): for the same tile area a CTA stages fewer operand
  // bytes per K block than with 128 rows and half the columns, and narrow tiles are bound by exactly those bytes
  bool half = d->cfg_cta_group == 3;
  const bool half_possible = allow_half && !r->a_mn && d->m > kBlockM / 2;
  if (half && !half_possible)
    return fail("tlb200: cfg_cta_group=3 (pairs of 64-row CTAs) needs a K-contiguous A, m > %d and an ordinary (not grouped, "
                "single-destination) contraction", kBlockM / 2);
  const int cfg_cg = (d->cfg_cta_group == 4 || half) ? 2 : d->cfg_cta_group;
  int cg = cfg_cg, bn = d->cfg_block_n, auto_split = 1, mc = d->cfg_cta_group == 4 ? 2 : 1;
  const int sms_all = sm_count > 0 ? sm_count : 148;
  const int mc_clusters = mc_clusters_of(sms_all);
  const bool mc_possible = allow_mc && !r->a_mn && d->m > kBlockM && mc_clusters >= 1;
  if (mc == 2 && !mc_possible)
    return fail("tlb200: cfg_cta_group=4 (A multicast across two pairs) needs a K-contiguous A, m > %d and an ordinary "
                "(not grouped, single-destination) contraction%s", kBlockM,
                g_no_multicast ? "; TLB200_NO_MULTICAST is set" : "");
  if (cg == 0 || bn == 0) {
    // Automatic plan: minimise a small cycle model calibrated on B200 (tools/sweep.py,
    // profiles/r1_notes.md) over {cta_group} x {block_n} x {split_k}.
    //   K-block cycles per CTA = 2*block_n / eff(cta_group, block_n)   (512 for a 256-wide tile)
    //   eff: how well the operand feed keeps the tensor pipe busy for that tile shape (measured at
    //   4096^3); wide tiles get a bonus for their lower power per flop but expose one epilogue per tile.
    // mc = 3 marks the pair tiles of 128 rows (64 per CTA): a CTA then stages 8 KB of A and block_n / 2 rows of B per
    // K block, and a narrow tile's K block costs its bytes at ~82 B/clk (profiles/r2_notes.md section 12)
    struct Cand { int cg, bn, mc; double eff; };
    // Tiles narrower than 256 columns: eff = 2 * block_n / (bytes one CTA stages per K block / 82 B per clock), the
    // measured main-loop pace since the stage barriers are grouped (245 / 269 / 319 / 406 cycles for 256x64, 256x128,
    // 128x64, 128x128; tools/tune_b2b.py is the check: which shapes still have a better explicit plan, and by how much).
    const Cand cands[] = {{2, 512, 1, 1.07}, {2, 256, 1, 1.0}, {1, 256, 1, 0.93}, {2, 128, 1, 0.95}, {1, 128, 1, 0.63},
                          {1, 64, 1, 0.40}, {2, 64, 1, 0.52}, {1, 32, 1, 0.26},
                          {2, 256, 2, kEffMc256}, {2, 128, 2, kEffMc128}, {2, 64, 2, kEffMc64},
                          {2, 64, 3, 0.85}, {2, 128, 3, 1.28}, {2, 256, 3, 1.70}};
    const int nkb = (d->k + r->span - 1) / r->span;
    double best = 1e300;
    int best_cg = 0, best_bn = 0, best_split = 1, best_mc = 1;
    for (const Cand& c0 : cands) {
      Cand c = c0;
      const bool c_half = c.mc == 3;
      if (c_half) {
        c.mc = 1;
        if (!half_possible || (d->cfg_cta_group != 0 && d->cfg_cta_group != 3) || d->cfg_epilogue >= 2 && (c.bn / 2) % (128 / r->ob)) continue;
      } else if (half) continue;
      if (c.mc == 2 && d->cfg_cta_group != 4 && (!kAutoMc || c.bn != 256)) continue;  // measured: only 256-wide tiles gain
      if (cfg_cg && c.cg != cfg_cg) continue;
      if (c.mc == 2 && (!mc_possible || d->cfg_cta_group == 2 || d->n <= c.bn)) continue;
      if (c.mc == 1 && d->cfg_cta_group == 4) continue;
      if (d->cfg_block_n && c.bn != d->cfg_block_n) continue;
      if (c.cg == 2 && ((d->m <= kBlockM && !c_half) || sms_all < 2)) continue;
      if (c.bn == 512 && (r->es != 2 || d->n < 512)) continue;  // 8-bit operands: double-buffered tiles measure faster
      if (c.bn > 32 && c.bn / 2 >= d->n && c.bn != d->cfg_block_n) continue;  // tile twice wider than the matrix
      const int step = r->b_mn ? r->span * c.cg : 32;
      if ((c.bn > 256 ? c.bn / 2 : c.bn) % step) continue;
      if (d->cfg_epilogue >= 2 && (c.bn > 256 ? c.bn / 2 : c.bn) % (128 / r->ob)) continue;  // staging slabs are 128 B wide
      const int c_rows = c_half ? kBlockM : kBlockM * c.cg;
      const long long tiles = static_cast<long long>((d->m + c_rows - 1) / c_rows) *
                              (((d->n + c.bn - 1) / c.bn + c.mc - 1) / c.mc);  // work items: c.mc column tiles each
      const long long clusters = c.mc == 2 ? mc_clusters : sms_all / c.cg;
      for (int split = 1; split <= 32; split *= 2) {
        if (split > 1 && (!allow_split || d->ab_dtype == TLB200_INT8 || d->reserved[1] == 1 || nkb / split < 8 || tiles * split > 2 * clusters)) break;
        if (d->reserved[1] > 1 && split > d->reserved[1]) break;
        const long long items = tiles * split;
        const long long waves = (items + clusters - 1) / clusters;
        const double kb = static_cast<double>((nkb + split - 1) / split);
        // K-block cycles: 2 * block_n / eff for the 256- and 512-wide tiles (calibrated on whole runs); narrower tiles
        // and the 128-row pairs from what bounds them (profiles/r2_notes.md sections 12, 13): the bytes one CTA stages
        // at ~82 B/clk, the tensor pipe (MACs, or its operand reads at 128 B/clk), the barrier hand-shake per group.
        double kb_cycles = 2.0 * c.bn / c.eff;
        if (c.mc == 1 && (c.bn < 256 || c_half)) {
          const int rows_cta = c_half ? kBlockM / 2 : kBlockM;
          const int a_rows = (c.cg == 1 && d->m < kBlockM) ? (d->m + 7) / 8 * 8 : rows_cta;  // short problems fetch only their rows
          const int b_rows = c.bn / c.cg;
          const double bytes = (a_rows + b_rows) * 128.0 / 82.0;
          const double mma = rows_cta * c.bn / 256.0, reads = (rows_cta + b_rows) / 4.0;
          const double pipe = 4.0 * (mma > reads ? mma : reads);
          const double shake = 300.0 / (c.bn <= 32 ? 4 : c.bn <= 128 ? 2 : 1);
          kb_cycles = bytes > pipe ? bytes : pipe;
          if (shake > kb_cycles) kb_cycles = shake;
        }
        double tile = kb * kb_cycles + (c.bn > 256 ? 8000.0 : 0.0);
        double total = 3000.0 + waves * tile + (c.bn > 256 ? 0.0 : 12.0 * c.bn) + (c.cg == 2 ? ((c.bn >= 256 && !c_half) ? 3000.0 : 500.0) : 0.0);  // pairs start later (cluster launch and sync)
        if (split == 1 && allow_split && allow_tail && d->reserved[1] != 1 && c.mc == 1 && c.bn <= 256 && d->ab_dtype != TLB200_INT8 &&
            items > clusters && items % clusters != 0) {
          // the tail split (decided below) shortens the last, partial wave to 1/S of a wave plus a small reduction
          const long long rem = items % clusters;
          long long s_tail = clusters / rem;
          if (s_tail > 8) s_tail = 8;
          if (s_tail > nkb / 8) s_tail = nkb / 8;
          if (s_tail >= 2) {
            const double cost = 9000.0 + static_cast<double>(rem) * kBlockM * c.cg * c.bn * (s_tail * 4.0 + 4.0) / 2500.0;
            const double with_tail = total - tile * (1.0 - 1.0 / static_cast<double>(s_tail)) + cost;
            if (with_tail < total) total = with_tail;
          }
        }
        if (split > 1) total += 3000.0 + (static_cast<double>(split) * 4.0 + 2.0) * d->m * d->n / 2500.0;
        if (total < best) { best = total; best_cg = c.cg; best_bn = c.bn; best_split = split; best_mc = c_half ? 3 : c.mc; }
      }
    }
    if (best_cg == 0) {  // nothing matched the constraints: fall back to the smallest legal tile
      best_cg = cfg_cg ? cfg_cg : 1;
      best_mc = d->cfg_cta_group == 4 ? 2 : 1;
      best_bn = d->cfg_block_n ? d->cfg_block_n : (r->b_mn ? r->span * best_cg : 32);
    }
    cg = best_cg; bn = best_bn; auto_split = best_split; mc = best_mc;
    if (mc == 3) { mc = 1; half = true; }
  }
  r->half = half ? 1 : 0;
  if (sms_all < 2) cg = 1;
  const int n_step = r->b_mn ? r->span * cg : 32;
  r->n_sub = bn > 256 ? 2 : 1;
  r->umma_n = bn / r->n_sub;
  if (bn < n_step || bn > 512 || bn % r->n_sub || r->umma_n % n_step)
    return fail("tlb200: cfg_block_n=%d invalid for cta_group=%d%s (need a multiple of %d in [%d,256], or of %d in (256,512])",
                bn, cg, r->b_mn ? ", N-major B" : "", n_step, n_step, 2 * n_step);
  if (mc == 2 && (r->n_sub != 1 || cg != 2))
    return fail("tlb200: cfg_cta_group=4 needs block_n <= 256");
  if (half && (r->n_sub != 1 || cg != 2 || mc != 1 || bn % 64))
    return fail("tlb200: cfg_cta_group=3 needs block_n a multiple of 64, at most 256");
  r->mc = mc;
  r->acc_stages = (2 * bn <= 512) ? 2 : 1;
  int ew = d->reserved[0];  // cfg_epi_warps
  if (ew == 0) {  // 8 warps drain a 512-wide tile unless its halves do not split into whole 128-byte slabs
    const int half = r->umma_n / 2;
    ew = (bn > 256 && half % 32 == 0 && half % (128 / r->ob) == 0) ? 8 : 4;
  }
  if (ew != 4 && ew != 8) return fail("tlb200: cfg_epi_warps must be 0, 4 or 8 (got %d)", ew);
  if (((half ? r->umma_n / 2 : r->umma_n) / (ew / 4)) % 32) return fail("tlb200: block_n=%d cannot be split over %d epilogue warps", bn, ew);
  r->epi_warps = ew;
  r->b_rows_per_cta = r->umma_n / cg;
  // A problem shorter than one tile (decode-time shapes) fetches only the rows it has, rounded up to the 8-row swizzle
  // atom, and its A stages are that small: more stages fit, and with cold operands the bytes in flight are the pace.
  r->short_m = allow_half && cg == 1 && !r->a_mn && d->m < kBlockM;
  r->a_stage_bytes = half ? kABytesPerStage / 2 : r->short_m ? (d->m + 7) / 8 * 8 * kStageRowBytes : kABytesPerStage;
  const int per_stage = r->a_stage_bytes + r->n_sub * r->b_rows_per_cta * kStageRowBytes;
  int stages = d->cfg_stages;
  // Eight epilogue warps get two staging slabs each (so a slab can be refilled while the TMA store of the previous
  // one is still reading it) whenever at least three pipeline stages still fit, or the stage count the caller asked for.
  r->epi_slabs = kEpiSlabs;
  if (ew == 8) {
    const int fit16 = (kSmemLimit - smem_fixed(16)) / per_stage;
    if (stages == 0 ? fit16 >= 3 : stages <= fit16) r->epi_slabs = 16;
  }
  const int max_stages = (kSmemLimit - smem_fixed(r->epi_slabs)) / per_stage;
  if (stages == 0) stages = max_stages < kMaxStages ? max_stages : kMaxStages;
  if (stages < 2 || stages > kMaxStages || stages > max_stages)
    return fail("tlb200: cfg_stages=%d out of range [2,%d]", stages, max_stages < kMaxStages ? max_stages : kMaxStages);
  p.cta_group = cg;
  p.block_m = half ? kBlockM : kBlockM * cg;
  p.block_n = bn;
  p.block_k = r->span;
  p.stages = stages;
  p.tiles_m = (d->m + p.block_m - 1) / p.block_m;
  p.tiles_n = (d->n + bn - 1) / bn;
  r->num_k_blocks = (d->k + r->span - 1) / r->span;
  const long long tiles = static_cast<long long>(p.tiles_m) * ((p.tiles_n + mc - 1) / mc);  // work items per K slice
  const int sms = sm_count > 0 ? sm_count : 148;
  long long clusters = mc == 2 ? mc_clusters : sms / cg;
  // Split-K (examples/gemm_splitk in the reference): when the output has too few tiles to occupy the
  // machine and K is deep, K is cut into slices computed by different clusters into an fp32
  // workspace and summed by a second kernel in slice order (deterministic).
  int split = d->reserved[1];  // cfg_split_k: 0 auto, 1 off, n = n slices (fewer if K is short), -1 = tail split only
  const bool force_tail = split == -1;
  if (force_tail) split = 1;
  const bool tail_off = d->reserved[1] == 1;
  if (split == 0) {
    split = auto_split;
    if (split == 1 && allow_split && d->ab_dtype != TLB200_INT8 && (d->cfg_cta_group != 0 && d->cfg_block_n != 0) && tiles * 2 <= clusters && r->num_k_blocks >= 16) {
      long long s2 = clusters / tiles;  // fully explicit tile: simple fill rule
      if (s2 > r->num_k_blocks / 8) s2 = r->num_k_blocks / 8;
      if (s2 > 32) s2 = 32;
      if (s2 >= 2) split = static_cast<int>(s2);
    }
  } else if (split > 1 && (d->cfg_cta_group == 0 || d->cfg_block_n == 0)) {
    split = auto_split > 1 ? auto_split : split;
  }
  if (split > r->num_k_blocks) split = r->num_k_blocks > 0 ? r->num_k_blocks : 1;
  r->kb_per_split = split > 1 ? (r->num_k_blocks + split - 1) / split : (r->num_k_blocks > 0 ? r->num_k_blocks : 1);
  r->split_k = split > 1 ? (r->num_k_blocks + r->kb_per_split - 1) / r->kb_per_split : 1;
  r->m_pad = p.tiles_m * p.block_m;
  r->ws_ld = (static_cast<long long>(d->n) + 3) / 4 * 4;
  r->split_bytes = r->split_k > 1 ? static_cast<long long>(r->split_k) * r->m_pad * r->ws_ld * 4 : 0;
  r->ws_bytes = r->a_copy_bytes + r->b_copy_bytes + r->split_bytes;
  p.split_k = r->split_k;
  // Tail split: with more than one wave of whole tiles, the last wave may hold only a few tiles.  Cutting just those
  // along K over all clusters shortens it from one wave to 1/S of one, at the price of a small reduction kernel.
  r->tail_begin = r->tail_tiles = r->tail_split = r->tail_kb = 0;
  r->tail_bytes = 0;
  if (allow_split && allow_tail && !tail_off && r->split_k == 1 && r->n_sub == 1 && mc == 1 && d->ab_dtype != TLB200_INT8 &&
      tiles > clusters && tiles % clusters != 0 && d->k > 0) {
    const long long rem = tiles % clusters;
    long long s_tail = clusters / rem;
    if (s_tail > 8) s_tail = 8;
    if (s_tail > r->num_k_blocks / 8) s_tail = r->num_k_blocks / 8;
    if (s_tail >= 2) {
      const double wave = static_cast<double>(r->num_k_blocks) * 2.0 * bn;           // tensor cycles of one wave
      const double saved = wave * (1.0 - 1.0 / static_cast<double>(s_tail));
      const double cost = 9000.0 + static_cast<double>(rem) * p.block_m * bn * (s_tail * 4.0 + 2.0 + r->ob) / 2500.0;
      if (force_tail || saved > 1.25 * cost) {
        r->tail_kb = static_cast<int>((r->num_k_blocks + s_tail - 1) / s_tail);
        r->tail_split = (r->num_k_blocks + r->tail_kb - 1) / r->tail_kb;
        r->tail_tiles = static_cast<int>(rem);
        r->tail_begin = static_cast<int>(tiles - rem);
        r->tail_bytes = (static_cast<long long>(r->tail_split) * rem * p.block_m * bn * 4 + 255) / 256 * 256;
      }
    }
  }
  p.workspace_bytes = r->ws_bytes + r->tail_bytes;
  const long long work_items = r->tail_tiles ? r->tail_begin + static_cast<long long>(r->tail_tiles) * r->tail_split : tiles * r->split_k;
  if (clusters > work_items) clusters = work_items;
  if (clusters < 1) clusters = 1;
  p.grid = static_cast<int>(clusters) * cg * mc;
  p.multicast = mc;
  p.threads = 128 + 32 * r->epi_warps;
  p.smem_bytes = smem_fixed(r->epi_slabs) + stages * per_stage;
  p.raster = d->cfg_raster > 0 ? d->cfg_raster : ((bn > 256 || r->es == 1) ? 16 : 16 / cg);
  int epi = d->cfg_epilogue;
  const bool accumulate = (d->epilogue & TLB200_EPI_ACCUMULATE) != 0;
  if (accumulate) {
    if (d->epilogue & TLB200_EPI_RELU) return fail("tlb200: TLB200_EPI_ACCUMULATE cannot be combined with TLB200_EPI_RELU");
    if (d->c_dtype == TLB200_INT32) return fail("tlb200: TLB200_EPI_ACCUMULATE is not offered for int32 output");
    if (epi >= 2) return fail("tlb200: TLB200_EPI_ACCUMULATE uses the predicated-store epilogue (cfg_epilogue 0 or 1)");
    epi = 1;
  }
  // A TMA store clips at the tensor's last column only in whole 16-byte units (measured: tools/pad_probe.py), so a
  // row whose byte length is not a multiple of 16 must take the predicated stores or the cells right of column n
  // in a padded C would be overwritten.
  const bool tma_ok = ((d->ldc * r->ob) % 16 == 0) && ((reinterpret_cast<uintptr_t>(c_ptr) % 16) == 0) &&
                      ((static_cast<long long>(d->n) * r->ob) % 16 == 0);
  // The TMA-store epilogue writes whole 128-byte-wide slabs: the tile must be a whole number of them.
  const bool slab_ok = ((half ? r->umma_n / 2 : r->umma_n) / (r->epi_warps / 4)) % (128 / r->ob) == 0;
  if (epi == 0) epi = (tma_ok && slab_ok) ? 2 : 1;
  if (epi >= 2 && !tma_ok) return fail("tlb200: the staged epilogues (cfg_epilogue 2, 3) need a 16-byte aligned C, row pitch and row length (n * element size)");
  if (epi >= 2 && !slab_ok)
    return fail("tlb200: the staged epilogues (cfg_epilogue 2, 3) need block_n (%d) to be a multiple of %d for %d-byte outputs", bn, 128 / r->ob, r->ob);
  if (epi < 1 || epi > 3) return fail("tlb200: cfg_epilogue must be 0, 1, 2 or 3");
  p.epilogue = epi;
  r->epi_tma = epi >= 2;
  r->epi_lsu = epi == 3;
  const int acc_cols = (half ? bn / 2 : bn) * r->acc_stages;
  p.tmem_cols = acc_cols <= 32 ? 32 : acc_cols <= 64 ? 64 : acc_cols <= 128 ? 128 : acc_cols <= 256 ? 256 : 512;
  p.umma_n = r->umma_n;
  p.epi_warps = r->epi_warps;
  p.instr_desc = instr_desc_impl(p.block_m, r->umma_n, 32 / r->es, d->ab_dtype, d->ab_dtype == TLB200_INT8 ? TLB200_INT32 : TLB200_FLOAT32,
                                 !r->a_mn, !r->b_mn, 1, 1);
  if (p.instr_desc == 0) return fail("tlb200: cannot encode instruction descriptor");
  return 0;
}

CUtensorMapDataType tm_dtype(int dt) {
  switch (dt) {
    case TLB200_FLOAT16: return CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
    case TLB200_BFLOAT16: return CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
    case TLB200_FLOAT32: return CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
    case TLB200_INT32: return CU_TENSOR_MAP_DATA_TYPE_INT32;
    default: return CU_TENSOR_MAP_DATA_TYPE_UINT8;
  }
}

// 2-D row-major tensor [rows, cols] (cols contiguous, pitch `ld` elements) -> tiled map with a
// [box_rows x box_cols] box and the 128B swizzle.
int encode_2d(CUtensorMap* tm, int dt, const void* base, long long rows, long long cols, long long ld,
              int box_cols, int box_rows, const char* what) {
  const int eb = dtype_bytes(dt);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * eb};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)};
  cuuint32_t estr[2] = {1, 1};
  CUresult rc = g_encode(tm, tm_dtype(dt), 2, const_cast<void*>(base), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                         CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (rc != CUDA_SUCCESS)
    return fail("tlb200: cuTensorMapEncodeTiled(%s) failed with CUresult %d (base=%p rows=%lld cols=%lld ld=%lld box=%dx%d)",
                what, (int)rc, base, rows, cols, ld, box_rows, box_cols);
  return 0;
}

template <int kOut>
__global__ void fill_zero_kernel(void* c, long long ldc, int m, int n) {
  const long long total = static_cast<long long>(m) * n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / n, col = i - r * n;
    store_scalar<kOut>(c, r * ldc + col, 0.f);
  }
}

// Launch a plain kernel as a programmatic dependent of the previous launch in `stream`: its blocks may become
// resident while that kernel still runs and proceed at their griddepcontrol.wait the moment it completes, which
// takes the launch latency of the reduction kernels off the critical path.
template <typename... KArgs, typename... Args>
cudaError_t launch_dependent(void (*kernel)(KArgs...), int grid, int block, cudaStream_t stream, Args... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(block, 1, 1);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Epilogue vectors as the reduction kernels take them (null = absent); same arithmetic, in the same order, as
// epi_apply_vectors in the GEMM kernel: v = ((alpha * sum) * row_scale[m]) * col_scale[n] + bias.
struct ReduceVectors {
  const float* row_scale;
  const float* col_scale;
  const float* bias;
  int bias_per_row;
};
__device__ __forceinline__ float reduce_apply_vectors(float v, const ReduceVectors& vec, long long row, int col) {
  if (vec.row_scale) v = v * vec.row_scale[row];
  if (vec.col_scale) v = v * vec.col_scale[col];
  if (vec.bias) v = v + vec.bias[vec.bias_per_row ? row : col];
  return v;
}

// Split-K second pass: C[m,n] = cast(relu?(vectors(alpha * sum_s ws[s*m_pad + m, n]))), slices summed in order.
template <int kOut>
__global__ void splitk_reduce_kernel(const float* __restrict__ ws, long long ws_ld, int m_pad, int split_k, void* c,
                                     long long ldc, int m, int n, float alpha, int relu, int accumulate, ReduceVectors vec) {
  griddep_wait();  // the partial sums come from the previous kernel
  const int n4 = (n + 3) / 4;
  const long long total = static_cast<long long>(m) * n4;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / n4;
    const int col = static_cast<int>(i - row * n4) * 4;
    float4 acc = *reinterpret_cast<const float4*>(ws + row * ws_ld + col);  // ws_ld is a multiple of 4
    for (int s = 1; s < split_k; ++s) {
      const float4 v = *reinterpret_cast<const float4*>(ws + (static_cast<long long>(s) * m_pad + row) * ws_ld + col);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float o[4] = {acc.x * alpha, acc.y * alpha, acc.z * alpha, acc.w * alpha};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (col + j < n) {
        o[j] = reduce_apply_vectors(o[j], vec, row, col + j);
        if (relu) o[j] = fmaxf(o[j], 0.f);
        if (accumulate) o[j] += load_out<kOut>(c, row * ldc + col + j);
        store_scalar<kOut>(c, row * ldc + col + j, o[j]);
      }
    }
  }
}

int launch_splitk_reduce(const float* ws, long long ws_ld, int m_pad, int split_k, void* c, long long ldc, int m, int n,
                         float alpha, bool relu, bool accumulate, int out_kind, cudaStream_t stream, const ReduceVectors& vec) {
  const long long total = static_cast<long long>(m) * ((n + 3) / 4);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  if (blocks < 1) blocks = 1;
  const int g = static_cast<int>(blocks), rl = relu ? 1 : 0, ac = accumulate ? 1 : 0;
  cudaError_t e = cudaSuccess;
  switch (out_kind) {
    case kOutF16: e = launch_dependent(splitk_reduce_kernel<kOutF16>, g, 256, stream, ws, ws_ld, m_pad, split_k, c, ldc, m, n, alpha, rl, ac, vec); break;
    case kOutBF16: e = launch_dependent(splitk_reduce_kernel<kOutBF16>, g, 256, stream, ws, ws_ld, m_pad, split_k, c, ldc, m, n, alpha, rl, ac, vec); break;
    case kOutF32: e = launch_dependent(splitk_reduce_kernel<kOutF32>, g, 256, stream, ws, ws_ld, m_pad, split_k, c, ldc, m, n, alpha, rl, ac, vec); break;
    case kOutE4M3: e = launch_dependent(splitk_reduce_kernel<kOutE4M3>, g, 256, stream, ws, ws_ld, m_pad, split_k, c, ldc, m, n, alpha, rl, ac, vec); break;
    default: e = launch_dependent(splitk_reduce_kernel<kOutE5M2>, g, 256, stream, ws, ws_ld, m_pad, split_k, c, ldc, m, n, alpha, rl, ac, vec); break;
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return fail("tlb200: split-K reduce launch: %s", cudaGetErrorString(e));
  return 0;
}

// Tail split second pass: tile j of the tail (work item order of the GEMM kernel) is the sum over its K slices s of
// workspace tiles (s * tail_tiles + j), each [block_m x block_n] fp32, summed in slice order (deterministic).
template <int kOut>
__global__ void tail_reduce_kernel(const float* __restrict__ ws, int tail_begin, int tail_tiles, int tail_split, int block_m,
                                   int block_n, int tiles_m, int tiles_n, int raster, void* c, long long ldc, int m, int n,
                                   float alpha, int relu, int accumulate, ReduceVectors vec) {
  griddep_wait();  // the partial tiles come from the previous kernel
  const int n4 = block_n / 4;
  const long long per_tile = static_cast<long long>(block_m) * n4;
  const long long total = per_tile * tail_tiles;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int j = static_cast<int>(i / per_tile);
    const int rem = static_cast<int>(i - j * per_tile);
    const int r = rem / n4, c4 = (rem - r * n4) * 4;
    // tile index -> (row tile, column tile): the kernel's grouped raster (gemm_sm100.cuh, tile_coord)
    const int t = tail_begin + j;
    const int group = raster * tiles_n;
    const int g = t / group;
    const int first = g * raster;
    const int rows = min(raster, tiles_m - first);
    const int within = t - g * group;
    const int tm = first + within % rows, tn = within / rows;
    const long long row = static_cast<long long>(tm) * block_m + r;
    const int col = tn * block_n + c4;
    if (row >= m || col >= n) continue;
    float4 acc = *reinterpret_cast<const float4*>(ws + (static_cast<long long>(j) * block_m + r) * block_n + c4);
    for (int sl = 1; sl < tail_split; ++sl) {
      const float4 v = *reinterpret_cast<const float4*>(ws + ((static_cast<long long>(sl) * tail_tiles + j) * block_m + r) * block_n + c4);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    float o[4] = {acc.x * alpha, acc.y * alpha, acc.z * alpha, acc.w * alpha};
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      if (col + q < n) {
        o[q] = reduce_apply_vectors(o[q], vec, row, col + q);
        if (relu) o[q] = fmaxf(o[q], 0.f);
        if (accumulate) o[q] += load_out<kOut>(c, row * ldc + col + q);
        store_scalar<kOut>(c, row * ldc + col + q, o[q]);
      }
    }
  }
}

int launch_tail_reduce(const float* ws, const Resolved& r, const tlb200_gemm_plan& pl, void* c, long long ldc, int m, int n,
                       float alpha, bool relu, bool accumulate, int out_kind, cudaStream_t stream, const ReduceVectors& vec) {
  const long long total = static_cast<long long>(r.tail_tiles) * pl.block_m * (pl.block_n / 4);
  long long blocks = (total + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const int g = static_cast<int>(blocks < 1 ? 1 : blocks), rl = relu ? 1 : 0, ac = accumulate ? 1 : 0;
#define TLB_TAIL(K) launch_dependent(tail_reduce_kernel<K>, g, 256, stream, ws, r.tail_begin, r.tail_tiles, r.tail_split, pl.block_m, \
                                    pl.block_n, pl.tiles_m, pl.tiles_n, pl.raster, c, ldc, m, n, alpha, rl, ac, vec)
  cudaError_t e = cudaSuccess;
  switch (out_kind) {
    case kOutF16: e = TLB_TAIL(kOutF16); break;
    case kOutBF16: e = TLB_TAIL(kOutBF16); break;
    case kOutF32: e = TLB_TAIL(kOutF32); break;
    case kOutE4M3: e = TLB_TAIL(kOutE4M3); break;
    default: e = TLB_TAIL(kOutE5M2); break;
  }
#undef TLB_TAIL
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return fail("tlb200: tail-split reduce launch: %s", cudaGetErrorString(e));
  return 0;
}

// Copies a row-major [rows, cols] operand to a 128-byte-multiple row pitch so that TMA can address it.
template <typename T>
__global__ void repitch_kernel(const T* __restrict__ src, long long ld_src, T* __restrict__ dst, long long ld_dst,
                               long long rows, long long cols) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols, c = i - r * cols;
    dst[r * ld_dst + c] = src[r * ld_src + c];
  }
}

// dst[c, r] = src[r, c] for 4-byte elements, 32x32 tiles through shared memory (both sides coalesced).
__global__ void transpose32_kernel(const uint32_t* __restrict__ src, long long ld_src, uint32_t* __restrict__ dst,
                                   long long ld_dst, long long rows, long long cols) {
  __shared__ uint32_t tile[32][33];
  const long long tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  for (long long t = blockIdx.x; t < tiles_r * tiles_c; t += gridDim.x) {
    const long long r0 = (t / tiles_c) * 32, c0 = (t % tiles_c) * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y)
      if (r0 + i < rows && c0 + threadIdx.x < cols) tile[i][threadIdx.x] = src[(r0 + i) * ld_src + c0 + threadIdx.x];
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y)
      if (c0 + i < cols && r0 + threadIdx.x < rows) dst[(c0 + i) * ld_dst + r0 + threadIdx.x] = tile[threadIdx.x][i];
    __syncthreads();
  }
}

// rows/cols describe the SOURCE as stored; with `transpose` the copy is [cols, rows].
int launch_repitch(const void* src, long long ld_src, void* dst, long long ld_dst, long long rows, long long cols, int es,
                   bool transpose, cudaStream_t stream) {
  long long blocks = (rows * cols + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  if (transpose) {
    if (es != 4) return fail("tlb200: transposing copy is only used for 4-byte operands");
    long long tb = ((rows + 31) / 32) * ((cols + 31) / 32);
    if (tb > 148 * 16) tb = 148 * 16;
    transpose32_kernel<<<static_cast<int>(tb), dim3(32, 8), 0, stream>>>(static_cast<const uint32_t*>(src), ld_src,
                                                                        static_cast<uint32_t*>(dst), ld_dst, rows, cols);
  } else if (es == 4)
    repitch_kernel<uint32_t><<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const uint32_t*>(src), ld_src,
                                                                          static_cast<uint32_t*>(dst), ld_dst, rows, cols);
  else if (es == 2)
    repitch_kernel<uint16_t><<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const uint16_t*>(src), ld_src,
                                                                          static_cast<uint16_t*>(dst), ld_dst, rows, cols);
  else
    repitch_kernel<uint8_t><<<static_cast<int>(blocks), 256, 0, stream>>>(static_cast<const uint8_t*>(src), ld_src,
                                                                        static_cast<uint8_t*>(dst), ld_dst, rows, cols);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail("tlb200: operand re-pitch launch: %s", cudaGetErrorString(e));
  return 0;
}

GemmKernelFn pick_kernel(int cg, int kind, int n_sub) {
  if (cg == 2) return n_sub == 2 ? kernel_cg2_ns2(kind) : kernel_cg2_ns1(kind);
  return n_sub == 2 ? kernel_cg1_ns2(kind) : kernel_cg1_ns1(kind);
}
GemmKernelFn pick_kernel_mc(int kind) { return kernel_mc(kind); }
std::array<const void*, 20> all_kernels() {
  std::array<const void*, 20> k;
  int i = 0;
  for (int cg = 1; cg <= 2; ++cg)
    for (int kind = 0; kind < 4; ++kind)
      for (int ns = 1; ns <= 2; ++ns) k[i++] = reinterpret_cast<const void*>(pick_kernel(cg, kind, ns));
  for (int kind = 0; kind < 4; ++kind) k[i++] = reinterpret_cast<const void*>(pick_kernel_mc(kind));
  return k;
}

BlockScaledFn pick_blockscaled(int cg, int bn) { return kernel_blockscaled(cg, bn); }
constexpr int kChainSmemFixed = 1024 + (2 * kMaxStages + 8) * 8 + 16;

int out_kind_of(int dt) {
  switch (dt) {
    case TLB200_FLOAT16: return kOutF16;
    case TLB200_BFLOAT16: return kOutBF16;
    case TLB200_FLOAT8_E4M3: return kOutE4M3;
    case TLB200_FLOAT8_E5M2: return kOutE5M2;
    case TLB200_INT32: return kOutI32;
    default: return kOutF32;
  }
}

// TMA multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's
// memory hung the GPU.  The A-multicast plan is therefore used only when this predicate PROVES that every byte the
// tensor map of A can address is ordinary memory of device `dev`, allocated by this process with cudaMalloc /
// cuMemAlloc: first and last byte are device memory of `dev`, not managed, not from a stream-ordered pool, not a
// virtual-memory mapping (cuMemMap: symmetric memory, multi-GPU fabrics, expandable segments), exportable by this
// process (an imported IPC mapping is not), and both lie in ONE allocation that contains the whole extent.
// Anything it cannot prove, and any query error, answers "no": the ordinary plan runs instead.
bool own_device_extent(const void* ptr, long long bytes, int dev) {
  if (!ptr || bytes <= 0 || dev < 0) return false;
  if (!g_ptr_attr || !g_addr_range || !g_retain_handle || !g_release_handle) return false;
  const CUdeviceptr lo = reinterpret_cast<CUdeviceptr>(ptr);
  const CUdeviceptr hi = lo + static_cast<CUdeviceptr>(bytes) - 1;
  if (hi < lo) return false;
  unsigned long long id[2] = {0, 0};
  const CUdeviceptr ends[2] = {lo, hi};
  for (int i = 0; i < 2; ++i) {
    unsigned int mtype = 0, managed = 1, ipc_ok = 0;
    int ordinal = -1;
    CUmemoryPool pool = nullptr;
    if (g_ptr_attr(&mtype, CU_POINTER_ATTRIBUTE_MEMORY_TYPE, ends[i]) != CUDA_SUCCESS || mtype != CU_MEMORYTYPE_DEVICE) return false;
    if (g_ptr_attr(&ordinal, CU_POINTER_ATTRIBUTE_DEVICE_ORDINAL, ends[i]) != CUDA_SUCCESS || ordinal != dev) return false;
    if (g_ptr_attr(&managed, CU_POINTER_ATTRIBUTE_IS_MANAGED, ends[i]) != CUDA_SUCCESS || managed) return false;
    if (g_ptr_attr(&pool, CU_POINTER_ATTRIBUTE_MEMPOOL_HANDLE, ends[i]) != CUDA_SUCCESS || pool != nullptr) return false;
    if (g_ptr_attr(&ipc_ok, CU_POINTER_ATTRIBUTE_IS_LEGACY_CUDA_IPC_CAPABLE, ends[i]) != CUDA_SUCCESS || !ipc_ok) return false;
    if (g_ptr_attr(&id[i], CU_POINTER_ATTRIBUTE_BUFFER_ID, ends[i]) != CUDA_SUCCESS) return false;
    CUmemGenericAllocationHandle h = 0;
    if (g_retain_handle(&h, reinterpret_cast<void*>(ends[i])) == CUDA_SUCCESS) {  // a cuMemMap'd range: never trusted
      g_release_handle(h);
      return false;
    }
  }
  if (id[0] != id[1]) return false;
  CUdeviceptr base = 0;
  size_t size = 0;
  if (g_addr_range(&base, &size, lo) != CUDA_SUCCESS) return false;
  if (base > lo || static_cast<unsigned long long>(hi - base) >= static_cast<unsigned long long>(size)) return false;
  return true;
}

int current_device_ready(int* dev_out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail("tlb200: cudaGetDevice: %s", cudaGetErrorString(e));
  if (dev < 0 || dev >= kMaxDevices) return fail("tlb200: device ordinal %d out of range", dev);
  if (!g_dev_ready[dev].load()) {
    int rc = tlb200_init(dev);
    if (rc) return rc;
  }
  *dev_out = dev;
  return 0;
}

// The plan one call will run: resolve(), the fall-back to an unsplit schedule when an automatic split-K does not fit
// the workspace, and - after the LAST resolve, whatever path led to it - the proof that an A-multicast plan reads
// this GPU's own memory only (own_device_extent).  Without the proof the ordinary plan is resolved instead; an
// explicit cfg_cta_group = 4 is an error.  Shared by gemm_impl and tlb200_gemm_plan_query_call.
int plan_call(const tlb200_gemm_desc* d, int dev, const void* a, const void* b, const void* c, const void* ws,
              int64_t ws_bytes, int n_peer, int n_groups, Resolved* r) {
  // never with grouped row tiles (not paired across column tiles) nor on the multi-GPU entry point
  // the pair tiles of 128 rows: any single-destination contraction may ask for them, the automatic plan takes them for
  // ordinary (not grouped) ones only
  const bool plain = n_peer == 0 && (n_groups == 0 || d->cfg_cta_group == 3);
  bool allow_mc = n_groups == 0 && n_peer == 0 && !g_no_multicast;
  const int sms = g_sm_count[dev].load();
  for (int attempt = 0; attempt < 2; ++attempt) {
    if (int rc = resolve(d, sms, c, r, ws != nullptr, a, b, allow_mc, true, plain)) return rc;
    if (r->split_k > 1 && !ws) return fail("tlb200_gemm: cfg_split_k=%d needs a workspace: call tlb200_gemm_ex", d->reserved[1]);
    if (r->ws_bytes > 0) {
      if (reinterpret_cast<uintptr_t>(ws) % 256) return fail("tlb200_gemm_ex: workspace must be 256-byte aligned");
      if (ws_bytes < r->ws_bytes && r->split_k > 1 && d->reserved[1] <= 1) {
        tlb200_gemm_desc d1 = *d;  // automatic split does not fit the workspace: run unsplit
        d1.reserved[1] = 1;
        if (int rc = resolve(&d1, sms, c, r, true, a, b, allow_mc, true, plain)) return rc;
      }
      if (ws_bytes < r->ws_bytes)
        return fail("tlb200_gemm_ex: this call needs a %lld-byte workspace (%lld for re-pitched operands, %lld for split-K x%d), got %lld",
                    r->ws_bytes, r->a_copy_bytes + r->b_copy_bytes, r->split_bytes, r->split_k, (long long)ws_bytes);
    }
    if (r->mc != 2) return 0;
    // What the multicast loads would read: the caller's A, or its re-pitched copy at the start of the workspace.
    // The extent is everything the tensor map [m rows x k columns, pitch lda] can address.
    const void* a_src = r->copy_a ? ws : a;
    const long long a_ld = r->copy_a ? r->lda_c : d->lda;
    const long long a_extent = ((static_cast<long long>(d->m) - 1) * a_ld + d->k) * r->es;
    if (allow_mc && !r->a_mn && own_device_extent(a_src, a_extent, dev)) return 0;
    if (d->cfg_cta_group == 4)
      return fail("tlb200_gemm: cfg_cta_group=4 (A multicast) needs all of A in one cudaMalloc allocation of the current "
                  "device (not managed, pooled, virtual-memory-mapped, imported or peer memory)");
    allow_mc = false;  // resolve again: no multicast plan can come out of it
  }
  return fail("tlb200_gemm: internal error: plan still multicast after it was ruled out");
}

}  // namespace

extern "C" {

const char* tlb200_last_error(void) { return g_err; }
int tlb200_abi_version(void) { return TLB200_ABI_VERSION; }
int64_t tlb200_launch_count(void) { return g_launches.load(); }

uint32_t tlb200_instr_desc(int atom_m, int atom_n, int atom_k, int ab_dtype, int c_dtype,
                           int a_is_k_major, int b_is_k_major, int scale_in_a, int scale_in_b) {
  uint32_t d = instr_desc_impl(atom_m, atom_n, atom_k, ab_dtype, c_dtype, a_is_k_major != 0,
                               b_is_k_major != 0, scale_in_a, scale_in_b);
  if (d == 0) fail("tlb200_instr_desc: unsupported arguments");
  return d;
}

int tlb200_tcgen5_meta(int M, int N, int K, int ab, int c, int32_t* out5) {
  // Decision table of src/op/tcgen5_meta.h:21-92, restated.
  auto ok = [&](int am, int an, int ak, int ws, int two) {
    if (out5) { out5[0] = am; out5[1] = an; out5[2] = ak; out5[3] = ws; out5[4] = two; }
    return 1;
  };
  const int ws_ns[3] = {256, 128, 64};
  const bool half_in = ab == TLB200_FLOAT16 || ab == TLB200_BFLOAT16;
  const bool f8_in = ab == TLB200_FLOAT8_E4M3 || ab == TLB200_FLOAT8_E5M2;
  if (half_in && c == TLB200_FLOAT32) {
    if (K % 16) return 0;
    if (M % 128 == 0) {
      for (int an = 256; an >= 16; an -= 16)
        if (N % an == 0) return ok(128, an, 16, 0, 0);
      return 0;
    }
    const int am = (M % 64 == 0) ? 64 : (M % 32 == 0) ? 32 : 0;
    if (!am) return 0;
    for (int an : ws_ns)
      if (N % an == 0) return ok(am, an, 16, 1, 0);
    return 0;
  }
  if (f8_in && (c == TLB200_FLOAT32 || c == TLB200_FLOAT16)) {
    if (K % 32) return 0;
    if (M % 128 == 0) {
      for (int an : ws_ns)
        if (N % an == 0) return ok(128, an, 32, 1, 0);
      for (int an = 256; an >= 16; an -= 16)
        if (N % an == 0) return ok(128, an, 32, 0, 1);
      for (int an = 256; an >= 8; an -= 8)
        if (N % an == 0) return ok(128, an, 32, 0, 0);
      return 0;
    }
    if (M % 64 == 0) {
      for (int an : ws_ns)
        if (N % an == 0) return ok(64, an, 32, 1, 0);
      for (int an = 256; an >= 8; an -= 8)
        if (N % an == 0) return ok(64, an, 32, 0, 0);
      return 0;
    }
    if (M % 32 == 0) {
      for (int an : ws_ns)
        if (N % an == 0) return ok(32, an, 32, 1, 0);
      return 0;
    }
  }
  return 0;
}

uint64_t tlb200_smem_desc(uint32_t smem_addr, int lbo16, int sbo16, int base_offset,
                          int leading_is_absolute, int swizzle_mode) {
  return smem_desc_impl(smem_addr, lbo16, sbo16, base_offset, leading_is_absolute, swizzle_mode);
}

int tlb200_gemm_plan_query(const tlb200_gemm_desc* desc, int sm_count, tlb200_gemm_plan* out) {
  // The plan tlb200_gemm_ex follows when given a workspace.  Without pointers nothing is known about where A lives,
  // so the A-multicast plan is reported only when the caller asks for it by name (cfg_cta_group = 4).
  Resolved r;
  int rc = resolve(desc, sm_count, nullptr, &r, true, nullptr, nullptr, desc && desc->cfg_cta_group == 4 && !g_no_multicast);
  if (rc) return rc;
  if (out) *out = r.plan;
  return 0;
}

int tlb200_gemm_plan_query_call(const tlb200_gemm_desc* desc, const void* a, const void* b, const void* c, const void* ws,
                                int64_t ws_bytes, tlb200_gemm_plan* out) {
  g_err[0] = 0;
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  Resolved r;
  if (int rc = plan_call(desc, dev, a, b, c, ws, ws_bytes, 0, 0, &r)) return rc;
  if (out) *out = r.plan;
  return 0;
}

int tlb200_is_own_device_memory(const void* ptr, int64_t bytes) {
  g_err[0] = 0;
  int dev;
  if (current_device_ready(&dev)) return 0;
  return own_device_extent(ptr, bytes, dev) ? 1 : 0;
}

int tlb200_init(int device) {
  std::lock_guard<std::mutex> lk(g_init_mu);
  g_err[0] = 0;
  if (device < 0 || device >= kMaxDevices) return fail("tlb200_init: device ordinal %d out of range", device);
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail("tlb200_init: no CUDA device (%s); this library has no CPU fallback",
                e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device >= ndev) return fail("tlb200_init: device %d >= device count %d", device, ndev);
  int prev = -1;
  cudaGetDevice(&prev);
  if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("tlb200_init: cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
  int rc = 0;
  do {
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) { rc = fail("tlb200_init: %s", cudaGetErrorString(e)); break; }
    if (prop.major != 10) {
      rc = fail("tlb200_init: device %d is sm_%d%d; this library contains sm_100a code only", device, prop.major, prop.minor);
      break;
    }
    g_sm_count[device].store(prop.multiProcessorCount);
    if (!g_encode) {
      void* fn = nullptr;
      cudaDriverEntryPointQueryResult q;
      e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
      if (e != cudaSuccess || !fn || q != cudaDriverEntryPointSuccess) {
        rc = fail("tlb200_init: cannot bind cuTensorMapEncodeTiled (%s)", cudaGetErrorString(e));
        break;
      }
      g_encode = reinterpret_cast<EncodeTiledFn>(fn);
      // Optional: without them own_device_extent() answers "no" and the A-multicast plan is simply never used.
      auto bind = [](const char* name) -> void* {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint(name, &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) {
          cudaGetLastError();
          return nullptr;
        }
        return f;
      };
      g_ptr_attr = reinterpret_cast<PtrAttrFn>(bind("cuPointerGetAttribute"));
      g_addr_range = reinterpret_cast<AddrRangeFn>(bind("cuMemGetAddressRange"));
      g_retain_handle = reinterpret_cast<RetainHandleFn>(bind("cuMemRetainAllocationHandle"));
      g_release_handle = reinterpret_cast<ReleaseHandleFn>(bind("cuMemRelease"));
      const char* nm = getenv("TLB200_NO_MULTICAST");
      g_no_multicast = nm && nm[0] && nm[0] != '0';
      const char* kgs = getenv("TLB200_KGROUP");
      if (kgs && (kgs[0] == '1' || kgs[0] == '2' || kgs[0] == '4') && !kgs[1]) g_kgroup = kgs[0] - '0';
    }
    // Same step as the generated init(): raise the dynamic shared memory limit (wrapper.py:15-21).
    e = cudaSuccess;
    for (const void* fn : all_kernels()) {
      if ((e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit)) != cudaSuccess) break;
    }
    for (int cg = 1; cg <= 2 && e == cudaSuccess; ++cg)
      for (int bn = 128; bn <= 256 && e == cudaSuccess; bn += 128)
        e = cudaFuncSetAttribute(reinterpret_cast<const void*>(pick_blockscaled(cg, bn)),
                                 cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(reinterpret_cast<const void*>(kernel_chain()), cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemLimit);
    if (e != cudaSuccess) {
      rc = fail("tlb200_init: failed to set the allowed dynamic shared memory size to %d with error: %s", kSmemLimit, cudaGetErrorString(e));
      break;
    }
    {  // how many 4-CTA clusters (two pairs sharing A by multicast) fit this GPU at one CTA per SM
      cudaLaunchConfig_t qc;
      memset(&qc, 0, sizeof(qc));
      qc.gridDim = dim3(prop.multiProcessorCount / 4 * 4, 1, 1);
      qc.blockDim = dim3(256, 1, 1);
      qc.dynamicSmemBytes = kSmemLimit;
      cudaLaunchAttribute qa[1];
      qa[0].id = cudaLaunchAttributeClusterDimension;
      qa[0].val.clusterDim.x = 4; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
      qc.attrs = qa; qc.numAttrs = 1;
      int nclusters = 0;
      if (cudaOccupancyMaxActiveClusters(&nclusters, pick_kernel_mc(kKindF16), &qc) == cudaSuccess && nclusters > 0) {
        g_mc_clusters = nclusters;
        g_mc_clusters_sms = prop.multiProcessorCount;
      } else {
        cudaGetLastError();  // clear: the estimate in mc_clusters_of() is used instead
      }
    }
    HostPipe& hp = g_pipe[device];
    if (!hp.in) {
      bool ok = cudaStreamCreateWithFlags(&hp.in, cudaStreamNonBlocking) == cudaSuccess &&
                cudaStreamCreateWithFlags(&hp.out, cudaStreamNonBlocking) == cudaSuccess &&
                cudaEventCreateWithFlags(&hp.fork, cudaEventDisableTiming) == cudaSuccess &&
                cudaEventCreateWithFlags(&hp.drained, cudaEventDisableTiming) == cudaSuccess;
      for (int i = 0; ok && i < kMaxPanels; ++i)
        ok = cudaEventCreateWithFlags(&hp.loaded[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&hp.done[i], cudaEventDisableTiming) == cudaSuccess;
      if (!ok) { rc = fail("tlb200_init: cannot create copy streams/events: %s", cudaGetErrorString(cudaGetLastError())); break; }
    }
    g_dev_ready[device].store(true);
  } while (0);
  if (prev >= 0 && prev != device) cudaSetDevice(prev);
  return rc;
}

#ifdef TLB200_TRACE
static long long* g_trace = nullptr;
extern "C" void tlb200_debug_set_trace(long long* p) { g_trace = p; }
static int g_debug_mode = 0;
extern "C" void tlb200_debug_set_mode(int mode) { g_debug_mode = mode; }
#endif

static int gemm_impl(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, void* ws, int64_t ws_bytes,
                     void* stream_v, void* const* peers = nullptr, int n_peer = 0, const int32_t* group_rows = nullptr,
                     int n_groups = 0, const tlb200_epilogue_vectors* vec = nullptr) {
  g_err[0] = 0;
  // Epilogue vectors exist only for tlb200_gemm_fused (every other entry point passes none and ignores vec_flags).
  ReduceVectors rv = {nullptr, nullptr, nullptr, 0};
  if (vec && d) {
    if (d->vec_flags & TLB200_VEC_ROW_SCALE) rv.row_scale = vec->row_scale;
    if (d->vec_flags & TLB200_VEC_COL_SCALE) rv.col_scale = vec->col_scale;
    if (d->vec_flags & (TLB200_VEC_BIAS_N | TLB200_VEC_BIAS_M)) rv.bias = vec->bias;
    rv.bias_per_row = (d->vec_flags & TLB200_VEC_BIAS_M) ? 1 : 0;
  }
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  Resolved r;
  if (int rc = plan_call(d, dev, a, b, c, ws, ws_bytes, n_peer, n_groups, &r)) return rc;
  const bool split = r.split_k > 1;
  // The tail split is optional: it is used when the workspace also holds its partial tiles (256-aligned), and never
  // for grouped contractions or multiple destinations.
  const bool tail = r.tail_tiles > 0 && ws && n_groups == 0 && n_peer == 0 && reinterpret_cast<uintptr_t>(ws) % 256 == 0 &&
                    ws_bytes >= r.ws_bytes + r.tail_bytes;
  if (d->m == 0 || d->n == 0) return 0;
  if (!c || (d->k > 0 && (!a || !b))) return fail("tlb200_gemm: null tensor pointer (a=%p b=%p c=%p)", a, b, c);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  const int out_kind = out_kind_of(d->c_dtype);
  cudaError_t e;
  if (d->k == 0) {  // empty contraction: C = 0 (T.clear semantics), in every destination
    if (d->epilogue & TLB200_EPI_ACCUMULATE) return 0;  // C += 0
    if (rv.bias) return fail("tlb200_gemm_fused: k = 0 with a bias vector is not offered (C would be the broadcast bias)");
    const int blocks = 148 * 4;
    for (int g = -1; g < n_peer; ++g) {
      void* dst = g < 0 ? c : peers[g];
      switch (out_kind) {
        case kOutF16: fill_zero_kernel<kOutF16><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;
        case kOutBF16: fill_zero_kernel<kOutBF16><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;
        case kOutF32: fill_zero_kernel<kOutF32><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;
        case kOutE4M3: fill_zero_kernel<kOutE4M3><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;
        case kOutI32: fill_zero_kernel<kOutF32><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;  // same bits
        default: fill_zero_kernel<kOutE5M2><<<blocks, 256, 0, stream>>>(dst, d->ldc, d->m, d->n); break;
      }
      if ((e = cudaGetLastError()) != cudaSuccess) return fail("tlb200_gemm: fill launch: %s", cudaGetErrorString(e));
      g_launches++;
    }
    return 0;
  }

  const tlb200_gemm_plan& pl = r.plan;
  GemmKernelParams kp;
  memset(&kp, 0, sizeof(kp));
  const int bk = r.span;  // K elements per stage
  // Workspace layout: [re-pitched A][re-pitched B][split-K partial sums]
  char* wsb = static_cast<char*>(ws);
  if (r.copy_a) {
    // source shape as stored: [K, M] when transpose_A else [M, K]
    if (int rc = launch_repitch(a, d->lda, wsb, r.lda_c, d->trans_a ? d->k : d->m, d->trans_a ? d->m : d->k, r.es, r.tr_a, stream)) return rc;
    a = wsb;
    g_launches++;
  }
  if (r.copy_b) {
    if (int rc = launch_repitch(b, d->ldb, wsb + r.a_copy_bytes, r.ldb_c, d->trans_b ? d->n : d->k, d->trans_b ? d->k : d->n, r.es, r.tr_b, stream)) return rc;
    b = wsb + r.a_copy_bytes;
    g_launches++;
  }
  if (ws) ws = wsb + r.a_copy_bytes + r.b_copy_bytes;  // from here on: the split-K region
  const long long lda = r.lda_c, ldb = r.ldb_c;
  // A
  if (!r.a_mn) {
    // with A multicast every CTA fetches half of its 128-row box (the other half arrives from the other pair)
    // A problem shorter than one tile (decode-time shapes) fetches only the rows it has, rounded up to the 8-row
    // swizzle atom: the other rows of the stage feed accumulator rows that are never stored, and filling them
    // with zeros would still cost the SM's TMA write bandwidth (the bound for such shapes).
    int a_box_rows = r.half ? kBlockM / 2 : kBlockM / r.mc;
    if (r.short_m) a_box_rows = (d->m + 7) / 8 * 8;
    if (int rc = encode_2d(&kp.tm_a, d->ab_dtype, a, d->m, d->k, lda, bk, a_box_rows, "A[M,K]")) return rc;
    kp.a_boxes = 1; kp.a_box_bytes = a_box_rows * kStageRowBytes;
  } else {
    if (int rc = encode_2d(&kp.tm_a, d->ab_dtype, a, d->k, d->m, lda, r.span, bk, "A[K,M]")) return rc;
    kp.a_boxes = kBlockM / r.span; kp.a_box_bytes = bk * kStageRowBytes;
  }
  // B
  if (!r.b_mn) {
    if (int rc = encode_2d(&kp.tm_b, d->ab_dtype, b, static_cast<long long>(d->n) * (n_groups > 0 ? n_groups : 1), d->k, ldb, bk, r.b_rows_per_cta, "B[N,K]")) return rc;
    kp.b_boxes = 1; kp.b_box_bytes = r.b_rows_per_cta * kStageRowBytes;
  } else {
    if (int rc = encode_2d(&kp.tm_b, d->ab_dtype, b, static_cast<long long>(d->k) * (n_groups > 0 ? n_groups : 1), d->n, ldb, r.span, bk, "B[K,N]")) return rc;
    kp.b_boxes = r.b_rows_per_cta / r.span; kp.b_box_bytes = bk * kStageRowBytes;
  }
  // Split-K: the GEMM kernel stores fp32 partial tiles into the workspace, slice s at rows
  // [s*m_pad, s*m_pad + M) of a [split_k*m_pad, ws_ld] matrix (m_pad = whole tiles, so a slice's
  // out-of-range tile rows stay inside its own region).
  const bool ws_tma = (r.umma_n / (r.epi_warps / 4)) % 32 == 0;
  if (split) {
    if (ws_tma)
      if (int rc = encode_2d(&kp.tm_c, TLB200_FLOAT32, ws, static_cast<long long>(r.split_k) * r.m_pad, d->n, r.ws_ld, 32, 32, "workspace")) return rc;
  } else if (r.epi_tma) {
    if (int rc = encode_2d(&kp.tm_c, d->c_dtype, c, d->m, d->n, d->ldc, 128 / r.ob, 32, "C[M,N]")) return rc;
    for (int g = 0; g < n_peer; ++g)
      if (int rc = encode_2d(&kp.tm_c_peer[g], d->c_dtype, peers[g], d->m, d->n, d->ldc, 128 / r.ob, 32, "peer C[M,N]")) return rc;
  }
  if (tail) {  // region after the re-pitched operands (there is no uniform split-K region in this mode)
    float* tws = reinterpret_cast<float*>(static_cast<char*>(ws) + r.split_bytes);
    if (int rc = encode_2d(&kp.tm_ws, TLB200_FLOAT32, tws, static_cast<long long>(r.tail_split) * r.tail_tiles * pl.block_m, pl.block_n,
                           pl.block_n, 32, 32, "tail workspace")) return rc;
    kp.ws = tws;
    kp.tail_begin = r.tail_begin; kp.tail_tiles = r.tail_tiles; kp.tail_split = r.tail_split; kp.tail_kb = r.tail_kb;
  }
  kp.n_peer = n_peer;
  for (int g = 0; g < n_peer; ++g) kp.c_peer[g] = peers[g];
  kp.c = split ? ws : c;
  kp.ldc = split ? r.ws_ld : d->ldc;
  kp.split_k = r.split_k;
  kp.kb_per_split = r.kb_per_split;
  kp.split_row_pitch = split ? r.m_pad : 0;
  kp.m = d->m; kp.n = d->n;
  kp.num_k_blocks = r.num_k_blocks;
  kp.tiles_m = pl.tiles_m; kp.tiles_n = pl.tiles_n;
  int grid = pl.grid;
  if (n_groups > 0) {  // row tiles are counted per group; a tile never spans two groups
    kp.n_groups = n_groups;
    kp.b_group_rows = r.b_mn ? d->k : d->n;
    int row = 0, tile = 0;
    for (int g = 0; g < n_groups; ++g) {
      kp.group_row[g] = row;
      kp.group_tile[g] = tile;
      row += group_rows[g];
      tile += (group_rows[g] + pl.block_m - 1) / pl.block_m;
    }
    kp.group_row[n_groups] = row;
    kp.group_tile[n_groups] = tile;
    kp.tiles_m = tile;
    if (tile == 0) return 0;
    long long clusters = g_sm_count[dev].load() / pl.cta_group;
    const long long items = static_cast<long long>(tile) * pl.tiles_n;
    if (clusters > items) clusters = items;
    grid = static_cast<int>(clusters < 1 ? 1 : clusters) * pl.cta_group;
  }
  kp.raster = pl.raster;
  {
    // Reciprocals of the work-item decomposition's divisors (GemmKernelParams::magic_*): the high 64 bits of
    // x * ceil(2^64 / d) are x / d for every 32-bit x when d >= 2; 0 stands for d <= 1.
    const int mcf = r.mc == 2 ? 2 : 1;
    const long long tiles_n_c = (kp.tiles_n + mcf - 1) / mcf;
    const long long tiles_mn = kp.tiles_m * tiles_n_c;
    auto magic = [](long long dv) -> unsigned long long {
      return dv <= 1 ? 0ull : ~0ull / static_cast<unsigned long long>(dv) + 1ull;
    };
    kp.magic_tiles_mn = magic(tiles_mn);
    kp.magic_group = magic(kp.raster * tiles_n_c);
    kp.magic_raster = magic(kp.raster);
    kp.magic_last_rows = magic(kp.raster > 0 ? kp.tiles_m % kp.raster : 0);
    kp.magic_tail_tiles = magic(tail ? r.tail_tiles : 0);
  }
  kp.block_n = pl.block_n;
  kp.umma_n = r.umma_n;
  kp.n_sub = r.n_sub;
  kp.acc_stages = r.acc_stages;
  kp.epi_warps = r.epi_warps;
  kp.epi_slabs = r.epi_slabs;
  kp.stages = pl.stages;
  kp.k_elems_per_block = bk;
  kp.a_mn_major = r.a_mn; kp.b_mn_major = r.b_mn;
  kp.b_rows_per_cta = r.b_rows_per_cta;
  kp.b_sub_bytes = r.b_rows_per_cta * kStageRowBytes;
  kp.idesc = pl.instr_desc;
  // Shared-memory matrix descriptors, 128B swizzle (layout_type 2), SBO = 8 rows x 128 B.
  // K-major: LBO is ignored by the hardware for swizzled layouts (reference sets 16 B,
  // tcgen05_macro_generator.py:186-190); K advances 32 B inside the swizzle atom (:312-327).
  // MN-major: LBO = distance between 128 B spans along M/N = one TMA box; K advances by
  // UMMA_K rows of 128 B.
  const uint32_t hi = static_cast<uint32_t>(smem_desc_impl(0, 0, 1024 >> 4, 0, 0, 2) >> 32);
  kp.desc_hi_a = hi; kp.desc_hi_b = hi;
  const uint32_t umma_k = 32 / r.es;
  kp.desc_lbo_a = (r.a_mn ? static_cast<uint32_t>(kp.a_box_bytes >> 4) : 1u) << 16;
  kp.desc_lbo_b = (r.b_mn ? static_cast<uint32_t>(kp.b_box_bytes >> 4) : 1u) << 16;
  kp.kadv_a = r.a_mn ? (umma_k * kStageRowBytes) >> 4 : kUmmaKBytes >> 4;
  kp.kadv_b = r.b_mn ? (umma_k * kStageRowBytes) >> 4 : kUmmaKBytes >> 4;
  kp.out_kind = split ? static_cast<int>(kOutF32) : out_kind;
  kp.acc_int = d->ab_dtype == TLB200_INT8;
  kp.epi_tma = split ? ws_tma : r.epi_tma;
  kp.epi_lsu = (!split && r.epi_lsu) ? 1 : 0;
  kp.relu = (!split && (d->epilogue & TLB200_EPI_RELU)) ? 1 : 0;
  kp.accumulate = (!split && (d->epilogue & TLB200_EPI_ACCUMULATE)) ? 1 : 0;
  kp.alpha = split ? 1.0f : d->alpha;
  kp.l2_policy = d->reserved[2];
  if (!split) {  // with split-K the reduction kernel applies them to the summed slices
    kp.row_scale = rv.row_scale; kp.col_scale = rv.col_scale; kp.bias = rv.bias; kp.bias_per_row = rv.bias_per_row;
  }
  // Narrow tiles (one accumulator, <= 128 columns) are bound by the per-barrier hand-shake of the stage ring, not by
  // the tensor pipe: two K blocks share a barrier pair (four for <= 32 columns) when the ring keeps >= 3 groups.
  kp.half_m = r.half;
  kp.a_stage_bytes = r.a_stage_bytes;
  kp.kgroup = 1;
  if (pl.block_n <= 128 && r.n_sub == 1) {
    int kgw = pl.block_n <= 32 ? 4 : 2;
    if (g_kgroup > 0) kgw = g_kgroup;
    if (d->reserved2[0] > 0) kgw = d->reserved2[0];  // cfg_kgroup
    while (kgw > 1 && pl.stages / kgw < 3) kgw >>= 1;
    kp.kgroup = kgw;
  }
#ifdef TLB200_TRACE
  kp.trace = g_trace;
  kp.debug_mode = g_debug_mode;
#endif

  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(pl.threads, 1, 1);
  cfg.dynamicSmemBytes = pl.smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = pl.cta_group * r.mc;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  // programmatic dependent launch: this grid's set-up may overlap the previous kernel's tail (the kernel waits
  // with griddepcontrol.wait before it touches global memory)
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 2;
  e = cudaLaunchKernelEx(&cfg, r.mc == 2 ? pick_kernel_mc(mma_kind_of(d->ab_dtype)) : pick_kernel(pl.cta_group, mma_kind_of(d->ab_dtype), r.n_sub), kp);
  if (e != cudaSuccess)
    return fail("tlb200_gemm: launch failed: %s (grid=%d cluster=%d smem=%d)", cudaGetErrorString(e), pl.grid,
                pl.cta_group, pl.smem_bytes);
  g_launches++;
  if (tail) {
    if (int rc = launch_tail_reduce(static_cast<const float*>(kp.ws), r, pl, c, d->ldc, d->m, d->n, d->alpha,
                                    (d->epilogue & TLB200_EPI_RELU) != 0, (d->epilogue & TLB200_EPI_ACCUMULATE) != 0, out_kind, stream, rv))
      return rc;
    g_launches++;
  }
  if (split) {
    if (int rc = launch_splitk_reduce(static_cast<const float*>(ws), r.ws_ld, r.m_pad, r.split_k, c, d->ldc, d->m, d->n, d->alpha,
                                      (d->epilogue & TLB200_EPI_RELU) != 0, (d->epilogue & TLB200_EPI_ACCUMULATE) != 0, out_kind, stream, rv))
      return rc;
    g_launches++;
  }
  return 0;
}

int tlb200_gemm_fused(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, const tlb200_epilogue_vectors* vec,
                      void* ws, int64_t ws_bytes, void* stream_v) {
  if (!d) return fail("tlb200_gemm_fused: null descriptor");
  if (ws && ws_bytes < 0) return fail("tlb200_gemm_fused: negative workspace size");
  const uint32_t f = d->vec_flags;
  if (f && !vec) return fail("tlb200_gemm_fused: vec_flags = 0x%x but no vector struct was given", f);
  if ((f & TLB200_VEC_BIAS_N) && (f & TLB200_VEC_BIAS_M)) return fail("tlb200_gemm_fused: a bias is per column or per row, not both");
  if (f && d->c_dtype == TLB200_INT32) return fail("tlb200_gemm_fused: epilogue vectors are not offered for int32 output");
  if (f && (d->epilogue & TLB200_EPI_ACCUMULATE)) return fail("tlb200_gemm_fused: epilogue vectors cannot be combined with TLB200_EPI_ACCUMULATE");
  if (vec) {
    const void* ptrs[3] = {(f & TLB200_VEC_ROW_SCALE) ? vec->row_scale : (const float*)16, (f & TLB200_VEC_COL_SCALE) ? vec->col_scale : (const float*)16,
                           (f & (TLB200_VEC_BIAS_N | TLB200_VEC_BIAS_M)) ? vec->bias : (const float*)16};
    for (const void* p : ptrs)
      if (!p || reinterpret_cast<uintptr_t>(p) % 4) return fail("tlb200_gemm_fused: a vector named in vec_flags is null or not 4-byte aligned");
  }
  return gemm_impl(d, a, b, c, ws, ws_bytes, stream_v, nullptr, 0, nullptr, 0, f ? vec : nullptr);
}

int tlb200_gemm(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, void* stream_v) {
  return gemm_impl(d, a, b, c, nullptr, 0, stream_v);
}

int tlb200_gemm_grouped(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, const int32_t* group_rows,
                        int n_groups, void* stream_v) {
  if (!d) return fail("tlb200_gemm_grouped: null descriptor");
  if (!group_rows || n_groups < 1 || n_groups > kMaxGroups)
    return fail("tlb200_gemm_grouped: n_groups must be in [1,%d] (got %d)", kMaxGroups, n_groups);
  long long total = 0;
  for (int g = 0; g < n_groups; ++g) {
    if (group_rows[g] < 0) return fail("tlb200_gemm_grouped: group %d has a negative row count", g);
    total += group_rows[g];
  }
  if (total != d->m) return fail("tlb200_gemm_grouped: group rows sum to %lld but desc.m = %d", total, d->m);
  if (d->trans_a) return fail("tlb200_gemm_grouped: A must be [sum_rows, K] (transpose_A = 0)");
  if (d->reserved[1] > 1) return fail("tlb200_gemm_grouped: split-K is not offered for grouped contractions");
  tlb200_gemm_desc d1 = *d;
  d1.reserved[1] = 1;
  if (d1.cfg_block_n == 0 && d1.cfg_cta_group == 0) {  // tiles must not waste rows on small groups
    // A row tile never spans two groups, so every group is padded to whole tiles: pairs (256 rows) only when some group is
    // big enough to fill one and the padding they add over 128-row tiles stays under 15 % (a ragged mix of 8 ... 1000
    // rows per group pads 30 % more with pairs: 418 vs 369 us; profiles/r2_notes.md section 20).
    long long biggest = 0, pad256 = 0, pad128 = 0;
    for (int g = 0; g < n_groups; ++g) {
      biggest = group_rows[g] > biggest ? group_rows[g] : biggest;
      pad256 += (group_rows[g] + 2 * kBlockM - 1) / (2 * kBlockM) * (2 * kBlockM);
      pad128 += (group_rows[g] + kBlockM - 1) / kBlockM * kBlockM;
    }
    d1.cfg_cta_group = (biggest > 192 && pad256 * 100 <= pad128 * 115) ? 2 : 1;
  }
  return gemm_impl(&d1, a, b, c, nullptr, 0, stream_v, nullptr, 0, group_rows, n_groups);
}

int tlb200_gemm_bcast(const tlb200_gemm_desc* d, const void* a, const void* b, void* const* c_list, int n_c, void* stream_v) {
  if (!c_list || n_c < 1 || n_c > 1 + kMaxPeerDst) return fail("tlb200_gemm_bcast: n_c must be in [1,%d] (got %d)", 1 + kMaxPeerDst, n_c);
  for (int g = 0; g < n_c; ++g) {
    if (!c_list[g]) return fail("tlb200_gemm_bcast: destination %d is null", g);
    // one epilogue mode serves every destination: they must agree on 16-byte alignment
    if ((reinterpret_cast<uintptr_t>(c_list[g]) % 16) != (reinterpret_cast<uintptr_t>(c_list[0]) % 16))
      return fail("tlb200_gemm_bcast: destination %d is aligned differently from destination 0", g);
  }
  if (d && d->reserved[1] > 1) return fail("tlb200_gemm_bcast: split-K cannot be combined with multiple destinations");
  if (d && d->cfg_cta_group == 3 && n_c > 1) return fail("tlb200_gemm_bcast: cfg_cta_group=3 is not offered with multiple destinations");
  if (d && (d->epilogue & TLB200_EPI_ACCUMULATE)) return fail("tlb200_gemm_bcast: TLB200_EPI_ACCUMULATE is not offered with multiple destinations");
  // Stores into peer memory drain at NVLink speed, so a tile's epilogue takes several times longer than on one GPU.
  // A 512-wide tile fills TMEM (single-buffered accumulator) and would expose all of that to the tensor pipe: the
  // automatic plan is therefore restricted to double-buffered tiles (<= 256 columns) when there are peer destinations.
  tlb200_gemm_desc d1;
  if (d && n_c > 1 && d->cfg_block_n == 0 && d->m > kBlockM && d->n >= 256 && dtype_bytes(d->ab_dtype) == 2) {
    d1 = *d;
    d1.cfg_block_n = 256;
    d = &d1;
  }
  return gemm_impl(d, a, b, c_list[0], nullptr, 0, stream_v, c_list + 1, n_c - 1);
}

int tlb200_gemm_ex(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, void* ws, int64_t ws_bytes,
                   void* stream_v) {
  if (ws && ws_bytes < 0) return fail("tlb200_gemm_ex: negative workspace size");
  return gemm_impl(d, a, b, c, ws, ws_bytes, stream_v);
}

// ---------------------------------------------------------------------------------------------
// fp8 contraction with per-(row, 128-K) x per-(128-N, 128-K) scales (gemm_blockscaled_sm100.cuh).
int tlb200_gemm_blockscaled(const tlb200_blockscaled_desc* d, const void* a, const void* b, void* c,
                            const float* scales_a, const float* scales_b, void* stream_v) {
  g_err[0] = 0;
  if (!d) return fail("tlb200_gemm_blockscaled: null descriptor");
  if (d->abi_version != TLB200_ABI_VERSION)
    return fail("tlb200_gemm_blockscaled: descriptor abi_version %d != library %d", d->abi_version, TLB200_ABI_VERSION);
  if (d->m < 0 || d->n < 0 || d->k < 0) return fail("tlb200_gemm_blockscaled: negative shape %dx%dx%d", d->m, d->n, d->k);
  if (d->ab_dtype != TLB200_FLOAT8_E4M3 && d->ab_dtype != TLB200_FLOAT8_E5M2)
    return fail("tlb200_gemm_blockscaled: operands must be float8_e4m3 or float8_e5m2 (got dtype %d)", d->ab_dtype);
  if (d->c_dtype != TLB200_BFLOAT16 && d->c_dtype != TLB200_FLOAT16 && d->c_dtype != TLB200_FLOAT32)
    return fail("tlb200_gemm_blockscaled: output must be bfloat16, float16 or float32 (got dtype %d)", d->c_dtype);
  if (d->group_n != 128 || d->group_k != 128)
    return fail("tlb200_gemm_blockscaled: scale groups must be 128 x 128 (got group_n=%d group_k=%d)", d->group_n, d->group_k);
  if (d->m == 0 || d->n == 0) return 0;
  if (!c) return fail("tlb200_gemm_blockscaled: null C");
  if (d->ldc < d->n) return fail("tlb200_gemm_blockscaled: ldc %lld < n %d", (long long)d->ldc, d->n);
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  const int ob = dtype_bytes(d->c_dtype);
  const int out_kind = out_kind_of(d->c_dtype);
  if (d->k == 0) {  // empty sum
    cudaError_t e = cudaMemset2DAsync(c, static_cast<size_t>(d->ldc) * ob, 0, static_cast<size_t>(d->n) * ob, d->m, stream);
    if (e != cudaSuccess) return fail("tlb200_gemm_blockscaled: clearing C: %s", cudaGetErrorString(e));
    return 0;
  }
  if (!a || !b || !scales_a || !scales_b) return fail("tlb200_gemm_blockscaled: null operand or scale pointer");
  if (d->lda < d->k || d->ldb < d->k)
    return fail("tlb200_gemm_blockscaled: lda %lld / ldb %lld smaller than k %d", (long long)d->lda, (long long)d->ldb, d->k);
  // TMA needs 16-byte aligned bases and row pitches (copy.cc:1489-1499); there is no staging copy on this path.
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b)) % 16 || d->lda % 16 || d->ldb % 16)
    return fail("tlb200_gemm_blockscaled: A and B must be 16-byte aligned with row pitches that are multiples of 16 "
                "(lda=%lld ldb=%lld)", (long long)d->lda, (long long)d->ldb);
  if ((reinterpret_cast<uintptr_t>(scales_a) | reinterpret_cast<uintptr_t>(scales_b)) % 4)
    return fail("tlb200_gemm_blockscaled: scale pointers must be 4-byte aligned");

  int cg = d->cfg_cta_group, bn = d->cfg_block_n;
  if (cg != 0 && cg != 1 && cg != 2) return fail("tlb200_gemm_blockscaled: cfg_cta_group %d (0, 1 or 2)", cg);
  if (bn != 0 && bn != 128 && bn != 256) return fail("tlb200_gemm_blockscaled: cfg_block_n %d (0, 128 or 256)", bn);
  const int sms = g_sm_count[dev].load();
  if (bn == 0) {
    // 256-wide tiles halve the A traffic; fall back to 128 when that leaves SMs without a tile
    const long long t256 = static_cast<long long>((d->m + 127) / 128) * ((d->n + 255) / 256);
    bn = (d->n <= 128 || t256 < sms) ? 128 : 256;
  }
  // Pairs (cta_group::2: half the B traffic per CTA) are the default for 256-wide tiles when a pair plan still
  // occupies the machine.  The per-K-block hand-back of the partial-sum buffer crosses the cluster; with the relaxed
  // cluster arrive it costs what the local one does (with the release form pairs ran at half the single-CTA rate:
  // profiles/r1_notes.md, profiles/r2_notes.md).  8192^3: 2478 vs 2191 TFLOP/s back to back.
  if (cg == 0) {
    const long long pair_tiles = static_cast<long long>((d->m + 2 * kBlockM - 1) / (2 * kBlockM)) * ((d->n + bn - 1) / bn);
    cg = (bn == 256 && d->m > kBlockM && pair_tiles >= sms / 2) ? 2 : 1;
  }
  const int block_m = kBlockM * cg;
  const int stage_bytes = kABytesPerStage + (bn / cg) * kStageRowBytes;
  const int fixed = 1024 + kBsEpiWarps * kEpiBufBytes + (2 * kMaxStages + 8) * 8 + 16;
  int stages = d->cfg_stages > 0 ? d->cfg_stages : (kSmemLimit - fixed) / stage_bytes;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2 || fixed + stages * stage_bytes > kSmemLimit)
    return fail("tlb200_gemm_blockscaled: %d stages of %d bytes do not fit %d bytes of shared memory", stages, stage_bytes, kSmemLimit);

  BlockScaledParams kp;
  memset(&kp, 0, sizeof(kp));
  if (int rc = encode_2d(&kp.tm_a, d->ab_dtype, a, d->m, d->k, d->lda, kStageRowBytes, kBlockM, "A")) return rc;
  if (int rc = encode_2d(&kp.tm_b, d->ab_dtype, b, d->n, d->k, d->ldb, kStageRowBytes, bn / cg, "B")) return rc;
  // TMA stores clip at the tensor's last column only in whole 16-byte units: a row whose byte length is not a
  // multiple of 16 takes the predicated-store epilogue so nothing beyond column n is ever written.
  kp.epi_tma = (reinterpret_cast<uintptr_t>(c) % 16 == 0 && (d->ldc * ob) % 16 == 0 &&
                (static_cast<long long>(d->n) * ob) % 16 == 0) ? 1 : 0;
  if (kp.epi_tma)
    if (int rc = encode_2d(&kp.tm_c, d->c_dtype, c, d->m, d->n, d->ldc, 128 / ob, 32, "C")) return rc;
  kp.c = c;
  kp.ldc = d->ldc;
  kp.sa = scales_a; kp.sb = scales_b;
  kp.sa_sm = d->sa_stride_m; kp.sa_sk = d->sa_stride_k;
  kp.sb_sn = d->sb_stride_n; kp.sb_sk = d->sb_stride_k;
  kp.m = d->m; kp.n = d->n;
  kp.num_k_blocks = (d->k + kStageRowBytes - 1) / kStageRowBytes;
  kp.tiles_m = (d->m + block_m - 1) / block_m;
  kp.tiles_n = (d->n + bn - 1) / bn;
  kp.raster = d->cfg_raster > 0 ? d->cfg_raster : 8;
  kp.stages = stages;
  kp.b_rows_per_cta = bn / cg;
  kp.idesc = instr_desc_impl(block_m, bn, 32, d->ab_dtype, TLB200_FLOAT32, true, true, 1, 1);
  kp.desc_hi = static_cast<uint32_t>(smem_desc_impl(0, 0, 1024 >> 4, 0, 0, 2) >> 32);
  kp.out_kind = out_kind;
  kp.l2_policy = 0;
  long long clusters = sms / cg;
  const long long tiles = static_cast<long long>(kp.tiles_m) * kp.tiles_n;
  if (clusters > tiles) clusters = tiles;
  const int grid = static_cast<int>(clusters) * cg;
  const int smem = fixed + stages * stage_bytes;

  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid, 1, 1);
  cfg.blockDim = dim3(kBsThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = cg;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  cudaError_t e = cudaLaunchKernelEx(&cfg, pick_blockscaled(cg, bn), kp);
  if (e != cudaSuccess)
    return fail("tlb200_gemm_blockscaled: launch failed: %s (grid=%d cluster=%d smem=%d)", cudaGetErrorString(e), grid, cg, smem);
  g_launches++;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Two contractions back to back, the intermediate in tensor memory (gemm_chain_sm100.cuh).
int tlb200_gemm_chain(const tlb200_chain_desc* d, const void* a, const void* b1, const void* b2, void* c, void* stream_v) {
  g_err[0] = 0;
  if (!d) return fail("tlb200_gemm_chain: null descriptor");
  if (d->abi_version != TLB200_ABI_VERSION)
    return fail("tlb200_gemm_chain: descriptor abi_version %d != library %d", d->abi_version, TLB200_ABI_VERSION);
  if (d->reserved[0] || d->reserved[1] || d->reserved[2]) return fail("tlb200_gemm_chain: descriptor field reserved must be zero");
  if (d->m < 0 || d->k < 0 || d->n1 < 0 || d->n2 < 0) return fail("tlb200_gemm_chain: negative extent");
  if (d->ab_dtype != TLB200_FLOAT16 && d->ab_dtype != TLB200_BFLOAT16)
    return fail("tlb200_gemm_chain: operands must be float16 or bfloat16 (got dtype %d)", d->ab_dtype);
  if (d->c_dtype != TLB200_FLOAT16 && d->c_dtype != TLB200_BFLOAT16 && d->c_dtype != TLB200_FLOAT32)
    return fail("tlb200_gemm_chain: output must be float16, bfloat16 or float32 (got dtype %d)", d->c_dtype);
  if (d->n1 < 64 || d->n1 > 256 || d->n1 % 64)
    return fail("tlb200_gemm_chain: n1 = %d; the intermediate width must be 64, 128, 192 or 256 (it lives in tensor memory)", d->n1);
  if (d->lda < d->k || d->ldb1 < d->k || d->ldb2 < d->n1 || d->ldc < d->n2)
    return fail("tlb200_gemm_chain: a row pitch is smaller than its row (lda=%lld ldb1=%lld ldb2=%lld ldc=%lld)", (long long)d->lda,
                (long long)d->ldb1, (long long)d->ldb2, (long long)d->ldc);
  if (d->m == 0 || d->n2 == 0) return 0;
  if (!a || !b1 || !b2 || !c) return fail("tlb200_gemm_chain: null tensor pointer");
  if ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b1) | reinterpret_cast<uintptr_t>(b2)) % 16 ||
      (d->lda * 2) % 16 || (d->ldb1 * 2) % 16 || (d->ldb2 * 2) % 16)
    return fail("tlb200_gemm_chain: A, B1, B2 must be 16-byte aligned with row pitches that are multiples of 16 bytes");
  if (d->k == 0) return fail("tlb200_gemm_chain: k = 0 is not offered (the result would be all zeros)");
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  ChainParams kp;
  memset(&kp, 0, sizeof(kp));
  if (int rc = encode_2d(&kp.tm_a, d->ab_dtype, a, d->m, d->k, d->lda, 64, kBlockM, "A[M,K]")) return rc;
  if (int rc = encode_2d(&kp.tm_b1, d->ab_dtype, b1, d->n1, d->k, d->ldb1, 64, d->n1, "B1[N1,K]")) return rc;
  if (int rc = encode_2d(&kp.tm_b2, d->ab_dtype, b2, d->n2, d->n1, d->ldb2, 64, kChainOWidth, "B2[N2,N1]")) return rc;
  kp.c = c;
  kp.ldc = d->ldc;
  kp.m = d->m; kp.n1 = d->n1; kp.n2 = d->n2;
  kp.kb1 = (d->k + 63) / 64;
  kp.kb2 = d->n1 / 64;
  kp.tiles_m = (d->m + kBlockM - 1) / kBlockM;
  kp.tiles_n2 = (d->n2 + kChainOWidth - 1) / kChainOWidth;
  const int max_stages = (kSmemLimit - kChainSmemFixed) / kChainStageBytes;
  kp.stages = d->cfg_stages > 0 ? d->cfg_stages : max_stages;
  if (kp.stages < 2 || kp.stages > max_stages || kp.stages > kMaxStages)
    return fail("tlb200_gemm_chain: cfg_stages=%d out of range [2,%d]", kp.stages, max_stages);
  kp.idesc1 = instr_desc_impl(kBlockM, d->n1, 16, d->ab_dtype, TLB200_FLOAT32, true, true, 1, 1);
  kp.idesc2 = instr_desc_impl(kBlockM, kChainOWidth, 16, d->ab_dtype, TLB200_FLOAT32, true, true, 1, 1);
  if (!kp.idesc1 || !kp.idesc2) return fail("tlb200_gemm_chain: cannot encode the instruction descriptors");
  kp.desc_hi = static_cast<uint32_t>(smem_desc_impl(0, 0, 1024 >> 4, 0, 0, 2) >> 32);
  kp.out_kind = out_kind_of(d->c_dtype);
  kp.mid_bf16 = d->ab_dtype == TLB200_BFLOAT16;
  kp.relu_mid = d->relu_mid != 0;
  kp.alpha1 = d->alpha1; kp.alpha2 = d->alpha2;
  int grid = g_sm_count[dev].load();
  if (grid > kp.tiles_m) grid = kp.tiles_m;
  const int smem = kChainSmemFixed + kp.stages * kChainStageBytes;
  ChainFn chain = kernel_chain();
  chain<<<grid, kChainThreads, smem, static_cast<cudaStream_t>(stream_v)>>>(kp);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail("tlb200_gemm_chain: launch failed: %s (grid=%d smem=%d)", cudaGetErrorString(e), grid, smem);
  g_launches++;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free split-K: the partial sums live in an array that is part of this library's device
// image.  `g_static_scratch` below is only a host-side handle for that array: it is not a device
// address and must never be handed to a kernel.  The real address differs per GPU and is obtained
// once per device with cudaGetSymbolAddress while that device is current.
static __device__ __align__(256) unsigned char g_static_scratch[TLB200_STATIC_SCRATCH_BYTES];
static void* g_static_scratch_ptr[kMaxDevices] = {nullptr};
static std::mutex g_static_scratch_mu;

int tlb200_gemm_static_scratch(const tlb200_gemm_desc* d, const void* a, const void* b, void* c, void* stream_v) {
  g_err[0] = 0;
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  void* scratch = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_scratch_mu);
    if (!g_static_scratch_ptr[dev]) {  // first call on this device: resolve the array's device address
      void* p = nullptr;
      cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
      if (e != cudaSuccess || !p)
        return fail("tlb200_gemm_static_scratch: cannot resolve the static scratch array on device %d: %s", dev,
                    cudaGetErrorString(e));
      g_static_scratch_ptr[dev] = p;
    }
    scratch = g_static_scratch_ptr[dev];
  }
  // Size check on the plan this very call would run (same pointers, same fall-backs as gemm_impl: an AUTOMATIC
  // split-K that does not fit simply runs unsplit).  Only scratch the caller asked for by name - a forced
  // cfg_split_k, or operands that must be re-pitched - can exceed the array, and that is an error here.
  Resolved r;
  if (plan_call(d, dev, a, b, c, scratch, static_cast<int64_t>(TLB200_STATIC_SCRATCH_BYTES), 0, 0, &r)) {
    Resolved want;  // what it would have needed, for the message
    if (resolve(d, g_sm_count[dev].load(), c, &want, true, a, b, false) == 0 &&
        want.ws_bytes > static_cast<long long>(TLB200_STATIC_SCRATCH_BYTES))
      return fail("tlb200_gemm_static_scratch: this plan (split-K x%d of %dx%d) needs %lld bytes of scratch but the static "
                  "scratch array holds %u; lower cfg_split_k or call tlb200_gemm_ex with a workspace of your own",
                  want.split_k, d->m, d->n, want.ws_bytes, static_cast<unsigned>(TLB200_STATIC_SCRATCH_BYTES));
    return -1;  // plan_call's own message stands
  }
  return gemm_impl(d, a, b, c, scratch, static_cast<int64_t>(TLB200_STATIC_SCRATCH_BYTES), stream_v);
}

int64_t tlb200_gemm_workspace_bytes(const tlb200_gemm_desc* d) {
  int dev = -1;
  cudaGetDevice(&dev);
  const int sms = (dev >= 0 && dev < kMaxDevices && g_sm_count[dev].load() > 0) ? g_sm_count[dev].load() : 148;
  // Whether a call takes the A-multicast plan depends on where its A lives: size for the larger of the two plans.
  // The tail part is optional: a smaller workspace simply runs without the tail split.
  Resolved r;
  long long need = -1;
  for (int with_mc = 0; with_mc < 2; ++with_mc) {
    if (with_mc ? g_no_multicast : (d && d->cfg_cta_group == 4)) continue;  // a plan family this descriptor cannot take
    if (resolve(d, sms, nullptr, &r, true, nullptr, nullptr, with_mc != 0) == 0 && r.ws_bytes + r.tail_bytes > need)
      need = r.ws_bytes + r.tail_bytes;
  }
  if (need < 0) return -1;  // the error text of the last resolve stands
  g_err[0] = 0;
  return need;
}

static long long extent_bytes(long long rows, long long cols, long long ld, int eb) {
  if (rows <= 0 || cols <= 0) return 0;
  return ((rows - 1) * ld + cols) * eb;
}
static long long round256(long long x) { return (x + 255) & ~255ll; }
// Row pitch (elements) of an operand's device copy in the host-buffer path: the host pitch when TMA can address it,
// else the row length rounded up to 128 bytes (the copy to the device is then a pitched 2-D copy).
static long long host_path_pitch(long long ld, long long cols, int es) {
  return (ld * es) % 16 == 0 ? ld : (cols * es + 127) / 128 * 128 / es;
}

int64_t tlb200_gemm_host_workspace_bytes(const tlb200_gemm_desc* d) {
  if (!d) return -1;
  const int es = dtype_bytes(d->ab_dtype), ob = dtype_bytes(d->c_dtype);
  if (!es || !ob) return -1;
  const long long a_rows = d->trans_a ? d->k : d->m, a_cols = d->trans_a ? d->m : d->k;
  const long long b_rows = d->trans_b ? d->n : d->k, b_cols = d->trans_b ? d->k : d->n;
  const long long a = extent_bytes(a_rows, a_cols, host_path_pitch(d->lda, a_cols, es), es);
  const long long b = extent_bytes(b_rows, b_cols, host_path_pitch(d->ldb, b_cols, es), es);
  const long long c = extent_bytes(d->m, d->n, d->ldc, ob);
  return round256(a) + round256(b) + round256(c);
}

int tlb200_gemm_host(const tlb200_gemm_desc* d, const void* ha, const void* hb, void* hc, void* ws,
                     int64_t ws_bytes, void* stream_v) {
  g_err[0] = 0;
  int dev;
  if (int rc = current_device_ready(&dev)) return rc;
  const int64_t need = tlb200_gemm_host_workspace_bytes(d);
  if (need < 0) return fail("tlb200_gemm_host: bad descriptor");
  if (ws_bytes < need || (!ws && need > 0)) return fail("tlb200_gemm_host: workspace %lld B < required %lld B", (long long)ws_bytes, (long long)need);
  if (reinterpret_cast<uintptr_t>(ws) % 256) return fail("tlb200_gemm_host: workspace must be 256-byte aligned");
  if (d->m == 0 || d->n == 0) return 0;
  if (d->epilogue & TLB200_EPI_ACCUMULATE)
    return fail("tlb200_gemm_host: TLB200_EPI_ACCUMULATE is not offered with host buffers (C is write-only on this path)");
  if (!hc || (d->k > 0 && (!ha || !hb))) return fail("tlb200_gemm_host: null host pointer");
  const int es = dtype_bytes(d->ab_dtype), ob = dtype_bytes(d->c_dtype);
  const long long a_rows = d->trans_a ? d->k : d->m, a_cols = d->trans_a ? d->m : d->k;
  const long long b_rows = d->trans_b ? d->n : d->k, b_cols = d->trans_b ? d->k : d->n;
  const long long lda_d = host_path_pitch(d->lda, a_cols, es), ldb_d = host_path_pitch(d->ldb, b_cols, es);
  const long long ab = extent_bytes(a_rows, a_cols, lda_d, es);
  const long long bb = extent_bytes(b_rows, b_cols, ldb_d, es);
  char* da = static_cast<char*>(ws);
  char* db = da + round256(ab);
  char* dc = db + round256(bb);
  cudaStream_t stream = static_cast<cudaStream_t>(stream_v);
  HostPipe& hp = g_pipe[dev];
  cudaError_t e;
#define TLB_CK(call, what) do { if ((e = (call)) != cudaSuccess) return fail("tlb200_gemm_host: %s: %s", what, cudaGetErrorString(e)); } while (0)
  // Everything is ordered after the caller's stream.
  TLB_CK(cudaEventRecord(hp.fork, stream), "fork");
  TLB_CK(cudaStreamWaitEvent(hp.in, hp.fork, 0), "fork");
  TLB_CK(cudaStreamWaitEvent(hp.out, hp.fork, 0), "fork");
  // Shrinking panels: about a third of what remains, between 256 and 2048 rows (8192 rows go as 2048, 2048, 1408, 896,
  // 640, 384, 256, 256, 256), so that what can only start when the last input byte has arrived is small.
  auto cut = [](int total, int* rows) {
    int n = 0, rem = total;
    while (rem > 0 && n < kMaxPanels / 2 - 1) {
      int r = (rem / 3 + 127) / 128 * 128;
      r = r < 256 ? 256 : (r > 2048 ? 2048 : r);
      if (r > rem) r = rem;
      rows[n++] = r;
      rem -= r;
    }
    if (rem > 0) rows[n++] = rem;
    return n;
  };
  auto h2d_rows = [&](char* dst, long long ld_d, const void* src, long long ld_h, long long r0, long long rows, long long cols,
                      const char* what) -> cudaError_t {
    (void)what;
    const char* hs = static_cast<const char*>(src) + r0 * ld_h * es;
    char* ds = dst + r0 * ld_d * es;
    // an operand whose host pitch TMA cannot address is re-pitched by the copy itself
    if (ld_d == ld_h) return cudaMemcpyAsync(ds, hs, extent_bytes(rows, cols, ld_d, es), cudaMemcpyHostToDevice, hp.in);
    return cudaMemcpy2DAsync(ds, ld_d * es, hs, ld_h * es, cols * es, rows, cudaMemcpyHostToDevice, hp.in);
  };
  int ev = 0;  // arrival number: loaded[ev] / done[ev]
  auto multiply = [&](const tlb200_gemm_desc& pd, long long a_off, long long b_off, long long c_off, long long rows, long long cols) -> int {
    TLB_CK(cudaEventRecord(hp.loaded[ev], hp.in), "record");
    TLB_CK(cudaStreamWaitEvent(stream, hp.loaded[ev], 0), "wait");
    if (int rc = tlb200_gemm(&pd, da + a_off, db + b_off, dc + c_off, stream_v)) return rc;
    TLB_CK(cudaEventRecord(hp.done[ev], stream), "record");
    TLB_CK(cudaStreamWaitEvent(hp.out, hp.done[ev], 0), "wait");
    if (cols == d->n && d->ldc == d->n) {
      TLB_CK(cudaMemcpyAsync(static_cast<char*>(hc) + c_off, dc + c_off, extent_bytes(rows, cols, d->ldc, ob), cudaMemcpyDeviceToHost, hp.out), "D2H C");
    } else {
      TLB_CK(cudaMemcpy2DAsync(static_cast<char*>(hc) + c_off, d->ldc * ob, dc + c_off, d->ldc * ob, cols * ob, rows, cudaMemcpyDeviceToHost, hp.out), "D2H C");
    }
    ++ev;
    return 0;
  };
  if (!d->trans_a && d->trans_b && d->k > 0 && d->m >= 2048 && d->n >= 2048) {
    // A[M,K] and B[N,K], both cut into row panels that arrive ALTERNATELY (A0, B0, A1, B1, ...).  With a fraction t of
    // each operand on the device a fraction t*t of C can be formed, so products - and their way back over PCIe - start
    // after the first pair of panels instead of after all of B: both directions of the link stay busy almost from the
    // start and the copy out never becomes a tail of its own.  Every arrival triggers ONE rectangular product with
    // everything of the other operand that is already there: A_i x B[0:y)^T -> C[rows_i, 0:y), B_j x A[0:x)^T ->
    // C[0:x, cols_j).
    int ra[kMaxPanels / 2], rb[kMaxPanels / 2];
    const int na = cut(d->m, ra), nb = cut(d->n, rb);
    long long x = 0, y = 0;  // rows of A / B on the device
    for (int i = 0, j = 0; i < na || j < nb;) {
      if (i < na && (x <= y || j >= nb)) {  // next A panel
        TLB_CK(h2d_rows(da, lda_d, ha, d->lda, x, ra[i], d->k, "H2D A"), "H2D A");
        if (y > 0) {
          tlb200_gemm_desc pd = *d;
          pd.lda = lda_d; pd.ldb = ldb_d; pd.m = ra[i]; pd.n = static_cast<int>(y);
          if (int rc = multiply(pd, x * lda_d * es, 0, x * d->ldc * ob, ra[i], y)) return rc;
        }
        x += ra[i++];
      } else {                               // next B panel
        TLB_CK(h2d_rows(db, ldb_d, hb, d->ldb, y, rb[j], d->k, "H2D B"), "H2D B");
        if (x > 0) {
          tlb200_gemm_desc pd = *d;
          pd.lda = lda_d; pd.ldb = ldb_d; pd.m = static_cast<int>(x); pd.n = rb[j];
          if (int rc = multiply(pd, 0, y * ldb_d * es, y * ob, x, rb[j])) return rc;
        }
        y += rb[j++];
      }
    }
  } else {
    // Other layouts: B (needed by every panel) goes first, then row panels of A / C when A is [M,K]: panel p+1 is copied
    // in while panel p is multiplied and panel p-1 is copied out.
    if (bb) TLB_CK(h2d_rows(db, ldb_d, hb, d->ldb, 0, b_rows, b_cols, "H2D B"), "H2D B");
    int rows_of[kMaxPanels];
    int panels = 1;
    rows_of[0] = d->m;
    if (!d->trans_a && d->k > 0 && d->m >= 2048) panels = cut(d->m, rows_of);
    long long r0 = 0;
    for (int p = 0; p < panels; r0 += rows_of[p], ++p) {
      tlb200_gemm_desc pd = *d;
      pd.lda = lda_d; pd.ldb = ldb_d;
      if (panels > 1) {
        pd.m = rows_of[p];
        if (d->k > 0) TLB_CK(h2d_rows(da, lda_d, ha, d->lda, r0, rows_of[p], d->k, "H2D A"), "H2D A");
      } else if (d->k > 0 && ab) {
        TLB_CK(h2d_rows(da, lda_d, ha, d->lda, 0, a_rows, a_cols, "H2D A"), "H2D A");
      }
      if (int rc = multiply(pd, panels > 1 ? r0 * lda_d * es : 0, 0, r0 * d->ldc * ob, pd.m, d->n)) return rc;
    }
  }
  TLB_CK(cudaEventRecord(hp.drained, hp.out), "record");
  TLB_CK(cudaStreamWaitEvent(stream, hp.drained, 0), "join");
  TLB_CK(cudaStreamSynchronize(stream), "synchronize");
#undef TLB_CK
  return 0;
}

}  // extern "C"
