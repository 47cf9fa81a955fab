// gt_mxf4.cu -- the `ground-truth` GEMM with 4-bit operands: tcgen05.mma kind::mxf4 (block-scaled E2M1, all scale
// factors 1.0).  Identity: Hamming(q, d) = pop(q) - sum_i (2 q_i - 1) d_i.
//
// Encoding.  E2M1 holds +-{0.5, 1, 2} exactly: the data bit at nibble position u becomes {0.5, 1, 2, 2}, the query bit
// +-{2, 1, 0.5, 0.5}, so every product is +-1 or 0 and the fp32 accumulator holds the integer pop(q) - Hamming exactly
// (|acc| <= 1024).  Both operands use the same physical nibble layout, so the K permutation this implies cancels.  The
// block scale factors (E8M0, one per 32 elements) are all 0x7F = 2^0: sixteen TMEM columns are filled with that byte
// once, and since every byte is equal the scale-factor layout does not matter.  Against the int8 formulation
// (gt_tcgen05.cu): K = 64 per instruction instead of 32, a data row expands to 512 bytes instead of 1024 (five logic
// ops per 32 elements), the query tile takes 64 KB instead of 128 KB.
//
// Structure.  One persistent CTA per SM, 14 warps: 0-3 epilogue (thread = query row; fused top-k, threshold shared
// across work items), 4-11 expansion in two groups (group g writes K-blocks 2g, 2g+1 of every tile, so one group
// stores while the other sits in its proxy fence), 12 MMA issue, 13 loader (one swizzling tensor-map TMA copy per
// data tile into a raw buffer: no thread waits on a global load).  Shared memory: query tile 64 KB + 4 stages x 30 KB
// + raw tile 30 KB.  TMEM: two 128 x 240 fp32 accumulators + the scale columns = 512 columns.  Lists, compaction and
// the exact merge follow gt_tcgen05.cu; DESIGN.md 4.3 has the measurements behind each of these choices.
// That is plan 0.  Two more plans of the same kernel (Plan<> below, HB_GT_PLAN) are exact and measured slower: CTA pairs
// issuing tcgen05.mma.cta_group::2, and four 120-row accumulators.
#include <cuda.h>  // CUtensorMap and the encode entry point's signature (no libcuda link: resolved through the runtime)

#include <algorithm>
#include <climits>
#include <cstdlib>

#include "gt_common.cuh"
#include "hb_internal.h"

#ifndef HB_GT_PROBE_MMAS
#define HB_GT_PROBE_MMAS 4
#endif
#ifndef HB_GT_PLAN_DEFAULT
#define HB_GT_PLAN_DEFAULT 0  // kernel plan unless HB_GT_PLAN is set (Plan<> below)
#endif
#ifndef HB_GT_MERGE_FENCE
#define HB_GT_MERGE_FENCE 1  // 120-row plans: one proxy fence per expansion warp and tile instead of one per stage
#endif
#ifndef HB_GT_BOUND_AHEAD
#define HB_GT_BOUND_AHEAD 1
#endif
#ifndef HB_GT_SUSPEND_NS
#define HB_GT_SUSPEND_NS 20000u  // upper bound of one mbarrier.try_wait suspension
#endif

namespace hb {
namespace {

constexpr int kTileM = 128;   // queries per work item (UMMA M)
constexpr int kTileN = 240;   // largest accumulator tile (UMMA N): 2 x 240 + 16 scale-factor columns <= 512
constexpr uint32_t kSfCol = 496;  // TMEM columns: accumulators from 0 (Plan<>), scale factors at 496..511
constexpr int kKBlocks = 4;   // 1024 bits = 4 K-blocks of 256 E2M1 (one 128-B swizzle atom row each)
constexpr int kEpiWarps = 4, kExpWarps = 8;
constexpr int kThreads = (kEpiWarps + kExpWarps + 2) * 32;  // + the MMA warp + the loader warp
constexpr uint32_t kABytes = kTileM * 512;            // 4 K-blocks x (128 rows x 128 B)
constexpr uint32_t kAKBlockBytes = kTileM * 128;
// the packed rows of one data tile as ONE tensor-map TMA copy delivers them: 128 B per row, 128-byte swizzle (the
// 16-byte piece c of row r sits at piece c ^ (r & 7)), so the same piece of eight consecutive rows lies in eight
// different bank groups.  (One 128-byte bulk copy per row at a padded pitch costs ~54 cycles per copy, 13 k cycles per
// tile; four large unswizzled copies need lane-rotated reads and selects to dodge the conflicts: 34.4 ms.)
constexpr uint32_t kRawPitch = 128;
// Kernel plans (HB_GT_PLAN; 0 is the default, the other two were built, are exact, and measured slower -- DESIGN.md 4.3).
//  0  one CTA, accumulator tiles of 240 rows, two of them in TMEM, a ring of one tile's four stages.
//  1  CTA pair (cta_group::2, a cluster of two on one TPC): M = 256 (each CTA its own 128 queries, accumulators in its
//     own TMEM) and each CTA loads and expands only its HALF of every 240-row tile; half-size stages, so the ring
//     holds two tiles.
//  2  one CTA, accumulator tiles of 120 rows, FOUR of them in TMEM (the MMA can run three tiles ahead of the
//     epilogue instead of one), a ring of two tiles.
template <int PLAN>
struct Plan {
  static constexpr bool pair = PLAN == 1;
  static constexpr uint32_t tile_n = PLAN == 2 ? 120 : kTileN;      // rows per accumulator tile (UMMA N)
  static constexpr uint32_t n_acc = PLAN == 2 ? 4 : 2;              // accumulators in TMEM
  static constexpr uint32_t acc_stride = PLAN == 2 ? 120 : 256;     // TMEM columns between them
  static constexpr uint32_t group = PLAN == 2 ? 40 : 48;            // accumulator columns an epilogue thread reads per step
  static constexpr uint32_t rows = pair ? tile_n / 2 : tile_n;      // data rows this CTA expands per tile
  static constexpr uint32_t rows_per_thread = rows > 128 ? 2 : 1;   // of an expansion group's 128 threads
  static constexpr uint32_t ring_tiles = rows > 128 ? 1 : 2;        // tiles the stage ring holds
  static constexpr uint32_t stages = 4 * ring_tiles;
  static constexpr uint32_t stage_bytes = rows * 128;
  static constexpr uint32_t raw_bytes = rows * kRawPitch;
  static constexpr uint32_t raw_off = kABytes + stages * stage_bytes;
  static constexpr uint32_t bar_off = raw_off + raw_bytes;
  // barriers: full[stages] empty[stages] tfull[n_acc] tempty[n_acc] a_full a_free raw_full raw_free, then the TMEM base slot
  static constexpr uint32_t bar_bytes = 16 * stages + 16 * n_acc + 32;
  static constexpr uint32_t smem_bytes = bar_off + bar_bytes + 16 + 1024;  // + alignment slack
  // instruction descriptor (cute::UMMA::InstrDescriptorBlockScaled layout): A = B = E2M1 (MXF4 format 1), both K-major,
  // scale format E8M0, scale-factor ids 0, N = tile_n, M = 128 (256 over a pair), K = 64
  static constexpr uint32_t idesc = (1u << 7) | (1u << 10) | (1u << 23) | ((tile_n >> 3) << 17) | ((((pair ? 2u : 1u) * kTileM) >> 4) << 24);
  static_assert(stage_bytes % 1024 == 0 && raw_off % 1024 == 0 && bar_off % 8 == 0, "smem layout");
  static_assert(smem_bytes <= 227 * 1024, "shared memory");
  static_assert((n_acc - 1) * acc_stride + tile_n <= kSfCol && tile_n <= acc_stride && tile_n % group == 0 && group > 32 && group <= 48 && group % 8 == 0,
                "TMEM plan");
};

using namespace gt;  // mbarrier / tcgen05 helpers, HB_TMEM_LD32 (gt_common.cuh)
constexpr auto mbar_wait = gt::mbar_wait<HB_GT_SUSPEND_NS>;
constexpr auto mbar_wait_cluster = gt::mbar_wait<HB_GT_SUSPEND_NS, true>;

#ifdef HB_GT_DIAG
// per-CTA phase clocks (diagnostic build only, tools/gt_diag.py): [0] kernel cycles, [1] MMA warp waiting for a free
// accumulator, [2] MMA warp waiting for a full stage, [3] expansion warp 0 waiting for an empty stage, [4] its fences,
// [5] epilogue warp 0 waiting for a full accumulator, [6] its inserts, [7] tiles, [8] expansion waiting for A free
__device__ long long g_gt4_diag[256 * 16];
// trace of CTA 0, tiles [kTr0, kTr0+32): at g_gt4_diag[2400 + 32*(tile-kTr0) + slot]
//   slot 0 MMA passes acc-free wait, 1 MMA issued last stage, 2+w epilogue warp w sees acc full, 6+w warp w releases acc,
//   10+kb expansion warp 0 arrives stage kb, 18+kb MMA passes full wait of stage kb
constexpr uint32_t kTr0 = 3000;
#define HB_TRACE(tile, slot) \
  do { if (blockIdx.x == 0 && (tile) >= kTr0 && (tile) < kTr0 + 32 && (threadIdx.x & 31) == 0) g_gt4_diag[2400 + 32 * ((tile) - kTr0) + (slot)] = clock64(); } while (0)
#define HB_DIAG_T0() const long long dg_t0 = clock64()
#define HB_DIAG_ADD(var) var += clock64() - dg_t0
#else
#define HB_TRACE(tile, slot)
#define HB_DIAG_T0()
#define HB_DIAG_ADD(var)
#endif

#define HB_TMEM_LD16(v, taddr)                                                                                          \
  asm volatile(                                                                                                         \
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "                                                                         \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"                                  \
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),     \
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])                        \
      : "r"(taddr)                                                                                                      \
      : "memory")

#define HB_TMEM_LD8(v, taddr)                                                                                           \
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"                          \
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])       \
               : "r"(taddr)                                                                                           \
               : "memory")

struct GtParams {
  const uint64_t* data;     // n rows x 16 u64
  const uint64_t* queries;  // nq rows x 16 u64
  uint64_t* partial;        // (n_chunks, nq, lcap) packed lists
  uint32_t* gthr;           // per query: smallest k-th best distance any work item has reached so far (bound on the global k-th)
  uint64_t n, key_offset;
  uint32_t nq, k;
  uint32_t lcap;            // slots per row list in `partial`: k (k <= 16, ascending) or 2k (k > 16, unordered append buffer)
  uint32_t n_qtiles, n_chunks, tiles_per_chunk, n_tiles;
};

// k > 16: the row's candidates go to an UNORDERED append buffer of 2k slots in global memory (one store per accepted
// row; the merge kernel needs no order).  When a buffer is about to fill, the whole warp compacts it: entries up to the
// k-th smallest move to the front, the tail is cleared, and the row's threshold becomes the k-th distance.  Amortised
// over the appends in between, against a dependent chain of loads per insert into a sorted list.  `buf` and the result
// are warp-uniform.
// Returns (k-th smallest distance << 32) | entries kept.  The bisection runs on the distance alone, between the
// smallest and largest distance in the buffer (7-10 passes instead of 43 over the packed value), and keeps every entry
// up to and including the k-th distance -- a few more than k when that distance is tied, which is harmless: the row's
// threshold is the distance, and rows arrive in increasing key order.  Only when the ties would leave less than
// max(8, k/4) free slots does it bisect the keys of the tied entries as well and keep exactly k.
__device__ __noinline__ uint64_t gt4_warp_compact(uint64_t* buf, uint32_t cap, uint32_t k) {
  const uint32_t lane = threadIdx.x & 31;
  uint32_t dmin = 0xFFFFFFFFu, dmax = 0;
  for (uint32_t i = lane; i < cap; i += 32) {
    const uint64_t e = buf[i];
    if (e != kNone) {
      dmin = min(dmin, (uint32_t)(e >> 32));
      dmax = max(dmax, (uint32_t)(e >> 32));
    }
  }
  uint32_t lo = __reduce_min_sync(0xFFFFFFFFu, dmin), hi = __reduce_max_sync(0xFFFFFFFFu, dmax);
  while (lo < hi) {  // smallest distance d with count(dist <= d) >= k
    const uint32_t mid = lo + ((hi - lo) >> 1);
    uint32_t c = 0;
    for (uint32_t i = lane; i < cap; i += 32) c += ((uint32_t)(buf[i] >> 32) <= mid) ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (c >= k) hi = mid;
    else lo = mid + 1;
  }
  uint32_t c_le = 0;
  for (uint32_t i = lane; i < cap; i += 32) c_le += ((uint32_t)(buf[i] >> 32) <= lo) ? 1u : 0u;
  c_le = __reduce_add_sync(0xFFFFFFFFu, c_le);
  uint64_t V = ((uint64_t)lo << 32) | 0xFFFFFFFFull;
  if (c_le + max(8u, k / 4) > cap) {  // heavy ties at the k-th distance: keep exactly k
    uint64_t l64 = (uint64_t)lo << 32, h64 = V;
    while (l64 < h64) {
      const uint64_t mid = l64 + ((h64 - l64) >> 1);
      uint32_t c = 0;
      for (uint32_t i = lane; i < cap; i += 32) c += (buf[i] <= mid) ? 1u : 0u;
      c = __reduce_add_sync(0xFFFFFFFFu, c);
      if (c >= k) h64 = mid;
      else l64 = mid + 1;
    }
    V = l64;
  }
  uint32_t woff = 0;
  for (uint32_t base = 0; base < cap; base += 32) {
    const uint32_t i = base + lane;
    const uint64_t e = (i < cap) ? buf[i] : kNone;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, e <= V);
    if (e <= V) buf[woff + __popc(m & ((1u << lane) - 1u))] = e;  // woff + rank <= i: never ahead of the reads
    woff += __popc(m);
    __syncwarp();
  }
  for (uint32_t i = woff + lane; i < cap; i += 32) buf[i] = kNone;
  __syncwarp();
  return ((uint64_t)lo << 32) | woff;
}

// The same for buffers of at most 256 slots (k <= 128), held in registers: one read of the buffer instead of one per pass.
__device__ __noinline__ uint64_t gt4_warp_compact_small(uint64_t* buf, uint32_t cap, uint32_t k) {
  const uint32_t lane = threadIdx.x & 31;
  uint64_t r[8];
  uint32_t dmin = 0xFFFFFFFFu, dmax = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t i = lane + 32 * j;
    r[j] = (i < cap) ? buf[i] : kNone;
    if (r[j] != kNone) {
      dmin = min(dmin, (uint32_t)(r[j] >> 32));
      dmax = max(dmax, (uint32_t)(r[j] >> 32));
    }
  }
  uint32_t lo = __reduce_min_sync(0xFFFFFFFFu, dmin), hi = __reduce_max_sync(0xFFFFFFFFu, dmax);
  while (lo < hi) {
    const uint32_t mid = lo + ((hi - lo) >> 1);
    uint32_t c = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) c += ((uint32_t)(r[j] >> 32) <= mid) ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (c >= k) hi = mid;
    else lo = mid + 1;
  }
  uint32_t c_le = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) c_le += ((uint32_t)(r[j] >> 32) <= lo) ? 1u : 0u;
  c_le = __reduce_add_sync(0xFFFFFFFFu, c_le);
  uint64_t V = ((uint64_t)lo << 32) | 0xFFFFFFFFull;
  if (c_le + max(8u, k / 4) > cap) {
    uint64_t l64 = (uint64_t)lo << 32, h64 = V;
    while (l64 < h64) {
      const uint64_t mid = l64 + ((h64 - l64) >> 1);
      uint32_t c = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) c += (r[j] <= mid) ? 1u : 0u;
      c = __reduce_add_sync(0xFFFFFFFFu, c);
      if (c >= k) h64 = mid;
      else l64 = mid + 1;
    }
    V = l64;
  }
  uint32_t woff = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) {  // every slot was read above, so the writes cannot overtake a read
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, r[j] <= V);
    if (r[j] <= V) buf[woff + __popc(m & ((1u << lane) - 1u))] = r[j];
    woff += __popc(m);
  }
  for (uint32_t i = woff + lane; i < cap; i += 32) buf[i] = kNone;
  __syncwarp();
  return ((uint64_t)lo << 32) | woff;
}

template <bool kRegList, int kPlan>
__global__ void __launch_bounds__(kThreads, 1) gt_mxf4_kernel(const GtParams p, const __grid_constant__ CUtensorMap tmap) {
  using PL = Plan<kPlan>;
  constexpr bool kPair = PL::pair;
  constexpr uint32_t TN = PL::tile_n, GR = PL::group, NACC = PL::n_acc;
  constexpr uint32_t kStages = PL::stages, kStageBytes = PL::stage_bytes;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // swizzle atoms need 1024-B alignment (the same offset in both CTAs of a pair)
  unsigned char* sm = smem_raw + (base - raw);
  const uint32_t sA = base, sB = base + kABytes, sRaw = base + PL::raw_off, sBar = base + PL::bar_off;
  const uint32_t bFull = sBar, bEmpty = sBar + 8 * kStages, bTFull = sBar + 16 * kStages, bTEmpty = bTFull + 8 * NACC, bAFull = bTEmpty + 8 * NACC,
                 bAFree = bAFull + 8, bRawFull = bAFree + 8, bRawFree = bRawFull + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + PL::bar_off + PL::bar_bytes);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // Pair: the leader (cluster rank 0) issues every MMA, so the barriers its MMA warp waits on -- stage full, accumulator
  // free, query tile ready -- live in the leader and take the arrivals of BOTH CTAs (remote arrives from the peer); the
  // barriers the MMA completions signal -- stage empty, accumulator full, query tile free -- exist in both CTAs and the
  // leader's commit is multicast to both.  The raw-tile barriers are private to each CTA.
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;
  constexpr uint32_t kCtas = kPair ? 2u : 1u;
  if (threadIdx.x == 0) {
    for (uint32_t s = 0; s < kStages; ++s) {
      mbar_init(bFull + 8 * s, kCtas * (kExpWarps / 2));
      mbar_init(bEmpty + 8 * s, 1);
    }
    for (uint32_t a = 0; a < NACC; ++a) {
      mbar_init(bTFull + 8 * a, 1);
      mbar_init(bTEmpty + 8 * a, kPair ? kCtas * kEpiWarps : kEpiWarps * 32);
    }
    mbar_init(bAFull, kCtas * kExpWarps);
    mbar_init(bAFree, 1);
    mbar_init(bRawFull, 1);
    mbar_init(bRawFree, kExpWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    if (kPair) {
      asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
    } else {
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  // scale factors: E8M0 0x7F = 2^0 in every byte of TMEM columns 496..511, all 128 lanes; read by every MMA as SFA and SFB
  if (warp < 4) {
    const uint32_t one = 0x7F7F7F7Fu;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(
            tmem + ((warp * 32u) << 16) + kSfCol),
        "r"(one)
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (kPair) cluster_sync_all();  // both CTAs' barriers, TMEM and scale factors are in place before either touches the other
  else __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  // work items: (query tile, chunk of data tiles); a pair takes two adjacent query tiles, one per CTA
  const uint32_t n_qslots = kPair ? (p.n_qtiles + 1) / 2 : p.n_qtiles;
  const uint32_t n_items = n_qslots * p.n_chunks;
  const uint32_t item0 = kPair ? blockIdx.x >> 1 : blockIdx.x, item_step = kPair ? gridDim.x >> 1 : gridDim.x;
  auto qtile_of = [&](uint32_t item) { return kPair ? (item % n_qslots) * 2 + rank : item % n_qslots; };
  // the leader's copies of the barriers both CTAs arrive on (shared::cluster addresses; own CTA when not paired)
  const uint32_t bFullL = kPair ? mapa_u32(bFull, 0) : bFull, bTEmptyL = kPair ? mapa_u32(bTEmpty, 0) : bTEmpty,
                 bAFullL = kPair ? mapa_u32(bAFull, 0) : bAFull;
#if defined(HB_GT_PROBE_STS) && HB_GT_PROBE_STS < 8
  for (uint32_t i = threadIdx.x; i < kStages * kStageBytes / 16; i += blockDim.x)
    asm volatile("st.shared.v4.b32 [%0], {%1,%1,%1,%1};" ::"r"(sB + i * 16), "r"(0u) : "memory");
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
#endif

  if (warp < kEpiWarps) {
    // ===================== epilogue: thread = one query row of the accumulator =====================
    const uint32_t row = warp * 32 + lane;
    uint32_t tcount = 0;
    [[maybe_unused]] long long dg_tfull = 0, dg_ins = 0;
    for (uint32_t item = item0; item < n_items; item += item_step) {
      const uint32_t qtile = qtile_of(item), chunk = item / n_qslots;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      const uint32_t q = qtile * kTileM + row;
      const bool live = q < p.nq;
      uint64_t* list = p.partial + ((size_t)chunk * p.nq + (live ? q : 0)) * p.lcap;
      // k <= 16: the row's ascending list lives in registers, right-aligned (rl[16-k..15]; the slots before it hold 0,
      // which no element displaces), so the k-th best is always rl[15] and every index is static
      uint64_t rl[kRegList ? 16 : 1];
      uint32_t cnt = 0;  // k > 16: filled slots of the row's append buffer
      int popq = 0;
      if (live) {
        const uint4* qp = reinterpret_cast<const uint4*>(p.queries + (size_t)q * 16);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint4 w = __ldg(qp + j);
          popq += __popc(w.x) + __popc(w.y) + __popc(w.z) + __popc(w.w);
        }
        if (!kRegList)
          for (uint32_t i = 0; i < p.lcap; ++i) list[i] = kNone;
      }
      if (kRegList) {
#pragma unroll
        for (int i = 0; i < 16; ++i) rl[i] = ((uint32_t)i + p.k < 16u) ? 0ull : kNone;
      }
      // The accumulator must exceed T to beat this item's current k-th best (strict: keys ascend inside an item), and
      // must exceed Tg to be at most the best k-th distance any item has reached for this query (ties kept: the
      // other item's keys may be larger).  Rows failing either test cannot be in the global top-k.
      float T = live ? -INFINITY : INFINITY;  // accumulators are exact small integers in fp32: pop(q) - Hamming
      uint32_t* gq = p.gthr + (live ? q : 0);
      const uint64_t row_end = min(p.n, (uint64_t)t1 * TN);
      // the shared bound is read one tile ahead: when the epilogue is what the kernel waits for, an accumulator is
      // already full when the loop comes round and the L2 round trip of this load would be paid on every tile
      [[maybe_unused]] uint32_t g_ahead = *reinterpret_cast<volatile uint32_t*>(gq);
      for (uint32_t t = t0; t < t1; ++t, ++tcount) {
        const uint32_t acc = tcount % NACC, use = tcount / NACC;
#if HB_GT_BOUND_AHEAD
        const uint32_t g_now = g_ahead;
#else
        const uint32_t g_now = *reinterpret_cast<volatile uint32_t*>(gq);
#endif
        {
          HB_DIAG_T0();
          mbar_wait(bTFull + 8 * acc, use & 1);
          HB_DIAG_ADD(dg_tfull);
        }
#if HB_GT_BOUND_AHEAD
        g_ahead = *reinterpret_cast<volatile uint32_t*>(gq);
#endif
        HB_TRACE(tcount, 2 + warp);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const float Tg = (g_now == 0xFFFFFFFFu) ? -INFINITY : (float)(popq - (int)g_now) - 1.0f;
        float Te = fmaxf(T, Tg);
        const uint32_t taddr = tmem + ((warp * 32u) << 16) + acc * PL::acc_stride;
        const uint64_t tile_row0 = (uint64_t)t * TN;
#pragma unroll 1
        for (int g = 0; g < (int)(TN / GR); ++g) {
          uint32_t h0 = 0, h1 = 0;
          bool hit;
          float m8[GR / 8];  // maxima of the group's six octets of columns
          {
            uint32_t v[GR];
            HB_TMEM_LD32(v, taddr + g * GR);
            if constexpr (GR == 48) HB_TMEM_LD16((v + 32), taddr + g * GR + 32);
            else HB_TMEM_LD8((v + 32), taddr + g * GR + 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int j = 0; j < GR / 8; ++j) {
              float a = __uint_as_float(v[8 * j]);
#pragma unroll
              for (int c = 1; c < 8; ++c) a = fmaxf(a, __uint_as_float(v[8 * j + c]));
              m8[j] = a;
            }
            float mx = m8[0];
#pragma unroll
            for (int j = 1; j < GR / 8; ++j) mx = fmaxf(mx, m8[j]);
            hit = __any_sync(0xFFFFFFFFu, mx > Te);
            if (kRegList && hit) {  // warp-uniform: which columns pass for which lane
#pragma unroll
              for (int c = 0; c < 32; ++c) h0 |= (uint32_t)(__uint_as_float(v[c]) > Te) << c;
#pragma unroll
              for (int c = 0; c < GR - 32; ++c) h1 |= (uint32_t)(__uint_as_float(v[32 + c]) > Te) << c;
            }
          }
          if (!kRegList && hit) {
            // k > 16: for every octet of columns some lane wants, all lanes re-read their eight values of it from TMEM
            // (the accumulator is still ours) and append the ones that pass to their own buffers, in parallel and with
            // static register indices.  A buffer that could overflow inside the octet is compacted first.
            HB_DIAG_T0();
            const uint64_t r0 = tile_row0 + (uint64_t)g * GR;
            const uint32_t nvalid = (uint32_t)min((uint64_t)GR, row_end > r0 ? row_end - r0 : 0);
            uint32_t fb = 0;
#pragma unroll
            for (int j = 0; j < GR / 8; ++j) fb |= (uint32_t)(m8[j] > Te) << j;
            uint32_t um = __reduce_or_sync(0xFFFFFFFFu, fb);
            while (um) {
              const uint32_t c0 = 8u * ((uint32_t)__ffs((int)um) - 1u);
              um &= um - 1;
              if (c0 >= nvalid) break;
              uint32_t fm = __ballot_sync(0xFFFFFFFFu, cnt + 8 > p.lcap);
              while (fm) {
                const int src = __ffs((int)fm) - 1;
                fm &= fm - 1;
                uint64_t* buf = reinterpret_cast<uint64_t*>(__shfl_sync(0xFFFFFFFFu, reinterpret_cast<unsigned long long>(list), src));
                __syncwarp();
                const uint64_t kth = p.lcap <= 256 ? gt4_warp_compact_small(buf, p.lcap, p.k) : gt4_warp_compact(buf, p.lcap, p.k);
                if ((int)lane == src) {
                  cnt = (uint32_t)kth;
                  atomicMin(gq, (uint32_t)(kth >> 32));
                  T = (float)(popq - (int)(uint32_t)(kth >> 32));
                  Te = fmaxf(T, Tg);
                }
              }
              uint32_t w[8];
              asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                           : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
                           : "r"(taddr + g * GR + c0)
                           : "memory");
              asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const float x = __uint_as_float(w[i]);
                if (x > Te && c0 + i < nvalid)
                  list[cnt++] = ((uint64_t)(uint32_t)(popq - (int)x) << 32) | (uint32_t)(p.key_offset + r0 + c0 + i);
              }
            }
            HB_DIAG_ADD(dg_ins);
          }
          if (kRegList && hit) {
            // Rare path, kept warp-uniform: for every column some lane wants, each lane re-reads its own value of that
            // column from TMEM (the accumulator is still ours) and the lanes that pass run one copy of the insert.
            HB_DIAG_T0();
            const uint64_t r0 = tile_row0 + (uint64_t)g * GR;
            const uint32_t nvalid = (uint32_t)min((uint64_t)GR, row_end > r0 ? row_end - r0 : 0);
            uint64_t m = (uint64_t)__reduce_or_sync(0xFFFFFFFFu, h0) | ((uint64_t)__reduce_or_sync(0xFFFFFFFFu, h1) << 32);
            if (nvalid < GR) m &= (1ull << nvalid) - 1ull;
            {
              while (m) {
                const uint32_t col = (uint32_t)__ffsll((long long)m) - 1u;
                m &= m - 1;
                uint32_t xb;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(xb) : "r"(taddr + g * GR + col) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                const float x = __uint_as_float(xb);
                if (x > Te) {
                  const uint32_t hd = (uint32_t)(popq - (int)x);
                  const uint32_t key = (uint32_t)(p.key_offset + r0 + col);
                  if (kRegList) {
                    const uint64_t e = ((uint64_t)hd << 32) | key;
#pragma unroll
                    for (int i = 15; i > 0; --i) {
                      if (rl[i - 1] > e) rl[i] = rl[i - 1];
                      else if (rl[i] > e) rl[i] = e;
                    }
                    if (rl[0] > e) rl[0] = e;
                    const uint64_t last = rl[15];
                    if (last != kNone) {
                      atomicMin(gq, (uint32_t)(last >> 32));
                      T = (float)(popq - (int)(uint32_t)(last >> 32));
                    }
                  } else {
                    list[cnt++] = ((uint64_t)hd << 32) | key;
                  }
                  Te = fmaxf(T, Tg);
                }
                if (!kRegList) {  // warp-uniform: compact every buffer that just filled
                  uint32_t fm = __ballot_sync(0xFFFFFFFFu, cnt == p.lcap);
                  while (fm) {
                    const int src = __ffs((int)fm) - 1;
                    fm &= fm - 1;
                    uint64_t* buf = reinterpret_cast<uint64_t*>(__shfl_sync(0xFFFFFFFFu, reinterpret_cast<unsigned long long>(list), src));
                    const uint64_t kth = p.lcap <= 256 ? gt4_warp_compact_small(buf, p.lcap, p.k) : gt4_warp_compact(buf, p.lcap, p.k);
                    if ((int)lane == src) {
                      cnt = (uint32_t)kth;
                      atomicMin(gq, (uint32_t)(kth >> 32));
                      T = (float)(popq - (int)(uint32_t)(kth >> 32));
                      Te = fmaxf(T, Tg);
                    }
                  }
                }
              }
            }
            HB_DIAG_ADD(dg_ins);
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        if (kPair) {  // one arrival per warp on the leader's barrier
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(bTEmptyL + 8 * acc);
        } else {
          mbar_arrive(bTEmpty + 8 * acc);
        }
        HB_TRACE(tcount, 6 + warp);
      }
      if (kRegList && live) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if ((uint32_t)i + p.k >= 16u) list[(uint32_t)i + p.k - 16u] = rl[i];
      }
    }
#ifdef HB_GT_DIAG
    if (warp == 0 && lane == 0) {
      g_gt4_diag[blockIdx.x * 16 + 5] = dg_tfull;
      g_gt4_diag[blockIdx.x * 16 + 6] = dg_ins;
      g_gt4_diag[blockIdx.x * 16 + 7] = tcount;
    }
#endif
  } else if (warp < kEpiWarps + kExpWarps) {
    // ===================== expansion: packed bits -> E2M1 operands in shared memory =====================
    const uint32_t ew = warp - kEpiWarps;
    const uint32_t et = ew * 32 + lane;  // 0..255: thread of the expansion role (the query-tile fill is indexed by it)
    const uint32_t grp = ew >> 2, wi = ew & 3u;  // a group = one warp on each SM sub-partition (warp id % 4)
    uint32_t it = 0;
    [[maybe_unused]] long long dg_empty = 0, dg_fence = 0, dg_afree = 0, dg_top = 0, dg_sts = 0;
    uint32_t xt = 0;  // tiles so far: parity of the raw buffer
    for (uint32_t item = item0; item < n_items; item += item_step, ++it) {
      const uint32_t qtile = qtile_of(item), chunk = item / n_qslots;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      // ---- A: the query tile.  Word c of a K-block's eight fills 16-byte chunk c; output word u of the chunk holds,
      // in nibble j, the element for bit 4j + u of the word: +-2, +-1, +-0.5, +-0.5 (sign bit = nibble bit 3, set when the
      // query bit is clear), matching the data scales 0.5, 1, 2, 2 below ----
      {
        HB_DIAG_T0();
        mbar_wait(bAFree, (it & 1) ^ 1);
        HB_DIAG_ADD(dg_afree);
      }
      {
        const uint32_t arow = et >> 1, kb0 = (et & 1) * 2;
        const uint32_t q = qtile * kTileM + arow;
        const bool live = q < p.nq;
        const uint4* qp = reinterpret_cast<const uint4*>(p.queries + (size_t)(live ? q : 0) * 16);
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const uint32_t kb = kb0 + kk;
          const uint4 qa = __ldg(qp + 2 * kb), qb = __ldg(qp + 2 * kb + 1);
          const uint32_t w[8] = {qa.x, qa.y, qa.z, qa.w, qb.x, qb.y, qb.z, qb.w};
          const uint32_t rowaddr = sA + kb * kAKBlockBytes + arow * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const uint32_t mag = (u == 0) ? 0x44444444u : (u == 1) ? 0x22222222u : 0x11111111u;
              const uint32_t clr = ((w[c] >> u) & 0x11111111u) ^ 0x11111111u;
              o[u] = live ? (mag | (clr << 3)) : 0u;
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + (((uint32_t)c ^ (arow & 7u)) << 4)), "r"(o[0]), "r"(o[1]),
                         "r"(o[2]), "r"(o[3])
                         : "memory");
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) {
          if (kPair) mbar_arrive_cluster(bAFullL);
          else mbar_arrive(bAFull);
        }
      }
      // ---- B: data tiles ----
      // The loader warp's bulk copies put the tile's packed rows in the raw buffer; each thread takes the 64 bytes its
      // group needs of two rows (64 wi + lane and 64 wi + 32 + lane) into registers and the buffer goes back to the
      // loader as soon as the values are in registers.  Global loads issued by
      // these warps would be waited for by that fence, a full memory latency per tile (28 -> 40 ms).
      // Two groups of four warps fill alternate pairs of stages: group 0 the K-blocks 0,1 of every tile, group 1 the
      // K-blocks 2,3, so one group stores while the other sits in its fence.  Eight consecutive lanes read and write
      // eight consecutive rows: no bank conflicts on either side.
      // 120-row plans: one row per thread (rows 32 wi + lane) and the ring holds two tiles.
      constexpr int kRowsPerThread = PL::rows_per_thread;
      for (uint32_t t = t0; t < t1; ++t, ++xt) {
        uint32_t cur[kRowsPerThread][16];
        {
          HB_DIAG_T0();
          mbar_wait(bRawFull, xt & 1);
          HB_DIAG_ADD(dg_top);
        }
#pragma unroll
        for (int h = 0; h < kRowsPerThread; ++h) {
          const uint32_t lrow = 32 * kRowsPerThread * wi + 32 * h + lane;  // lrow & 7 == lane & 7; rows past the data end arrive as zeros
          const uint32_t ra = sRaw + (lrow < PL::rows ? lrow : 0u) * kRawPitch;
#pragma unroll
          for (int e = 0; e < 4; ++e)
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                         : "=r"(cur[h][4 * e]), "=r"(cur[h][4 * e + 1]), "=r"(cur[h][4 * e + 2]), "=r"(cur[h][4 * e + 3])
                         : "r"(ra + (((4u * grp + e) ^ (lane & 7u)) << 4))
                         : "memory");
        }
        // hand the buffer back once the values have arrived (the XOR of all of them is a dependency on every load; it
        // cannot be all ones for sixteen words of real data and the store below never happens)
        {
          uint32_t dep = 0;
#pragma unroll
          for (int h = 0; h < kRowsPerThread; ++h)
#pragma unroll
            for (int e = 0; e < 16; ++e) dep |= cur[h][e];
          if (dep == 0xFFFFFFFFu && t == 0xFFFFFFFFu) asm volatile("st.shared.b32 [%0], %1;" ::"r"(sRaw), "r"(dep) : "memory");
        }
        // (the copy for the next tile then has this whole tile's expansion time to land)
        __syncwarp();
        if (lane == 0) mbar_arrive(bRawFree);
        // group g fills K-blocks 2g, 2g + 1 of the tile's four stages; the ring position and its phase follow the tile count
        const uint32_t sbase = (xt % PL::ring_tiles) * 4, sphase = (xt / PL::ring_tiles) & 1u;
        if constexpr (kRowsPerThread == 1 && HB_GT_MERGE_FENCE) {
          // half-size stages: both of the group's K-blocks behind ONE proxy fence (the fence, not the stores, is what an
          // expansion warp spends its time in), so a row costs as many fences as it does with 240-row stages
          const uint32_t st0 = sbase + 2 * grp;
          {
            HB_DIAG_T0();
            mbar_wait(bEmpty + 8 * st0, sphase ^ 1);
            mbar_wait(bEmpty + 8 * (st0 + 1), sphase ^ 1);
            HB_DIAG_ADD(dg_empty);
          }
          const uint32_t lrow = 32 * wi + lane;
          if (lrow < PL::rows) {
#pragma unroll
            for (int sg = 0; sg < 2; ++sg) {
              const uint32_t rowaddr = sB + (st0 + sg) * kStageBytes + lrow * 128;
#pragma unroll
              for (int c = 0; c < 8; ++c) {
                const uint32_t w = cur[0][8 * sg + c];
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + (((uint32_t)c ^ (lane & 7u)) << 4)), "r"(w & 0x11111111u),
                             "r"(w & 0x22222222u), "r"(w & 0x44444444u), "r"((w >> 1) & 0x44444444u)
                             : "memory");
              }
            }
          }
          {
            HB_DIAG_T0();
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            HB_DIAG_ADD(dg_fence);
          }
          if (lane == 0) {
            if (kPair) {
              mbar_arrive_cluster(bFullL + 8 * st0);
              mbar_arrive_cluster(bFullL + 8 * (st0 + 1));
            } else {
              mbar_arrive(bFull + 8 * st0);
              mbar_arrive(bFull + 8 * (st0 + 1));
            }
          }
        } else {
#pragma unroll
        for (int sg = 0; sg < 2; ++sg) {
          const uint32_t kb = 2 * grp + sg, stage = sbase + kb;
          {
            HB_DIAG_T0();
            mbar_wait(bEmpty + 8 * stage, sphase ^ 1);
            HB_DIAG_ADD(dg_empty);
          }
          [[maybe_unused]] const long long dg_s0 = clock64();
#ifndef HB_GT_PROBE_STS
#define HB_GT_PROBE_STS 8  // probe builds (wrong results): fewer stores or fewer MMAs, to time the other side alone
#endif
#pragma unroll
          for (int h = 0; h < kRowsPerThread; ++h) {
            const uint32_t lrow = 32 * kRowsPerThread * wi + 32 * h + lane;
            const uint32_t rowaddr = sB + stage * kStageBytes + lrow * 128;
            if (lrow < PL::rows) {
#pragma unroll
              for (int c = 0; c < HB_GT_PROBE_STS; ++c) {  // E2M1 codes 0001 = 0.5, 0010 = 1, 0100 = 2; bit 3 of each nibble reuses 2
                const uint32_t w = cur[h][8 * sg + c];
                asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + (((uint32_t)c ^ (lane & 7u)) << 4)), "r"(w & 0x11111111u),
                             "r"(w & 0x22222222u), "r"(w & 0x44444444u), "r"((w >> 1) & 0x44444444u)
                             : "memory");
              }
            }
          }
#ifdef HB_GT_DIAG
          dg_sts += clock64() - dg_s0;
#endif
          {
            HB_DIAG_T0();
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            HB_DIAG_ADD(dg_fence);
          }
          if (lane == 0) {
            if (kPair) mbar_arrive_cluster(bFullL + 8 * stage);
            else mbar_arrive(bFull + 8 * stage);
          }
          if (wi == 0) HB_TRACE(xt, 10 + kb);
        }
        }  // (separate fences)
      }
    }
#ifdef HB_GT_DIAG
    if (ew == 0 && lane == 0) {
      g_gt4_diag[blockIdx.x * 16 + 3] = dg_empty;
      g_gt4_diag[blockIdx.x * 16 + 4] = dg_fence;
      g_gt4_diag[blockIdx.x * 16 + 8] = dg_afree;
      g_gt4_diag[blockIdx.x * 16 + 9] = dg_top;
      g_gt4_diag[blockIdx.x * 16 + 10] = dg_sts;
    }
#endif
  } else if (warp == kEpiWarps + kExpWarps) {
    // ===================== MMA issue: one elected thread =====================
    uint32_t stage = 0, sphase = 0, tcount = 0, it = 0;
    [[maybe_unused]] long long dg_tempty = 0, dg_full = 0;
    [[maybe_unused]] const long long dg_start = clock64();
    const uint64_t da0 = umma_desc(sA), db0 = umma_desc(sB);
    // pair: only the leader issues; the descriptors name the same offsets in both CTAs' shared memory
    for (uint32_t item = item0; item < n_items && rank == 0; item += item_step, ++it) {
      const uint32_t chunk = item / n_qslots;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      if (kPair) mbar_wait_cluster(bAFull, it & 1);
      else mbar_wait(bAFull, it & 1);
      for (uint32_t t = t0; t < t1; ++t, ++tcount) {
        const uint32_t acc = tcount % NACC, ause = tcount / NACC;
        {
          HB_DIAG_T0();
          if (kPair) mbar_wait_cluster(bTEmpty + 8 * acc, (ause & 1) ^ 1);
          else mbar_wait(bTEmpty + 8 * acc, (ause & 1) ^ 1);
          HB_DIAG_ADD(dg_tempty);
        }
        HB_TRACE(tcount, 0);
        const uint32_t tmem_d = tmem + acc * PL::acc_stride;
#pragma unroll
        for (int kb = 0; kb < kKBlocks; ++kb) {
          {
            HB_DIAG_T0();
            if (kPair) mbar_wait_cluster(bFull + 8 * stage, sphase);
            else mbar_wait(bFull + 8 * stage, sphase);
            HB_DIAG_ADD(dg_full);
          }
          HB_TRACE(tcount, 18 + kb);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (elect_one_sync()) {
            const uint64_t da = da0 + (uint64_t)(kb * (kAKBlockBytes >> 4)), db = db0 + (uint64_t)(stage * (kStageBytes >> 4));
#pragma unroll
            for (int j = 0; j < HB_GT_PROBE_MMAS; ++j) {  // K = 64 E2M1 per instruction: +32 B inside the swizzle atom = +2 in the encoded address
              if (kPair) umma_mxf4_pair(tmem_d, da + 2 * j, db + 2 * j, PL::idesc, tmem + kSfCol, (kb | j) != 0);
              else umma_mxf4(tmem_d, da + 2 * j, db + 2 * j, PL::idesc, tmem + kSfCol, (kb | j) != 0);
            }
            if (kPair) {
              umma_commit_pair(bEmpty + 8 * stage);
              if (kb == kKBlocks - 1) {
                umma_commit_pair(bTFull + 8 * acc);
                if (t == t1 - 1) umma_commit_pair(bAFree);
              }
            } else {
              umma_commit(bEmpty + 8 * stage);
              if (kb == kKBlocks - 1) {
                umma_commit(bTFull + 8 * acc);
                if (t == t1 - 1) umma_commit(bAFree);
              }
            }
          }
          __syncwarp();
          if (kb == kKBlocks - 1) HB_TRACE(tcount, 1);
          if (++stage == kStages) {
            stage = 0;
            sphase ^= 1;
          }
        }
      }
    }
    // pair: the last multicast commits have landed in both CTAs before either leaves (the final cluster barrier follows)
    if (kPair && rank == 0 && it > 0) mbar_wait(bAFree, (it - 1) & 1);
#ifdef HB_GT_DIAG
    if (lane == 0) {
      g_gt4_diag[blockIdx.x * 16 + 0] = clock64() - dg_start;
      g_gt4_diag[blockIdx.x * 16 + 1] = dg_tempty;
      g_gt4_diag[blockIdx.x * 16 + 2] = dg_full;
    }
#endif
  } else {
    // ===================== loader: the tile's packed rows into the raw buffer, one TMA copy =====================
    // (async proxy end to end: the copies complete on raw_full's transaction count, no thread waits on a load)
    uint32_t xt = 0;
    for (uint32_t item = item0; item < n_items; item += item_step) {
      const uint32_t chunk = item / n_qslots;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      for (uint32_t t = t0; t < t1; ++t, ++xt) {
        mbar_wait(bRawFree, (xt & 1) ^ 1);
        if (lane == 0) {  // pair: this CTA's half of the tile's rows
          asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bRawFull), "r"(PL::raw_bytes) : "memory");
          asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(sRaw),
                       "l"(&tmap), "r"(0), "r"((int)(t * TN + rank * PL::rows)), "r"(bRawFull)
                       : "memory");
        }
        __syncwarp();
      }
    }
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  if (kPair) cluster_sync_all();  // neither CTA's shared memory or TMEM goes away while the other may still signal it
  else __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    if (kPair) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

}  // namespace

#ifdef HB_GT_DIAG
extern "C" int hb_debug_gt4_diag(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, g_gt4_diag, sizeof(long long) * 256 * 16);
}
#endif

namespace {
// cuTensorMapEncodeTiled through the runtime's driver entry point lookup (the library does not link libcuda)
using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult st = cudaDriverEntryPointSymbolNotFound;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &st) != cudaSuccess || st != cudaDriverEntryPointSuccess)
      f = nullptr;
    return reinterpret_cast<EncodeTiledFn>(f);
  }();
  return fn;
}
}  // namespace

bool ground_truth_mxf4_supported(const void* d_data, const void* d_queries, uint64_t n, uint64_t nq, uint32_t k) {
  return n >= 1 && nq >= 1 && k >= 1 && k <= 1024 && (reinterpret_cast<uintptr_t>(d_data) & 31) == 0 &&
         (reinterpret_cast<uintptr_t>(d_queries) & 15) == 0 && n <= 0xFFFFFFFFull - kTileN;
}

void ground_truth_mxf4_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                             void* d_packed, cudaStream_t stream, uint32_t* d_bound) {
  HB_REQUIRE(ground_truth_mxf4_supported(d_data, d_queries, n, nq, k), HB_ERR_UNSUPPORTED,
             "tcgen05 ground truth: needs 32-byte aligned data, 1 <= k <= 1024");
  HB_REQUIRE(key_offset + n <= 0xFFFFFFFFull, HB_ERR_UNSUPPORTED, "keys must stay below 2^32");
  HB_REQUIRE(nq <= 0x7FFFFF00ull, HB_ERR_UNSUPPORTED, "too many queries in one call");
  int dev = 0, sms = 0;
  HB_CUDA(cudaGetDevice(&dev));
  HB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  GtParams p{};
  p.data = static_cast<const uint64_t*>(d_data);
  p.queries = static_cast<const uint64_t*>(d_queries);
  p.n = n;
  p.key_offset = key_offset;
  p.nq = (uint32_t)nq;
  p.k = k;
  p.n_qtiles = (uint32_t)((nq + kTileM - 1) / kTileM);
  const char* plan_str = std::getenv("HB_GT_PLAN");  // read per call: the tests switch plans inside one process
  const int plan_env = plan_str ? std::atoi(plan_str) : HB_GT_PLAN_DEFAULT;
  int plan = (plan_env >= 0 && plan_env <= 2) ? plan_env : HB_GT_PLAN_DEFAULT;
  if (plan == 1 && (sms < 2 || p.n_qtiles < 2)) plan = 0;
  const bool pair = plan == 1;
  const uint32_t tile_n = plan == 2 ? Plan<2>::tile_n : Plan<0>::tile_n;
  p.n_tiles = (uint32_t)((n + tile_n - 1) / tile_n);
  // CTA pairs: two adjacent query tiles per work item, one per CTA; the pair shares the chunk's data tiles
  const uint32_t n_qslots = pair ? (p.n_qtiles + 1) / 2 : p.n_qtiles;
  const uint32_t workers = pair ? (uint32_t)sms / 2 : (uint32_t)sms;
  // chunks of data tiles: enough items to fill the grid several times over, picked so the last wave is nearly full
  uint32_t best_chunks = 1;
  double best_eff = -1.0;
  const uint32_t max_chunks = std::min<uint32_t>(p.n_tiles, 96);
  for (uint32_t c = 1; c <= max_chunks; ++c) {
    const uint32_t tpc = (p.n_tiles + c - 1) / c;
    const uint32_t chunks = (p.n_tiles + tpc - 1) / tpc;
    const uint64_t items = (uint64_t)chunks * n_qslots;
    const uint64_t waves = (items + workers - 1) / workers;
    // cost model: waves x (tiles per chunk + per-item overhead: writing the query tile and refilling the item's lists,
    // in tile times -- register lists (k <= 16) refill in ~25, lists in global memory in ~4 per entry)
    const double overhead = (4.0 + (k <= 16 ? 50.0 : 0.5 * k)) * (double)Plan<0>::tile_n / tile_n;
    const double cost = (double)waves * (tpc + overhead);
    const double eff = ((double)p.n_tiles * n_qslots / workers) / cost;
    if (eff > best_eff + 1e-9) {
      best_eff = eff;
      best_chunks = chunks;
    }
  }
  p.n_chunks = best_chunks;
  p.tiles_per_chunk = (p.n_tiles + p.n_chunks - 1) / p.n_chunks;
  p.n_chunks = (p.n_tiles + p.tiles_per_chunk - 1) / p.tiles_per_chunk;

  uint64_t* partial = nullptr;
#ifndef HB_GT_LCAP_MULT
#define HB_GT_LCAP_MULT 2  // append-buffer slots per row = this x k (measured: 3 and 4 are not faster, profiles/r01k_*)
#endif
  p.lcap = k <= 16 ? k : HB_GT_LCAP_MULT * k;
  const size_t list_bytes = (size_t)p.n_chunks * nq * p.lcap * 8;
  HB_CUDA(cudaMallocAsync(&partial, list_bytes + nq * 4, stream));
  p.partial = partial;
  if (d_bound) {
    p.gthr = d_bound;  // carried over from the caller's earlier slices
  } else {
    p.gthr = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(partial) + list_bytes);
    HB_CUDA(cudaMemsetAsync(p.gthr, 0xFF, nq * 4, stream));
  }
  // the packed rows as a 2-D tensor of 32-bit words, n rows x 32; one box = one data tile, 128-byte swizzled, rows past
  // the end filled with zeros
  CUtensorMap tmap;
  {
    EncodeTiledFn enc = encode_tiled_fn();
    if (!enc) {
      cudaFreeAsync(partial, stream);
      HB_REQUIRE(false, HB_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available from this driver");
    }
    const cuuint64_t gdim[2] = {32, n}, gstride[1] = {128};
    const cuuint32_t box[2] = {32, plan == 0 ? Plan<0>::rows : plan == 1 ? Plan<1>::rows : Plan<2>::rows}, estride[2] = {1, 1};
    const CUresult r = enc(&tmap, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void*>(d_data), gdim, gstride, box, estride,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
      cudaFreeAsync(partial, stream);
      HB_REQUIRE(false, HB_ERR_CUDA, "cuTensorMapEncodeTiled failed");
    }
  }
  const uint32_t n_items = n_qslots * p.n_chunks;
  cudaError_t le = cudaSuccess;
  auto launch = [&](auto kernel, uint32_t smem, uint32_t cluster) {
    le = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (le != cudaSuccess) return;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(cluster * std::min<uint32_t>(n_items, workers));
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr{};
    attr.id = cudaLaunchAttributeClusterDimension;
    attr.val.clusterDim.x = cluster;
    attr.val.clusterDim.y = attr.val.clusterDim.z = 1;
    cfg.attrs = &attr;
    cfg.numAttrs = cluster > 1 ? 1 : 0;
    le = cudaLaunchKernelEx(&cfg, kernel, p, tmap);
  };
  if (plan == 1) {
    if (k <= 16) launch(gt_mxf4_kernel<true, 1>, Plan<1>::smem_bytes, 2);
    else launch(gt_mxf4_kernel<false, 1>, Plan<1>::smem_bytes, 2);
  } else if (plan == 2) {
    if (k <= 16) launch(gt_mxf4_kernel<true, 2>, Plan<2>::smem_bytes, 1);
    else launch(gt_mxf4_kernel<false, 2>, Plan<2>::smem_bytes, 1);
  } else {
    if (k <= 16) launch(gt_mxf4_kernel<true, 0>, Plan<0>::smem_bytes, 1);
    else launch(gt_mxf4_kernel<false, 0>, Plan<0>::smem_bytes, 1);
  }
  if (le != cudaSuccess) {
    cudaFreeAsync(partial, stream);
    HB_CUDA(le);
  }
  count_launches();
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    try {
      topk_merge_device(partial, p.n_chunks, nq, k, d_packed, stream, /*sorted=*/k <= 16, p.lcap);
    } catch (...) {
      cudaFreeAsync(partial, stream);
      throw;
    }
  }
  cudaFreeAsync(partial, stream);
  HB_CUDA(e);
}

}  // namespace hb
