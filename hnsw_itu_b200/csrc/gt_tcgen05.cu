// gt_tcgen05.cu -- the `ground-truth` path as a tcgen05 / TMEM int8 GEMM with a fused per-row top-k.
//
// Replaces Bruteforce::search (src/index/bruteforce.rs:37-47) + MinK::min_k
// (src/collections/mod.rs:52-77) + the CLI's per-row sort (cli/src/main.rs:738-741) for batches of
// queries: Hamming(q, d) = pop(q) - sum_i (2 q_i - 1) d_i, so one signed-int8 GEMM
//     acc[query][row] = sum_i A[query][i] * B[row][i],   A = +-scale_i (query bit), B = d_i * 64 / scale_i
// gives acc = 64 * (pop(q) - Hamming) exactly (int32 accumulators, |acc| <= 65536).  The position-dependent
// power-of-two scale lets the data operand be produced from the packed bits with ONE logic op per four
// int8 values (`w & (0x01010101 << s)`); the K dimension is permuted identically on both operands, which
// leaves every dot product unchanged.
//
// One persistent CTA per SM, 13 warps:
//   warps 0-3   epilogue: tcgen05.ld the accumulator tile (thread = query row), one 3-input max per two
//               values against the row's current k-th best, rare single-thread insert into the row's list
//   warps 4-11  expansion: 256-bit loads of packed rows (prefetched one tile ahead), bits -> int8 in the
//               128-byte-swizzled K-major layout the UMMA shared-memory descriptor names
//   warp 12     one thread issues tcgen05.mma.cta_group::1.kind::i8 (M=128, N=256, K=32), commits to mbarriers
// Shared memory: the query tile expanded once per work item (128 KB) + 3 stages of one K-block of the data
// tile (32 KB each).  TMEM: two 128x256 int32 accumulators (all 512 columns).
// Work item = (128-query tile, chunk of data tiles); per-item ascending lists are merged exactly by
// topk_merge_kernel (bruteforce.cu).  Rows inside an item are visited in increasing key order, so a candidate
// whose distance equals the current k-th best always has the larger key and a strict compare is the
// reference's (distance, key) order (src/index/mod.rs:90-97).
#include <algorithm>
#include <climits>

#include "gt_common.cuh"
#include "hb_internal.h"

namespace hb {
namespace {

constexpr int kTileM = 128;   // queries per work item (UMMA M)
constexpr int kTileN = 256;   // data rows per accumulator tile (UMMA N)
constexpr int kKBlocks = 8;   // 1024 bits = 8 K-blocks of 128 int8 (one 128-B swizzle atom row each)
constexpr int kStages = 3;
constexpr int kEpiWarps = 4, kExpWarps = 8;
constexpr int kThreads = (kEpiWarps + kExpWarps + 1) * 32;
constexpr uint32_t kABytes = kTileM * 1024;           // 8 K-blocks x (128 rows x 128 B)
constexpr uint32_t kAKBlockBytes = kTileM * 128;
constexpr uint32_t kStageBytes = kTileN * 128;
constexpr uint32_t kBarOffset = kABytes + kStages * kStageBytes;
constexpr uint32_t kSmemBytes = kBarOffset + 128 + 1024;  // + barriers + alignment slack

// instruction descriptor (cute::UMMA::InstrDescriptor layout): D = s32, A = B = signed 8 bit, both K-major, N = 256, M = 128
constexpr uint32_t kIdesc = (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(kTileN >> 3) << 17) | ((uint32_t)(kTileM >> 4) << 24);

using namespace gt;  // mbarrier / tcgen05 helpers, HB_TMEM_LD32 (gt_common.cuh)
constexpr auto mbar_wait = gt::mbar_wait<0>;

#ifdef HB_GT_DIAG
// per-CTA phase clocks (diagnostic build only, tools/gt_diag.py): [0] kernel cycles, [1] MMA warp waiting for a free
// accumulator, [2] MMA warp waiting for a full stage, [3] expansion warp 0 waiting for an empty stage, [4] its fences,
// [5] epilogue warp 0 waiting for a full accumulator, [6] its inserts, [7] tiles, [8] expansion waiting for A free
__device__ long long g_gt_diag[256 * 16];
// trace of CTA 0, tiles [kTr0, kTr0+32): at g_gt_diag[2400 + 32*(tile-kTr0) + slot]
//   slot 0 MMA passes acc-free wait, 1 MMA issued last stage, 2+w epilogue warp w sees acc full, 6+w warp w releases acc,
//   10+kb expansion warp 0 arrives stage kb, 18+kb MMA passes full wait of stage kb
constexpr uint32_t kTr0 = 3000;
#define HB_TRACE(tile, slot) \
  do { if (blockIdx.x == 0 && (tile) >= kTr0 && (tile) < kTr0 + 32 && (threadIdx.x & 31) == 0) g_gt_diag[2400 + 32 * ((tile) - kTr0) + (slot)] = clock64(); } while (0)
#define HB_DIAG_T0() const long long dg_t0 = clock64()
#define HB_DIAG_ADD(var) var += clock64() - dg_t0
#else
#define HB_TRACE(tile, slot)
#define HB_DIAG_T0()
#define HB_DIAG_ADD(var)
#endif

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

template <bool kRegList>
__global__ void __launch_bounds__(kThreads, 1) gt_mma_kernel(const GtParams p) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;  // swizzle atoms need 1024-B alignment
  unsigned char* sm = smem_raw + (base - raw);
  const uint32_t sA = base, sB = base + kABytes, sBar = base + kBarOffset;
  // barriers: full[3] empty[3] tfull[2] tempty[2] a_full a_free
  const uint32_t bFull = sBar, bEmpty = sBar + 8 * kStages, bTFull = sBar + 16 * kStages, bTEmpty = bTFull + 16, bAFull = bTEmpty + 16,
                 bAFree = bAFull + 8;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + kBarOffset + 120);

  const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(bFull + 8 * s, kExpWarps);
      mbar_init(bEmpty + 8 * s, 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(bTFull + 8 * a, 1);
      mbar_init(bTEmpty + 8 * a, kEpiWarps * 32);
    }
    mbar_init(bAFull, kExpWarps);
    mbar_init(bAFree, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;

  const uint32_t n_items = p.n_qtiles * p.n_chunks;

  if (warp < kEpiWarps) {
    // ===================== epilogue: thread = one query row of the accumulator =====================
    const uint32_t row = warp * 32 + lane;
    uint32_t tcount = 0;
    [[maybe_unused]] long long dg_tfull = 0, dg_ins = 0;
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x) {
      const uint32_t qtile = item % p.n_qtiles, chunk = item / p.n_qtiles;
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
      int T = live ? INT_MIN : INT_MAX;
      uint32_t* gq = p.gthr + (live ? q : 0);
      const uint64_t row_end = min(p.n, (uint64_t)t1 * kTileN);
      for (uint32_t t = t0; t < t1; ++t, ++tcount) {
        const uint32_t acc = tcount & 1, use = tcount >> 1;
        const uint32_t g_now = *reinterpret_cast<volatile uint32_t*>(gq);
        {
          HB_DIAG_T0();
          mbar_wait(bTFull + 8 * acc, use & 1);
          HB_DIAG_ADD(dg_tfull);
        }
        HB_TRACE(tcount, 2 + warp);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int Tg = (g_now == 0xFFFFFFFFu) ? INT_MIN : 64 * (popq - (int)g_now) - 64;
        int Te = max(T, Tg);
        const uint32_t taddr = tmem + ((warp * 32u) << 16) + acc * kTileN;
        const uint64_t tile_row0 = (uint64_t)t * kTileN;
#pragma unroll 1
        for (int g = 0; g < kTileN / 64; ++g) {
          uint32_t h0 = 0, h1 = 0;
          bool hit;
          {
            int v[64];
            HB_TMEM_LD32(v, taddr + g * 64);
            HB_TMEM_LD32((v + 32), taddr + g * 64 + 32);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            int mx = v[0];
#pragma unroll
            for (int c = 1; c < 64; c += 1) mx = max(mx, v[c]);
            hit = __any_sync(0xFFFFFFFFu, mx > Te);
            if (hit) {  // warp-uniform: which columns pass for which lane
#pragma unroll
              for (int c = 0; c < 32; ++c) {
                h0 |= (uint32_t)(v[c] > Te) << c;
                h1 |= (uint32_t)(v[32 + c] > Te) << c;
              }
            }
          }
          if (hit) {
            // Rare path, kept warp-uniform: for every column some lane wants, each lane re-reads its own value of that
            // column from TMEM (the accumulator is still ours) and the lanes that pass run one copy of the insert.
            HB_DIAG_T0();
            const uint64_t r0 = tile_row0 + (uint64_t)g * 64;
            const uint32_t nvalid = (uint32_t)min((uint64_t)64, row_end > r0 ? row_end - r0 : 0);
            uint64_t m = (uint64_t)__reduce_or_sync(0xFFFFFFFFu, h0) | ((uint64_t)__reduce_or_sync(0xFFFFFFFFu, h1) << 32);
            if (nvalid < 64) m &= (1ull << nvalid) - 1ull;
            {
              while (m) {
                const uint32_t col = (uint32_t)__ffsll((long long)m) - 1u;
                m &= m - 1;
                int x;
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(x) : "r"(taddr + g * 64 + col) : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (x > Te) {
                  const uint32_t key = (uint32_t)(p.key_offset + r0 + col);
                  if (kRegList) {
                    const uint64_t e = ((uint64_t)(uint32_t)(popq - (x >> 6)) << 32) | key;
#pragma unroll
                    for (int i = 15; i > 0; --i) {
                      if (rl[i - 1] > e) rl[i] = rl[i - 1];
                      else if (rl[i] > e) rl[i] = e;
                    }
                    if (rl[0] > e) rl[0] = e;
                    const uint64_t last = rl[15];
                    if (last != kNone) {
                      atomicMin(gq, (uint32_t)(last >> 32));
                      T = 64 * (popq - (int)(uint32_t)(last >> 32));
                    }
                  } else {
                    list[cnt++] = ((uint64_t)(uint32_t)(popq - (x >> 6)) << 32) | key;
                  }
                  Te = max(T, Tg);
                }
                if (!kRegList) {  // warp-uniform: compact every buffer that just filled
                  uint32_t fm = __ballot_sync(0xFFFFFFFFu, cnt == p.lcap);
                  while (fm) {
                    const int src = __ffs((int)fm) - 1;
                    fm &= fm - 1;
                    uint64_t* buf = reinterpret_cast<uint64_t*>(__shfl_sync(0xFFFFFFFFu, reinterpret_cast<unsigned long long>(list), src));
                    const uint64_t kth = warp_compact(buf, p.lcap, p.k);
                    if ((int)lane == src) {
                      cnt = p.k;
                      atomicMin(gq, (uint32_t)(kth >> 32));
                      T = 64 * (popq - (int)(uint32_t)(kth >> 32));
                      Te = max(T, Tg);
                    }
                  }
                }
              }
            }
            HB_DIAG_ADD(dg_ins);
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        mbar_arrive(bTEmpty + 8 * acc);
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
      g_gt_diag[blockIdx.x * 16 + 5] = dg_tfull;
      g_gt_diag[blockIdx.x * 16 + 6] = dg_ins;
      g_gt_diag[blockIdx.x * 16 + 7] = tcount;
    }
#endif
  } else if (warp < kEpiWarps + kExpWarps) {
    // ===================== expansion: packed bits -> int8 operands in shared memory =====================
    const uint32_t ew = warp - kEpiWarps;
    const uint32_t et = ew * 32 + lane;  // 0..255: row of the data tile
    uint32_t stage = 0, sphase = 0, it = 0;
    [[maybe_unused]] long long dg_empty = 0, dg_fence = 0, dg_afree = 0;
    [[maybe_unused]] uint32_t xt = 0;
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const uint32_t qtile = item % p.n_qtiles, chunk = item / p.n_qtiles;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      // prefetch the first data tile's row while the query tile is written
      uint64_t cur[16], nxt[16];
      {
        const uint64_t r = (uint64_t)t0 * kTileN + et;
        if (r < p.n) {
          const uint64_t* src = p.data + r * 16;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];"
                         : "=l"(cur[4 * j]), "=l"(cur[4 * j + 1]), "=l"(cur[4 * j + 2]), "=l"(cur[4 * j + 3])
                         : "l"(src + 4 * j));
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) cur[j] = 0;
        }
      }
      // ---- A: the query tile, +-(64 >> s) per bit (s = 7 shares scale 1 with s = 6's data scale 64) ----
      {
        HB_DIAG_T0();
        mbar_wait(bAFree, (it & 1) ^ 1);
        HB_DIAG_ADD(dg_afree);
      }
      {
        const uint32_t arow = et >> 1, kb0 = (et & 1) * 4;
        const uint32_t q = qtile * kTileM + arow;
        const bool live = q < p.nq;
        const uint4* qp = reinterpret_cast<const uint4*>(p.queries + (size_t)(live ? q : 0) * 16);
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          const uint32_t kb = kb0 + kk;
          uint4 qw = __ldg(qp + kb);
          const uint32_t w[4] = {qw.x, qw.y, qw.z, qw.w};
          const uint32_t rowaddr = sA + kb * kAKBlockBytes + arow * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int s = 4 * (c & 1) + u;
              const uint32_t sc = (s < 7) ? (64u >> s) : 1u;
              const uint32_t m = (w[c >> 1] >> s) & 0x01010101u;
              o[u] = live ? __vsub4(m * (2u * sc), sc * 0x01010101u) : 0u;
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + (((uint32_t)c ^ (arow & 7u)) << 4)), "r"(o[0]), "r"(o[1]),
                         "r"(o[2]), "r"(o[3])
                         : "memory");
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(bAFull);
      }
      // ---- B: data tiles ----
      // The next tile's row is fetched at the top of the tile (four 256-bit loads).  Every stage's proxy fence is a
      // MEMBAR that waits for outstanding loads, so the load latency is exposed once per tile, overlapped with the
      // wait for the first free stage.
      for (uint32_t t = t0; t < t1; ++t) {
        const uint64_t rn = (uint64_t)(t + 1) * kTileN + et;
        if (t + 1 < t1 && rn < p.n) {
          const uint64_t* src = p.data + rn * 16;
#pragma unroll
          for (int j = 0; j < 4; ++j)
            asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];"
                         : "=l"(nxt[4 * j]), "=l"(nxt[4 * j + 1]), "=l"(nxt[4 * j + 2]), "=l"(nxt[4 * j + 3])
                         : "l"(src + 4 * j));
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) nxt[j] = 0;
        }
#pragma unroll
        for (int kb = 0; kb < kKBlocks; ++kb) {
          {
            HB_DIAG_T0();
            mbar_wait(bEmpty + 8 * stage, sphase ^ 1);
            HB_DIAG_ADD(dg_empty);
          }
          const uint32_t w[4] = {(uint32_t)cur[2 * kb], (uint32_t)(cur[2 * kb] >> 32), (uint32_t)cur[2 * kb + 1],
                                 (uint32_t)(cur[2 * kb + 1] >> 32)};
          const uint32_t rowaddr = sB + stage * kStageBytes + et * 128;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            uint32_t o[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int s = 4 * (c & 1) + u;
              o[u] = (s < 7) ? (w[c >> 1] & (0x01010101u << s)) : ((w[c >> 1] >> 1) & 0x40404040u);
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(rowaddr + (((uint32_t)c ^ (et & 7u)) << 4)), "r"(o[0]), "r"(o[1]),
                         "r"(o[2]), "r"(o[3])
                         : "memory");
          }
          {
            HB_DIAG_T0();
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            HB_DIAG_ADD(dg_fence);
          }
          if (lane == 0) mbar_arrive(bFull + 8 * stage);
          if (ew == 0) HB_TRACE(xt, 10 + kb);
          if (++stage == kStages) {
            stage = 0;
            sphase ^= 1;
          }
        }
#pragma unroll
        for (int j = 0; j < 16; ++j) cur[j] = nxt[j];
#ifdef HB_GT_DIAG
        ++xt;
#endif
      }
    }
#ifdef HB_GT_DIAG
    if (ew == 0 && lane == 0) {
      g_gt_diag[blockIdx.x * 16 + 3] = dg_empty;
      g_gt_diag[blockIdx.x * 16 + 4] = dg_fence;
      g_gt_diag[blockIdx.x * 16 + 8] = dg_afree;
    }
#endif
  } else {
    // ===================== MMA issue: one elected thread =====================
    uint32_t stage = 0, sphase = 0, tcount = 0, it = 0;
    [[maybe_unused]] long long dg_tempty = 0, dg_full = 0;
    [[maybe_unused]] const long long dg_start = clock64();
    const uint64_t da0 = umma_desc(sA), db0 = umma_desc(sB);
    for (uint32_t item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
      const uint32_t chunk = item / p.n_qtiles;
      const uint32_t t0 = chunk * p.tiles_per_chunk, t1 = min(p.n_tiles, t0 + p.tiles_per_chunk);
      mbar_wait(bAFull, it & 1);
      for (uint32_t t = t0; t < t1; ++t, ++tcount) {
        const uint32_t acc = tcount & 1, ause = tcount >> 1;
        {
          HB_DIAG_T0();
          mbar_wait(bTEmpty + 8 * acc, (ause & 1) ^ 1);
          HB_DIAG_ADD(dg_tempty);
        }
        HB_TRACE(tcount, 0);
        const uint32_t tmem_d = tmem + acc * kTileN;
#pragma unroll
        for (int kb = 0; kb < kKBlocks; ++kb) {
          {
            HB_DIAG_T0();
            mbar_wait(bFull + 8 * stage, sphase);
            HB_DIAG_ADD(dg_full);
          }
          HB_TRACE(tcount, 18 + kb);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (elect_one_sync()) {
            const uint64_t da = da0 + (uint64_t)(kb * (kAKBlockBytes >> 4)), db = db0 + (uint64_t)(stage * (kStageBytes >> 4));
#pragma unroll
            for (int j = 0; j < 4; ++j)  // K = 32 int8 per instruction: +32 B inside the swizzle atom = +2 in the encoded address
              umma_i8(tmem_d, da + 2 * j, db + 2 * j, kIdesc, (kb | j) != 0);
            umma_commit(bEmpty + 8 * stage);
            if (kb == kKBlocks - 1) {
              umma_commit(bTFull + 8 * acc);
              if (t == t1 - 1) umma_commit(bAFree);
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
#ifdef HB_GT_DIAG
    if (lane == 0) {
      g_gt_diag[blockIdx.x * 16 + 0] = clock64() - dg_start;
      g_gt_diag[blockIdx.x * 16 + 1] = dg_tempty;
      g_gt_diag[blockIdx.x * 16 + 2] = dg_full;
    }
#endif
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

}  // namespace

#ifdef HB_GT_DIAG
extern "C" int hb_debug_gt_diag(long long* out) {
  return (int)cudaMemcpyFromSymbol(out, g_gt_diag, sizeof(long long) * 256 * 16);
}
#endif

bool ground_truth_mma_supported(const void* d_data, const void* d_queries, uint64_t n, uint64_t nq, uint32_t k) {
  return n >= 1 && nq >= 1 && k >= 1 && k <= 1024 && (reinterpret_cast<uintptr_t>(d_data) & 31) == 0 &&
         (reinterpret_cast<uintptr_t>(d_queries) & 15) == 0 && n <= 0xFFFFFFFFull - kTileN;
}

void ground_truth_mma_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                             void* d_packed, cudaStream_t stream) {
  HB_REQUIRE(ground_truth_mma_supported(d_data, d_queries, n, nq, k), HB_ERR_UNSUPPORTED,
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
  p.n_tiles = (uint32_t)((n + kTileN - 1) / kTileN);
  // chunks of data tiles: enough items to fill the grid several times over, picked so the last wave is nearly full
  uint32_t best_chunks = 1;
  double best_eff = -1.0;
  const uint32_t max_chunks = std::min<uint32_t>(p.n_tiles, 96);
  for (uint32_t c = 1; c <= max_chunks; ++c) {
    const uint32_t tpc = (p.n_tiles + c - 1) / c;
    const uint32_t chunks = (p.n_tiles + tpc - 1) / tpc;
    const uint64_t items = (uint64_t)chunks * p.n_qtiles;
    const uint64_t waves = (items + sms - 1) / sms;
    // cost model: waves x (tiles per chunk + per-item overhead: writing the query tile and refilling the item's lists,
    // in tile times -- register lists (k <= 16) refill in ~25, lists in global memory in ~4 per entry)
    const double overhead = 2.0 + (k <= 16 ? 25.0 : 0.25 * k);
    const double cost = (double)waves * (tpc + overhead);
    const double eff = ((double)p.n_tiles * p.n_qtiles / sms) / cost;
    if (eff > best_eff + 1e-9) {
      best_eff = eff;
      best_chunks = chunks;
    }
  }
  p.n_chunks = best_chunks;
  p.tiles_per_chunk = (p.n_tiles + p.n_chunks - 1) / p.n_chunks;
  p.n_chunks = (p.n_tiles + p.tiles_per_chunk - 1) / p.tiles_per_chunk;

  uint64_t* partial = nullptr;
  p.lcap = k <= 16 ? k : 2 * k;
  const size_t list_bytes = (size_t)p.n_chunks * nq * p.lcap * 8;
  HB_CUDA(cudaMallocAsync(&partial, list_bytes + nq * 4, stream));
  p.partial = partial;
  p.gthr = reinterpret_cast<uint32_t*>(reinterpret_cast<unsigned char*>(partial) + list_bytes);
  HB_CUDA(cudaMemsetAsync(p.gthr, 0xFF, nq * 4, stream));
  const uint32_t n_items = p.n_qtiles * p.n_chunks;
  const uint32_t grid = std::min<uint32_t>(n_items, (uint32_t)sms);
  if (k <= 16) {
    HB_CUDA(cudaFuncSetAttribute(gt_mma_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    gt_mma_kernel<true><<<grid, kThreads, kSmemBytes, stream>>>(p);
  } else {
    HB_CUDA(cudaFuncSetAttribute(gt_mma_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemBytes));
    gt_mma_kernel<false><<<grid, kThreads, kSmemBytes, stream>>>(p);
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
