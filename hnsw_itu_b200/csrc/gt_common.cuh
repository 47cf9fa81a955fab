// gt_common.cuh -- tcgen05 / mbarrier / TMEM building blocks shared by the ground-truth kernels (gt_mxf4.cu: E2M1
// operands, gt_tcgen05.cu: int8 operands) and the tensor-pipe peak probe (gt_peak.cu).  sm_100a only.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace hb {
namespace gt {

constexpr uint64_t kNone = ~0ull;  // "no element" in a packed (distance << 32 | key) list

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Bounded spin: a protocol bug traps (the launch fails with an error) instead of hanging the GPU.
// SUSPEND_NS > 0: try_wait carries a suspend-time hint (the warp sleeps in the barrier unit instead of spinning) and
// the clock is sampled every 16 spins; 0: plain try_wait, clock every 1024 spins.
// CLUSTER: acquire at cluster scope (the arrivals come from the peer CTA of a pair as well).
template <uint32_t SUSPEND_NS, bool CLUSTER = false>
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  long long t0 = 0;
  constexpr uint32_t kEvery = SUSPEND_NS ? 15u : 1023u;
  for (uint32_t spins = 0; !done; ++spins) {
    if (SUSPEND_NS && CLUSTER) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2, %3;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(parity), "r"(SUSPEND_NS)
          : "memory");
    } else if (CLUSTER) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    } else if (SUSPEND_NS) {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(parity), "r"(SUSPEND_NS)
          : "memory");
    } else {
      asm volatile(
          "{\n\t.reg .pred p;\n\t"
          "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
          "selp.u32 %0, 1, 0, p;\n\t}"
          : "=r"(done)
          : "r"(bar), "r"(parity)
          : "memory");
    }
    if (!done && (spins & kEvery) == kEvery) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 6000000000ll) __trap();  // ~3 s: no wait in these kernels is longer than a few ms
    }
  }
}

// K-major, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart (cute::UMMA::SmemDescriptor layout, version 1)
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// tcgen05.mma.cta_group::1.kind::i8: D (s32, TMEM) (+)= A (smem, s8) x B (smem, s8)
__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// tcgen05.mma.cta_group::1.kind::mxf4.block_scale: D (f32, TMEM) (+)= A (E2M1) x B (E2M1), E8M0 scales of both read at tmem_sf
__device__ __forceinline__ void umma_mxf4(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t tmem_sf, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%5], p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(tmem_sf)
      : "memory");
}
// one lane of the (converged) warp; ptxas knows the guarded region is single-threaded, so the MMA needs no lane loop
__device__ __forceinline__ uint32_t elect_one_sync() {
  uint32_t pred = 0;
  asm volatile(
      "{\n\t.reg .b32 rx;\n\t.reg .pred px;\n\t"
      "elect.sync rx|px, %1;\n\t"
      "@px mov.s32 %0, 1;\n\t}"
      : "+r"(pred)
      : "r"(0xFFFFFFFFu));
  return pred;
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// ---- CTA pairs (cta_group::2): two CTAs of a cluster on one TPC issue one MMA over both SMs ----
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// the shared::cluster address of `addr` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// arrive on a barrier anywhere in the cluster; release at cluster scope (what this thread wrote is visible to the waiter)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {  // every thread of every CTA of the cluster
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// tcgen05.mma.cta_group::2.kind::mxf4.block_scale: M = 256 over the pair (each CTA's A rows and accumulator lanes are
// its own 128), B's N rows split in halves between the two CTAs' shared memory at the same offsets.  Leader CTA only.
__device__ __forceinline__ void umma_mxf4_pair(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t tmem_sf, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::mxf4.block_scale.block32 [%0], %1, %2, %3, [%5], [%5], p;\n\t}" ::"r"(tmem_d),
      "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "r"(tmem_sf)
      : "memory");
}
// the pair's commit: arrives on the barrier at this offset in both CTAs
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3)
               : "memory");
}

#define HB_TMEM_LD32(v, taddr)                                                                                          \
  asm volatile(                                                                                                         \
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "                                                                         \
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "                                         \
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"                         \
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),     \
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),          \
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),         \
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])                       \
      : "r"(taddr)                                                                                                      \
      : "memory")

// k > 16: a row's candidates go to an UNORDERED append buffer of 2k slots in global memory (one store per accepted
// row; the merge kernel needs no order).  When a buffer fills, the whole warp compacts it: the k-th smallest packed
// value V is found by bisection on the 43 significant bits (every lane counts its share, REDUX adds), entries <= V
// move to the front, the tail is cleared, and the row's threshold becomes V's distance.  Amortised over k appends,
// against a dependent chain of loads per insert into a sorted list.  `buf` and the result are warp-uniform.
static __device__ __noinline__ uint64_t warp_compact(uint64_t* buf, uint32_t cap, uint32_t k) {
  const uint32_t lane = threadIdx.x & 31;
  uint64_t lo = 0, hi = (1025ull << 32);  // smallest V with count(e <= V) >= k lies in [lo, hi]
  while (lo < hi) {
    const uint64_t mid = lo + ((hi - lo) >> 1);
    uint32_t c = 0;
    for (uint32_t i = lane; i < cap; i += 32) c += (buf[i] <= mid) ? 1u : 0u;
    c = __reduce_add_sync(0xFFFFFFFFu, c);
    if (c >= k) hi = mid;
    else lo = mid + 1;
  }
  uint32_t woff = 0;
  for (uint32_t base = 0; base < cap; base += 32) {
    const uint32_t i = base + lane;
    const uint64_t e = (i < cap) ? buf[i] : kNone;
    const uint32_t m = __ballot_sync(0xFFFFFFFFu, e <= lo);
    if (e <= lo) buf[woff + __popc(m & ((1u << lane) - 1u))] = e;  // woff + rank <= i: never ahead of the reads
    woff += __popc(m);
    __syncwarp();
  }
  for (uint32_t i = woff + lane; i < cap; i += 32) buf[i] = kNone;
  __syncwarp();
  return lo;
}

}  // namespace gt
}  // namespace hb
