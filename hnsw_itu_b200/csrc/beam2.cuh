// beam2.cuh -- second-generation warp-level beam search of the query path (sm_100a).
//
// Same algorithm and same results as beam.cuh (one warp runs one nsw::search, src/index/nsw.rs:133-184; bounded beam +
// boundary-tie stack + lossy recently-seen cache, proven against the unbounded restatement in
// tests/test_bounded_model.py), re-implemented to spend fewer issue slots per expansion -- the round-1 kernel was
// issue-bound (profiles/r01i_search_kernel_ncu_summary.json: 77 k warp instructions per query):
//
//   * BeamSplit<R>: the beam as two 32-bit register arrays (distance << 1 | expanded, key) instead of one packed
//     64-bit array: membership is one 32-bit compare per slot, the maximum / minimum are REDUX on the distance word
//     followed by REDUX on the keys that tie, and no 64-bit compare/select chains remain.
//   * hamming_rows(): 4 lanes per sketch row, each lane loads 2 x 16 B (bytes [16c,16c+16) and [64+16c,64+16c+16) of
//     the row, so every load instruction covers whole 32-B sectors) straight into registers -- no shared-memory
//     stage, no LDGSTS (ptxas pads every LDGSTS with three dummy LDS on sm_100a), and a 2-round butterfly instead of
//     3 rounds: ~25 instructions per 8 rows against ~59 before.
//   * SeenCache<TAG16>: the recently-seen cache with 16-bit tags.  Set index = low bits of the gid, tag = the remaining
//     high bits: the (set, tag) pair is a bijection of the gid, so a hit is still exact (never a false "seen") while
//     n_total < 0xFFFF << log2(sets); twice the entries per byte, and the 16 KB per CTA the stage used to take now
//     hold 4x the entries (2 048 per warp).  TAG16 = false keeps full 32-bit ids for larger indexes.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

#include "beam.cuh"

namespace hb {

constexpr uint32_t kIdsPad = 8;                       // ids are padded to a multiple of 8 rows with a valid gid
constexpr uint32_t kIdsWords2 = kIdsCap + kIdsPad;    // staged ids of one two-line segment + padding
constexpr uint32_t kEmpty32 = 0xFFFFFFFFu;

// how a sketch piece is loaded: 0 = ld.global.nc (L1 allocating), 1 = ld.global.nc.L1::no_allocate
#ifndef HB_LD_MODE
#define HB_LD_MODE 1
#endif
__device__ __forceinline__ uint4 ld_sketch16(const uint4* p) {
  uint4 v;
#if HB_LD_MODE == 1
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
#else
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
#endif
  return v;
}

// the query as this lane sees it: pieces c and 4 + c of the 8 x 16-B row (c = lane % 4)
struct QueryRegs {
  uint4 a, b;
};
__device__ __forceinline__ QueryRegs load_query(const uint4* __restrict__ queries, uint32_t q) {
  const uint32_t c = threadIdx.x & 3u;
  QueryRegs r;
  r.a = __ldg(queries + (size_t)q * 8 + c);
  r.b = __ldg(queries + (size_t)q * 8 + 4 + c);
  return r;
}

// Sketch::distance (cli/src/sketch.rs:18-29) of one row to the query; every lane returns the distance
__device__ __forceinline__ uint32_t hamming_one(const uint4* __restrict__ sk, uint32_t gid, const QueryRegs& q) {
  const uint32_t c = threadIdx.x & 3u;
  const uint4* src = sk + (size_t)gid * 8 + c;
  uint32_t d = popc_xor4(ld_sketch16(src), q.a) + popc_xor4(ld_sketch16(src + 4), q.b);
  d += __shfl_xor_sync(kFullMask, d, 1);
  d += __shfl_xor_sync(kFullMask, d, 2);
  return d;
}

// Distances of the n (1..32) sketches whose gids are staged in ids[0..n) to the query: lane i returns the distance of
// ids[i] (garbage for i >= n).  ids[] must be readable (valid gids) up to the next multiple of 8 past n.
// Lane (r8 = lane / 4, c = lane % 4) handles pieces c and 4 + c of rows 8 p + r8, p = 0 .. ceil(n / 8) - 1.
__device__ __forceinline__ uint32_t hamming_rows(const uint4* __restrict__ sk, const uint32_t* ids, uint32_t n, const QueryRegs& q) {
  const uint32_t lane = threadIdx.x & 31, r8 = lane >> 2, c = lane & 3u;
  uint4 a[4], b[4];
#pragma unroll
  for (int p = 0; p < 4; ++p)
    if (8u * p < n) {
      const uint4* src = sk + (size_t)ids[8u * p + r8] * 8 + c;
      a[p] = ld_sketch16(src);
      b[p] = ld_sketch16(src + 4);
    }
  uint32_t x[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) x[p] = (8u * p < n) ? popc_xor4(a[p], q.a) + popc_xor4(b[p], q.b) : 0u;
  // butterfly over the 4 lanes of a row: afterwards lane (r8, c) holds the full sum of row 8 c + r8
  const bool b0 = c & 1u, b1 = c & 2u;
  const uint32_t k0 = b0 ? x[1] : x[0], s0 = b0 ? x[0] : x[1];
  const uint32_t k1 = b0 ? x[3] : x[2], s1 = b0 ? x[2] : x[3];
  const uint32_t y0 = k0 + __shfl_xor_sync(kFullMask, s0, 1);  // pass b0
  const uint32_t y1 = k1 + __shfl_xor_sync(kFullMask, s1, 1);  // pass 2 + b0
  const uint32_t kk = b1 ? y1 : y0, ss = b1 ? y0 : y1;
  const uint32_t z = kk + __shfl_xor_sync(kFullMask, ss, 2);   // pass 2 b1 + b0 = c
  return __shfl_sync(kFullMask, z, ((lane & 7u) << 2) | (lane >> 3));  // row i sits in lane 4 (i % 8) + i / 8
}

// How the neighbour sketches travel: 1 (default) = cp.async into a per-warp stage of 32 rows, the next group of 32 issued
// before the accept replay of the current one; 0 = LDG into registers (hamming_rows above).  Measured on B200: a
// dependent 32-row gather by one warp costs ~2 050 cycles through cp.async against ~2 900 through LDG.128
// (profiles/r01c_gather_shape_microbench.log), and cp.async holds no registers while the rows are in flight, which is
// worth one more resident CTA per SM: 8.5-8.7 against 8.3 M queries/s (profiles/r02d_ab_gather_variants.log).
#ifndef HB_GATHER2
#define HB_GATHER2 1
#endif
constexpr uint32_t kStageRows2 = HB_GATHER2 == 1 ? 32u : 0u;
constexpr uint32_t kStageBytes2 = kStageRows2 * 128u;

// stage slot of (pass p, half h) for this lane: lane-private 16-B slots, consecutive lanes in consecutive slots
// (conflict-free LDS.128); every lane reads back exactly the pieces it copied, so cp.async.wait_group orders it all
__device__ __forceinline__ void gather_issue(const uint4* __restrict__ sk, const uint32_t* ids, uint32_t n, uint4* stage) {
  const uint32_t lane = threadIdx.x & 31, r8 = lane >> 2, c = lane & 3u;
  uint4* const mine = stage + lane;
#pragma unroll
  for (int p = 0; p < 4; ++p)
    if (8u * p < n) {
      const uint4* src = sk + (size_t)ids[8u * p + r8] * 8 + c;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(mine + (2 * p) * 32)), "l"(src) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(mine + (2 * p + 1) * 32)), "l"(src + 4) : "memory");
    }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ uint32_t gather_finish(uint32_t n, const QueryRegs& q, const uint4* stage) {
  const uint32_t lane = threadIdx.x & 31, c = lane & 3u;
  const uint4* const mine = stage + lane;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  uint32_t x[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) x[p] = (8u * p < n) ? popc_xor4(mine[(2 * p) * 32], q.a) + popc_xor4(mine[(2 * p + 1) * 32], q.b) : 0u;
  const bool b0 = c & 1u, b1 = c & 2u;
  const uint32_t k0 = b0 ? x[1] : x[0], s0 = b0 ? x[0] : x[1];
  const uint32_t k1 = b0 ? x[3] : x[2], s1 = b0 ? x[2] : x[3];
  const uint32_t y0 = k0 + __shfl_xor_sync(kFullMask, s0, 1);
  const uint32_t y1 = k1 + __shfl_xor_sync(kFullMask, s1, 1);
  const uint32_t kk = b1 ? y1 : y0, ss = b1 ? y0 : y1;
  const uint32_t z = kk + __shfl_xor_sync(kFullMask, ss, 2);
  return __shfl_sync(kFullMask, z, ((lane & 7u) << 2) | (lane >> 3));
}

// ---- recently-seen cache -----------------------------------------------------------------------------------------
// Two-entry sets, most recent first.  Lossy on purpose: a forgotten id is evaluated again and rejected by the beam's
// own membership test; an id is never reported seen unless it was stored (no false positives -- that would drop a
// node and break parity).
template <bool TAG16>
struct SeenCache;

template <>
struct SeenCache<true> {  // one 32-bit word per set = two 16-bit tags
  uint32_t* sets;
  uint32_t mask, shift;   // set = id & mask, tag = id >> shift
  static __host__ __device__ constexpr size_t bytes(uint32_t entries_log2) { return (size_t)2 << entries_log2; }
  // largest n_total the tags can represent exactly (tag 0xFFFF is the empty marker)
  static __host__ __device__ constexpr uint64_t max_ids(uint32_t entries_log2) { return (uint64_t)0xFFFFu << (entries_log2 - 1); }
  __device__ __forceinline__ void init(void* mem, uint32_t entries_log2) {
    sets = reinterpret_cast<uint32_t*>(mem);
    shift = entries_log2 - 1;
    mask = (1u << shift) - 1u;
  }
  __device__ __forceinline__ bool test_and_set(uint32_t id) const {
    uint32_t* s = sets + (id & mask);
    const uint32_t tag = id >> shift, e = *s;
    if ((e & 0xFFFFu) == tag || (e >> 16) == tag) return false;
    *s = (e << 16) | tag;
    return true;
  }
};

template <>
struct SeenCache<false> {  // one 64-bit word per set = two full ids (any n_total)
  uint2* sets;
  uint32_t shift;
  static __host__ __device__ constexpr size_t bytes(uint32_t entries_log2) { return (size_t)4 << entries_log2; }
  static __host__ __device__ constexpr uint64_t max_ids(uint32_t) { return ~0ull; }
  __device__ __forceinline__ void init(void* mem, uint32_t entries_log2) {
    sets = reinterpret_cast<uint2*>(mem);
    shift = 32u - (entries_log2 - 1);
  }
  __device__ __forceinline__ bool test_and_set(uint32_t id) const {
    uint2* s = sets + ((id * 0x9E3779B1u) >> shift);
    const uint2 e = *s;
    if (e.x == id || e.y == id) return false;
    *s = make_uint2(id, e.x);
    return true;
  }
};

__device__ __forceinline__ void fill_ff(void* mem, size_t bytes) {  // bytes is a multiple of 512
  const uint32_t lane = threadIdx.x & 31;
  uint4* m4 = reinterpret_cast<uint4*>(mem);
  const uint4 e4 = make_uint4(kEmpty32, kEmpty32, kEmpty32, kEmpty32);
  for (uint32_t i = lane; i < bytes / 16; i += 32) m4[i] = e4;
  __syncwarp();
}

// ---- the beam -------------------------------------------------------------------------------------------------------
// The reference's result heap `w` and the members of its candidate heap `cands` that are still in `w`
// (nsw.rs:145-146), as 32 R unsorted slots: slot i = register i / 32 of lane i % 32.  wd = distance << 1 | expanded
// (0xFFFFFFFF = empty: above every real value as unsigned, below as signed), wk = key.  An integer compare of
// (wd >> 1, wk) is the reference's Ord on Distance (src/index/mod.rs:90-97); keys are unique inside the beam.
template <int R>
struct BeamSplit {
  uint32_t wd[R], wk[R];
  uint32_t size;     // |w|
  uint32_t Fd;       // w.peek_max().distance, maintained while size >= ef (uniform across the warp)
  uint32_t spill_n;  // boundary-tie stack depth
  uint32_t spill_d;  // distance shared by the stack entries

  __device__ __forceinline__ void reset(uint32_t d0, uint32_t key) {
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll
    for (int r = 0; r < R; ++r) wd[r] = wk[r] = kEmpty32;
    if (lane == 0) {
      wd[0] = d0 << 1;
      wk[0] = key;
    }
    Fd = d0;
    size = 1;
    spill_n = 0;
    spill_d = 0;
  }
  __device__ __forceinline__ uint32_t F() const { return Fd; }

  __device__ __forceinline__ void refresh_F() {
    int m = (int)wd[0];
#pragma unroll
    for (int r = 1; r < R; ++r) m = max(m, (int)wd[r]);  // signed: empty slots (-1) lose
    Fd = (uint32_t)__reduce_max_sync(kFullMask, m) >> 1;
  }

  __device__ __forceinline__ bool contains(uint32_t key) const {
    bool hit = false;
#pragma unroll
    for (int r = 0; r < R; ++r) hit |= wk[r] == key;
    return __any_sync(kFullMask, hit);
  }

  // cands.pop_min() restricted to members of w: the smallest (distance, key) not yet expanded; marks it expanded.
  __device__ __forceinline__ bool pop_unexpanded(uint32_t& key) {
    uint32_t cd = kEmpty32;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (!(wd[r] & 1u)) cd = min(cd, wd[r]);
    const uint32_t md = __reduce_min_sync(kFullMask, cd);
    if (md == kEmpty32) return false;
    uint32_t ck = kEmpty32;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (wd[r] == md) ck = min(ck, wk[r]);
    const uint32_t mk = __reduce_min_sync(kFullMask, ck);
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (wk[r] == mk) wd[r] |= 1u;
    key = mk;
    return true;
  }

  // the key pop_unexpanded() would return next if nothing were inserted in between (kNoKey: none); nothing is marked
  __device__ __forceinline__ uint32_t peek_unexpanded() const {
    uint32_t cd = kEmpty32;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (!(wd[r] & 1u)) cd = min(cd, wd[r]);
    const uint32_t md = __reduce_min_sync(kFullMask, cd);
    uint32_t ck = kEmpty32;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (wd[r] == md) ck = min(ck, wk[r]);
    return md == kEmpty32 ? kNoKey : __reduce_min_sync(kFullMask, ck);
  }

  // w.push((d, key)) while |w| < ef (nsw.rs:171-172)
  __device__ __forceinline__ void push(uint32_t d, uint32_t key, uint32_t ef) {
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if ((size >> 5) == (uint32_t)r && (size & 31u) == lane) {
        wd[r] = d << 1;
        wk[r] = key;
      }
    ++size;
    if (size >= ef) refresh_F();
  }

  // w.push((d, key)); w.pop_max() on a full beam (nsw.rs:171-176).  Precondition: key not in w, d < F.
  // Returns the evicted key; ev_unexpanded = the evicted element had not been expanded (it is still in `cands`).
  __device__ __forceinline__ uint32_t replace_max(uint32_t d, uint32_t key, bool& ev_unexpanded) {
    uint32_t lk = 0;  // key + 1 of this lane's largest key at distance F (0 = none)
#pragma unroll
    for (int r = 0; r < R; ++r)
      if ((wd[r] >> 1) == Fd) lk = max(lk, wk[r] + 1u);
    const uint32_t ek = __reduce_max_sync(kFullMask, lk) - 1u;
    bool unexp = false;
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (wk[r] == ek) {
        unexp = !(wd[r] & 1u);
        wd[r] = d << 1;
        wk[r] = key;
      }
    ev_unexpanded = __any_sync(kFullMask, unexp);
    refresh_F();
    return ek;
  }

  // The n smallest elements in ascending (distance, key) order: element j moves to register j / 32 of lane j % 32
  // (drain_asc().take(n), hnsw.rs:296-299); everything else is dropped.
  __device__ __forceinline__ void sort_prefix(uint32_t n) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t od[R], ok[R];
    n = min(n, size);
#pragma unroll
    for (int ro = 0; ro < R; ++ro) {
      od[ro] = ok[ro] = kEmpty32;
      for (uint32_t j = 32u * ro; j < min(n, 32u * ro + 32u); ++j) {
        uint32_t cd = kEmpty32;
#pragma unroll
        for (int r = 0; r < R; ++r) cd = min(cd, wd[r] >> 1 | (wd[r] == kEmpty32 ? 0x80000000u : 0u));
        const uint32_t md = __reduce_min_sync(kFullMask, cd);
        uint32_t ck = kEmpty32;
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (wd[r] != kEmpty32 && (wd[r] >> 1) == md) ck = min(ck, wk[r]);
        const uint32_t mk = __reduce_min_sync(kFullMask, ck);
#pragma unroll
        for (int r = 0; r < R; ++r)
          if (wk[r] == mk) wd[r] = wk[r] = kEmpty32;
        if ((j & 31u) == lane) {
          od[ro] = md;  // plain distance from here on
          ok[ro] = mk;
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      wd[r] = od[r];
      wk[r] = ok[r];
    }
  }
  // after sort_prefix: rank of this lane's register r, and (distance, key) of rank i broadcast to every lane
  __device__ __forceinline__ uint32_t sorted_index(int r) const { return 32u * r + (threadIdx.x & 31); }
  __device__ __forceinline__ void at(uint32_t i, uint32_t& d, uint32_t& key) const {
    const uint32_t r = i >> 5;
    uint32_t td = wd[0], tk = wk[0];
#pragma unroll
    for (int j = 1; j < R; ++j)
      if (r == (uint32_t)j) {
        td = wd[j];
        tk = wk[j];
      }
    d = __shfl_sync(kFullMask, td, i & 31u);
    key = __shfl_sync(kFullMask, tk, i & 31u);
  }
};

// per-warp scratch of the v2 kernel in shared memory: [stage kStageBytes2][ids kIdsWords2][spill kSpillCap2][cache]
template <bool TAG16>
struct WarpScratch2 {
  SeenCache<TAG16> cache;
  uint4* stage;
  uint32_t* ids;
  uint32_t* spill;
};
constexpr uint32_t kSpillCap2 = 56;  // 8 padding ids + 56 stack entries = 64 words
template <bool TAG16>
__host__ __device__ constexpr size_t warp_scratch2_bytes(uint32_t entries_log2) {
  return SeenCache<TAG16>::bytes(entries_log2) + (size_t)(kIdsWords2 + kSpillCap2) * 4 + kStageBytes2;
}

// -DHB_DEBUG_BOUNDS: every gid, degree and staged count is checked before it is used as an address; the first failure is
// recorded (code, value) and the query ends with status 3.  compute-sanitizer is not available on the GPU pool, this build
// run through the parity tests is the memory-safety evidence (profiles/r02*_debug_bounds*.log).
#ifdef HB_DEBUG_BOUNDS
#define HB_CHK2(cond, code, val)                                              \
  do {                                                                        \
    if (!__all_sync(kFullMask, (cond))) {                                     \
      if (lane == 0 && atomicCAS(gv.dbg, 0u, (uint32_t)(code)) == 0u) gv.dbg[1] = (uint32_t)(val); \
      status = 3;                                                             \
      return;                                                                 \
    }                                                                         \
  } while (0)
#else
#define HB_CHK2(cond, code, val) do {} while (0)
#endif

// nsw::search(graph, q, ef, ep) (nsw.rs:133-184) on the graph that contains gid `ep`.  On return b holds w.
// status != 0: the boundary-tie stack overflowed (the caller reruns the query on the large-capacity kernel).
// COUNT = false drops the X/E/V/A work counters (four registers and their updates) from the timed path.
template <int R, bool TAG16, bool COUNT>
__device__ __forceinline__ void beam_search2(const GraphView& gv, const QueryRegs& q, uint32_t ef, uint32_t ep,
                                             const WarpScratch2<TAG16>& ws, BeamSplit<R>& b, SearchCounters& c, uint32_t& status) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  HB_CHK2(ep < gv.n_total, 11, ep);
  {
    const uint32_t d0 = hamming_one(gv.sk, ep, q);  // nsw.rs:140-141
    if (lane == 0) ws.cache.test_and_set(ep);        // nsw.rs:143-144
    b.reset(d0, ep);                                  // nsw.rs:145-146
    if (COUNT) ++c.V;
    __syncwarp();
  }
  for (;;) {
    // ---- c = cands.pop_min(); if c.distance > F break  (nsw.rs:148-154) ----
    uint32_t ckey;
    if (!b.pop_unexpanded(ckey)) {
      if (b.spill_n && b.spill_d == b.F() && b.size >= ef) {
        uint32_t kk = 0;
        if (lane == 0) kk = ws.spill[b.spill_n - 1];
        ckey = __shfl_sync(kFullMask, kk, 0);
        --b.spill_n;
      } else {
        break;
      }
    }
    if (COUNT) ++c.X;
    HB_CHK2(ckey < gv.n_total, 12, ckey);
    // ---- for e in graph.neighborhood(c.key)  (nsw.rs:156) ----
    const uint32_t* row = gv.row(ckey);
    uint32_t w0, w1;
    w0 = __ldg(row + lane);
    w1 = __ldg(row + 32 + lane);
    const uint32_t deg = __shfl_sync(kFullMask, w0, 0);
    HB_CHK2(gv.rowoff ? deg < 65536u : deg < gv.stride, 13, deg);
    if (COUNT) c.E += deg;
    for (uint32_t base = 0; base <= deg; base += 64) {
      // rows longer than one segment (hubs): pull the next segment's two lines into L2 while this one is worked on
      if (base + 64 <= deg && lane < 2) prefetch_l2(row + base + 64 + lane * 32);
      // ---- nsw.rs:157-161 through the recently-seen cache; order-preserving compaction of the unseen ids ----
      uint32_t n_new;
      {
        const uint32_t wi = base + lane;
        const bool isnew = (wi >= 1u) && (wi <= deg) && w0 != kNoKey && ws.cache.test_and_set(w0);  // 0xFFFFFFFF: builder tombstone
        const uint32_t nm = __ballot_sync(kFullMask, isnew);
        if (isnew) ws.ids[__popc(nm & lt_mask)] = w0;
        n_new = __popc(nm);
      }
      if (base + 32 <= deg) {
        const uint32_t wi = base + 32 + lane;
        const bool isnew = (wi <= deg) && w1 != kNoKey && ws.cache.test_and_set(w1);
        const uint32_t nm = __ballot_sync(kFullMask, isnew);
        if (isnew) ws.ids[n_new + __popc(nm & lt_mask)] = w1;
        n_new += __popc(nm);
      }
      HB_CHK2(n_new <= kIdsCap, 14, n_new);
      if (lane < kIdsPad) ws.ids[n_new + lane] = ckey;  // padding rows: any valid gid
      if (COUNT) c.V += n_new;
      __syncwarp();
#if HB_GATHER2 == 1
      if (n_new) gather_issue(gv.sk, ws.ids, min(32u, n_new), ws.stage);
#endif
      for (uint32_t g0 = 0; g0 < n_new; g0 += 32) {
        const uint32_t n = min(32u, n_new - g0);
        const uint32_t myid = ws.ids[g0 + lane];
        HB_CHK2(lane >= n || myid < gv.n_total, 15, myid);
        // distance_fn(point, query) for the group (nsw.rs:164-165)
#if HB_GATHER2 == 1
        const uint32_t myd = gather_finish(n, q, ws.stage);
        if (g0 + 32u < n_new) gather_issue(gv.sk, ws.ids + g0 + 32, n_new - g0 - 32u, ws.stage);  // in flight during the replay
#else
        const uint32_t myd = hamming_rows(gv.sk, ws.ids + g0, n, q);
#endif
        // ---- sequential accept in neighbour order (nsw.rs:162-176) ----
        // pm = lanes that still have to be looked at and currently pass !(d >= F && |w| >= ef)
        uint32_t pm = __ballot_sync(kFullMask, lane < n && (b.size < ef || myd < b.F()));
        while (pm) {
          const int first = __ffs(pm) - 1;
          pm &= pm - 1u;
          const uint32_t nd = __shfl_sync(kFullMask, myd, first), nk = __shfl_sync(kFullMask, myid, first);
          if (b.contains(nk)) continue;  // a revisit the cache no longer remembered
          if (COUNT) ++c.A;
          if (b.size < ef) {
            b.push(nd, nk, ef);
            if (pm && b.size >= ef) pm &= __ballot_sync(kFullMask, myd < b.F());  // the beam just filled: F now binds
            continue;
          }
          const uint32_t f_old = b.F();
          bool ev_unexp;
          const uint32_t ek = b.replace_max(nd, nk, ev_unexp);
          if (b.F() != f_old) {
            if (pm) pm &= __ballot_sync(kFullMask, myd < b.F());
          } else if (ev_unexp) {
            // evicted before it was expanded and still tied with the worst distance: it stays a live candidate
            if (b.spill_n && b.spill_d != f_old) b.spill_n = 0;
            b.spill_d = f_old;
            if (b.spill_n < kSpillCap2) {
              if (lane == 0) ws.spill[b.spill_n] = ek;
              ++b.spill_n;
            } else {
              status = 2;
            }
          }
        }
      }
      __syncwarp();  // ws.ids is rewritten by the next segment
      if (base + 64 <= deg) {
        w0 = (base + 64 + lane <= deg) ? __ldg(row + base + 64 + lane) : 0u;
        w1 = (base + 96 + lane <= deg) ? __ldg(row + base + 96 + lane) : 0u;
      }
    }
    if (status) break;
  }
}

}  // namespace hb
