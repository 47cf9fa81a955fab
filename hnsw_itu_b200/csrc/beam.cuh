// beam.cuh -- warp-level building blocks of the batched beam search (sm_100a), first generation.  The query path moved to
// beam2.cuh in round 2 (split 32-bit beam registers, 4-lane gather, 16-bit-tag cache); this file now serves the device
// builder (builder.cu: BeamRegs + beam_search with ef = ef_construction) and holds the types both share (GraphView, ...).
//
// One warp runs one nsw::search (src/index/nsw.rs:133-184).  The pieces:
//
//   * BeamRegs<R>: the reference's result heap `w` AND its candidate heap `cands`, held as ONE sorted
//     array of 32*R packed elements in REGISTERS (element i lives in lane i / R, register i % R).
//     An element is (distance << 33 | key << 1 | expanded); an integer compare of two elements is
//     the reference's Ord on Distance (src/index/mod.rs:90-97) because keys are unique.
//     "cands.pop_min()" = first element whose expanded bit is clear.  An element evicted from a full
//     beam can only be expanded later if its distance equals the then-current worst distance F (the
//     break is strict, nsw.rs:152), and such elements leave in decreasing order, so a STACK of keys
//     sharing one distance replays pop_min over them exactly (oracle: search_bounded).
//   * a direct-mapped cache of recently seen ids in shared memory instead of the reference's HashSet
//     (nsw.rs:143,157-161).  A hit is certainly visited and is skipped like the reference does.  A miss
//     may be a revisit; it costs one redundant distance evaluation but cannot change the result: a
//     visited node that is no longer in w was rejected or evicted while the beam was full, so its
//     distance is >= every later F and it is rejected again; a visited node still in w is found by the
//     "already in w" compare before the insert.  X, E, A and the results equal the reference's; only
//     V (distance evaluations performed) can exceed |visited|.
//   * hamming32(): distances of up to 32 sketches to the query.  A 128-B sketch row is read by 8 lanes
//     x LDG.128 (4 rows per instruction, up to 8 independent instructions in flight per lane),
//     XOR+POPC per lane, then a 3-round butterfly that reduces the 8x8 partial sums with 7 shuffles.
//   * the reference's accept / evict rule (nsw.rs:162-176) replayed IN NEIGHBOUR ORDER over the 32
//     distances: ballot(d < F or beam not full), take the first set lane, insert, refresh F, re-ballot.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>

namespace hb {

constexpr uint32_t kFullMask = 0xFFFFFFFFu;
constexpr uint64_t kBeamInf = ~0ull;  // empty slot: larger than any element, expanded bit set
constexpr uint32_t kIdsCap = 64;      // ids staged per segment = two 128-B lines of an adjacency row
constexpr uint32_t kNoKey = 0xFFFFFFFFu;

__device__ __forceinline__ uint64_t bm_pack(uint32_t d, uint32_t key) { return ((uint64_t)d << 33) | ((uint64_t)key << 1); }
__device__ __forceinline__ uint32_t bm_dist(uint64_t v) { return (uint32_t)(v >> 33); }
__device__ __forceinline__ uint32_t bm_key(uint64_t v) { return (uint32_t)(v >> 1); }

__device__ __forceinline__ uint32_t popc_xor4(uint4 a, uint4 b) {
  return __popc(a.x ^ b.x) + __popc(a.y ^ b.y) + __popc(a.z ^ b.z) + __popc(a.w ^ b.w);
}

// read-only view of the device index (index.cu): one id space over all graphs
struct GraphView {
  const uint4* sk;       // n_total rows of 128 B
  const uint32_t* adj;   // n_total rows of `stride` words: [0] = degree, then neighbour gids in canonical order
  uint32_t stride;
  const uint32_t* rowoff;  // compact layout: first 128-B line of every row (nullptr: fixed stride)
#ifdef HB_DEBUG_BOUNDS
  uint32_t* dbg;       // [0] first failed check (0 = none), [1] offending value, [2] query
  uint32_t n_total;
#endif
  __device__ __forceinline__ const uint32_t* row(uint32_t gid) const {
    return rowoff ? adj + (size_t)__ldg(rowoff + gid) * 32 : adj + (size_t)gid * stride;
  }
};

struct SearchCounters {
  uint32_t X = 0, E = 0, V = 0, A = 0;
};

// How hamming32() fetches sketch rows: 1 = cp.async (LDGSTS) 16 B per lane into a 4-KB shared-memory stage, then LDS
// (measured on B200, profiles/r01c_gather_shape_microbench.log: a dependent 32-row gather costs 2 040 cycles this way
// against 2 880 with eight LDG.128 per lane, and it needs no 32 data registers); 0 = LDG.128 into registers.
#ifndef HB_GATHER_CPASYNC
#define HB_GATHER_CPASYNC 1
#endif
constexpr uint32_t kStageBytes = HB_GATHER_CPASYNC ? 32u * 128u : 0u;
// 1 = the next group's sketch gather is issued before the accept replay of the current group (same results)
#ifndef HB_OVERLAP_REPLAY
#define HB_OVERLAP_REPLAY 1
#endif

// per-warp scratch in shared memory: [stage 4 KB][cache 4 << cache_log2][ids kIdsCap * 4]
struct WarpScratch {
  uint4* stage;          // 32 sketch rows (cp.async destination); lane (g, s) only ever touches rows 4j+g, piece s
  uint32_t* cache;       // 1 << cache_log2 recently seen ids (0xFFFFFFFF = empty)
  uint32_t cache_shift;  // 32 - cache_log2
  uint32_t* ids;         // kIdsCap staged neighbour ids (one two-line segment of an adjacency row, compacted)
};
__host__ __device__ constexpr size_t warp_scratch_bytes(uint32_t cache_log2) {
  return (size_t)kStageBytes + ((size_t)4 << cache_log2) + (size_t)kIdsCap * 4;
}
__device__ __forceinline__ WarpScratch make_warp_scratch(unsigned char* mine, uint32_t cache_log2) {
  WarpScratch ws;
  ws.stage = reinterpret_cast<uint4*>(mine);
  ws.cache = reinterpret_cast<uint32_t*>(mine + kStageBytes);
  ws.cache_shift = 32u - cache_log2;
  ws.ids = reinterpret_cast<uint32_t*>(mine + kStageBytes + ((size_t)4 << cache_log2));
  return ws;
}

template <int R>
struct BeamRegs {
  uint64_t w[R];
  uint64_t maxv;     // element at index size-1 (uniform across the warp)
  uint32_t size;     // |w|
  uint32_t spill_n;  // boundary-tie stack depth
  uint32_t spill_d;  // distance shared by the stack entries

  __device__ __forceinline__ void reset(uint64_t first) {
    const uint32_t lane = threadIdx.x & 31;
#pragma unroll
    for (int r = 0; r < R; ++r) w[r] = kBeamInf;
    if (lane == 0) w[0] = first;
    maxv = first;
    size = 1;
    spill_n = 0;
    spill_d = 0;
  }
  __device__ __forceinline__ uint32_t F() const { return bm_dist(maxv); }

  // element at index i (uniform i) broadcast to every lane
  __device__ __forceinline__ uint64_t at(uint32_t i) const {
    const uint32_t r = i % R;
    uint64_t t = w[0];
#pragma unroll
    for (int j = 1; j < R; ++j)
      if (r == (uint32_t)j) t = w[j];
    return __shfl_sync(kFullMask, t, i / R);
  }

  // is an element with this key (any flag) already in w?
  __device__ __forceinline__ bool contains(uint64_t v) const {
    bool hit = false;
#pragma unroll
    for (int r = 0; r < R; ++r) hit |= ((w[r] ^ v) >> 1) == 0;
    return __any_sync(kFullMask, hit);
  }

  // cands.pop_min() restricted to members of w: first element not yet expanded; marks it expanded.
  __device__ __forceinline__ bool pop_unexpanded(uint32_t& key) {
    uint64_t cv = kBeamInf;
#pragma unroll
    for (int r = R - 1; r >= 0; --r)
      if (!(w[r] & 1ull)) cv = w[r];
    const uint32_t m = __ballot_sync(kFullMask, cv != kBeamInf);
    if (!m) return false;
    const uint64_t c = __shfl_sync(kFullMask, cv, __ffs(m) - 1);
#pragma unroll
    for (int r = 0; r < R; ++r)
      if (w[r] == c) w[r] = c | 1ull;
    if (maxv == c) maxv = c | 1ull;
    key = bm_key(c);
    return true;
  }

  // the element pop_unexpanded() would return next, without marking it (used to predict the next expansion)
  __device__ __forceinline__ bool peek_unexpanded(uint32_t& key) const {
    uint64_t cv = kBeamInf;
#pragma unroll
    for (int r = R - 1; r >= 0; --r)
      if (!(w[r] & 1ull)) cv = w[r];
    const uint32_t m = __ballot_sync(kFullMask, cv != kBeamInf);
    if (!m) return false;
    key = __shfl_sync(kFullMask, bm_key(cv), __ffs(m) - 1);
    return true;
  }

  // the next two elements pop_unexpanded() would return if nothing were inserted in between (kNoKey = none)
  __device__ __forceinline__ void peek_unexpanded2(uint32_t& k1, uint32_t& k2) const {
    uint64_t c1 = kBeamInf;
#pragma unroll
    for (int r = R - 1; r >= 0; --r)
      if (!(w[r] & 1ull)) c1 = w[r];
    const uint32_t m = __ballot_sync(kFullMask, c1 != kBeamInf);
    k1 = k2 = kNoKey;
    if (!m) return;
    const uint32_t l1 = __ffs(m) - 1;
    const uint64_t first = __shfl_sync(kFullMask, c1, l1);
    k1 = bm_key(first);
    uint64_t c2 = kBeamInf;
#pragma unroll
    for (int r = R - 1; r >= 0; --r)
      if (!(w[r] & 1ull) && w[r] > first) c2 = w[r];
    const uint32_t m2 = __ballot_sync(kFullMask, c2 != kBeamInf);
    if (!m2) return;
    k2 = __shfl_sync(kFullMask, bm_key(c2), __ffs(m2) - 1);
  }

  __device__ __forceinline__ void sort_prefix(uint32_t) {}  // already sorted
  __device__ __forceinline__ uint32_t sorted_index(int r) const { return (threadIdx.x & 31) * R + r; }

  // w.push(v); if |w| > ef: w.pop_max()  (nsw.rs:171-176).  Precondition: v not in w, and (size < ef or v < maxv).
  // Returns the evicted element (kBeamInf when nothing was evicted).
  __device__ __forceinline__ uint64_t insert(uint64_t v, uint32_t ef) {
    const uint32_t lane = threadIdx.x & 31;
    uint32_t lt = 0;
#pragma unroll
    for (int r = 0; r < R; ++r) lt += (w[r] < v) ? 1u : 0u;
    const uint32_t m = __ballot_sync(kFullMask, lt < (uint32_t)R);  // never empty: an element >= v or an empty slot exists
    const uint32_t L = __ffs(m) - 1;
    const uint32_t pos = L * R + __shfl_sync(kFullMask, lt, L);
    const uint64_t up = __shfl_up_sync(kFullMask, w[R - 1], 1);
    const bool full = size >= ef;
#pragma unroll
    for (int r = R - 1; r >= 0; --r) {
      const uint32_t idx = lane * R + r;
      const uint64_t prev = (r == 0) ? up : w[r > 0 ? r - 1 : 0];
      uint64_t nv = (idx > pos) ? prev : (idx == pos ? v : w[r]);
      if (full && idx == ef) nv = kBeamInf;  // the old maximum moved here: pop_max()
      w[r] = nv;
    }
    uint64_t ev = kBeamInf;
    if (full) ev = maxv; else ++size;
    maxv = at(size - 1);
    return ev;
  }
};

#if HB_GATHER_CPASYNC
// hamming32() in two halves, so that the accept replay of one group of 32 runs while the next group's rows are in
// flight: issue = the cp.async copies (every lane copies exactly the 16-B pieces it reads back), finish = wait, XOR+POPC,
// butterfly.  finish() has consumed the stage (its shuffles need every piece) before the next issue() is reached.
__device__ __forceinline__ void hamming32_issue(const uint4* __restrict__ sk, const uint32_t* ids, uint32_t n, uint4* stage) {
  const uint32_t lane = threadIdx.x & 31, g = lane >> 3, s = lane & 7u;
  uint4* const mine = stage + g * 8 + s;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (4u * j < n) {
      const uint4* src = sk + (size_t)ids[min(4u * j + g, n - 1u)] * 8 + s;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(mine + j * 32)), "l"(src) : "memory");
    }
  asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ uint32_t hamming32_finish(uint32_t n, uint4 qv, const uint4* stage) {
  const uint32_t lane = threadIdx.x & 31, g = lane >> 3, s = lane & 7u;
  const uint4* const mine = stage + g * 8 + s;
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  uint32_t x[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) x[j] = (4u * j < n) ? popc_xor4(mine[j * 32], qv) : 0u;
  const bool b0 = s & 1u, b1 = s & 2u, b2 = s & 4u;
  uint32_t y[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t keep = b0 ? x[2 * i + 1] : x[2 * i], send = b0 ? x[2 * i] : x[2 * i + 1];
    y[i] = keep + __shfl_xor_sync(kFullMask, send, 1);
  }
  uint32_t z[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const uint32_t keep = b1 ? y[2 * i + 1] : y[2 * i], send = b1 ? y[2 * i] : y[2 * i + 1];
    z[i] = keep + __shfl_xor_sync(kFullMask, send, 2);
  }
  const uint32_t keep = b2 ? z[1] : z[0], send = b2 ? z[0] : z[1];
  const uint32_t t = keep + __shfl_xor_sync(kFullMask, send, 4);
  return __shfl_sync(kFullMask, t, 8u * (lane & 3u) + (lane >> 2));  // row i sits in lane 8 * (i % 4) + i / 4
}
#endif

// Distances of the n (1..32) sketches whose gids are staged in ids[0..n) to the query piece qv:
// lane i returns the distance of ids[i] (garbage for i >= n).  Lane = (g = lane / 8, s = lane % 8) owns the
// s-th 16-B piece of rows 4 * j + g, j = 0..7.
__device__ __forceinline__ uint32_t hamming32(const uint4* __restrict__ sk, const uint32_t* ids, uint32_t n, uint4 qv, uint4* stage) {
  const uint32_t lane = threadIdx.x & 31, g = lane >> 3, s = lane & 7u;
  uint32_t x[8];
#if HB_GATHER_CPASYNC
  // every lane copies exactly the 16-B pieces it reads back below, so cp.async.wait_group is the only ordering needed
  uint4* const mine = stage + g * 8 + s;
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (4u * j < n) {
      const uint4* src = sk + (size_t)ids[min(4u * j + g, n - 1u)] * 8 + s;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(mine + j * 32)), "l"(src) : "memory");
    }
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 8; ++j) x[j] = (4u * j < n) ? popc_xor4(mine[j * 32], qv) : 0u;
#else
  uint4 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (4u * j < n) v[j] = __ldg(sk + (size_t)ids[min(4u * j + g, n - 1u)] * 8 + s);
#pragma unroll
  for (int j = 0; j < 8; ++j) x[j] = (4u * j < n) ? popc_xor4(v[j], qv) : 0u;
#endif
  // butterfly: after the three rounds lane (g, s) holds the full sum of row 4 * s + g
  const bool b0 = s & 1u, b1 = s & 2u, b2 = s & 4u;
  uint32_t y[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint32_t keep = b0 ? x[2 * i + 1] : x[2 * i], send = b0 ? x[2 * i] : x[2 * i + 1];
    y[i] = keep + __shfl_xor_sync(kFullMask, send, 1);
  }
  uint32_t z[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const uint32_t keep = b1 ? y[2 * i + 1] : y[2 * i], send = b1 ? y[2 * i] : y[2 * i + 1];
    z[i] = keep + __shfl_xor_sync(kFullMask, send, 2);
  }
  const uint32_t keep = b2 ? z[1] : z[0], send = b2 ? z[0] : z[1];
  const uint32_t t = keep + __shfl_xor_sync(kFullMask, send, 4);
  // row i sits in lane 8 * (i % 4) + i / 4
  return __shfl_sync(kFullMask, t, 8u * (lane & 3u) + (lane >> 2));
}

__device__ __forceinline__ uint32_t hamming1(const uint4* __restrict__ sk, uint32_t gid, uint4 qv) {
  const uint32_t s = threadIdx.x & 7u;
  uint32_t d = popc_xor4(__ldg(sk + (size_t)gid * 8 + s), qv);
  d += __shfl_xor_sync(kFullMask, d, 1);
  d += __shfl_xor_sync(kFullMask, d, 2);
  d += __shfl_xor_sync(kFullMask, d, 4);
  return d;
}

__device__ __forceinline__ void cache_clear(const WarpScratch& ws) {
  const uint32_t lane = threadIdx.x & 31;
  uint4* c4 = reinterpret_cast<uint4*>(ws.cache);
  const uint4 e4 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
  const uint32_t n4 = (1u << (32u - ws.cache_shift)) >> 2;
  for (uint32_t i = lane; i < n4; i += 32) c4[i] = e4;
  __syncwarp();
}

// ---- prefetch policy (compile-time switches so variants can be A/B-timed on the B200) ----
//   HB_PREFETCH_ROW: 0 = off; 1 = the adjacency row of the predicted next candidate is loaded into registers
//                    one expansion ahead; 2 = additionally the row of the predicted second-next candidate is
//                    prefetched into L2 (so that the register load one expansion later is an L2 hit)
//   HB_PREFETCH_SK : 0 = off; n = n L2 prefetches (32-B sector stride 128/n B) per sketch row of the predicted next
//                    candidate's neighbours
#ifndef HB_PREFETCH_ROW
#define HB_PREFETCH_ROW 0
#endif
#ifndef HB_PREFETCH_SK
#define HB_PREFETCH_SK 0
#endif

// HB_DIAG_CLOCK (diagnostics build only): the counters are replaced by cycles spent per phase
// X = waiting for the adjacency row, E = cache check + compaction, V = sketch gather + reduction, A = accept replay
#ifdef HB_DIAG_CLOCK
#define HB_TICK(var) const long long var = clock64()
#define HB_ACC(field, t0, t1) (field) += (uint32_t)((t1) - (t0))
#else
#define HB_TICK(var) do {} while (0)
#define HB_ACC(field, t0, t1) do {} while (0)
#endif

__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

// Prefetch (into L2) the sketches of the neighbours held in the two row words w0 (row words 0..31) / w1 (32..63).
__device__ __forceinline__ void prefetch_row_sketches(const uint4* __restrict__ sk, uint32_t w0, uint32_t w1) {
#if HB_PREFETCH_SK > 0
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t ndeg = __shfl_sync(kFullMask, w0, 0);
  const bool v0 = lane >= 1u && lane <= ndeg && w0 != kNoKey;
  const bool v1 = 32u + lane <= ndeg && w1 != kNoKey;
#pragma unroll
  for (int i = 0; i < HB_PREFETCH_SK; ++i) {
    if (v0) prefetch_l2(reinterpret_cast<const char*>(sk + (size_t)w0 * 8) + i * (128 / HB_PREFETCH_SK));
    if (v1) prefetch_l2(reinterpret_cast<const char*>(sk + (size_t)w1 * 8) + i * (128 / HB_PREFETCH_SK));
  }
#endif
}

// nsw::search(graph, q, ef, ep) (nsw.rs:133-184) on the graph that contains gid `ep`.  On return b holds w.
// status != 0: the boundary-tie stack overflowed (the caller reruns the query on the large-capacity path).
//
// One expansion = one adjacency row, handled in SEGMENTS of two 128-B lines (64 row words): both lines are loaded
// together, every neighbour is tested against the recently-seen cache, the unseen ids of the whole segment are
// compacted IN NEIGHBOUR ORDER into ws.ids, and dense groups of 32 go through hamming32() + the sequential accept
// replay.  Rows must have a stride of at least 64 words (index.cu guarantees it).
//
// Latency hiding: the reference's loop is a chain of dependent random reads (row -> sketches -> accept -> next row).
// After popping candidate c_t the beam already names the most likely c_{t+1} (its next unexpanded element) and
// c_{t+2}; c_{t+1}'s row is loaded into registers now and consumed one expansion later, c_{t+2}'s row is prefetched
// into L2, and once c_{t+1}'s row words have arrived the sketches of its neighbours are prefetched into L2.  A wrong
// prediction (a neighbour accepted in this expansion overtakes them) falls back to plain loads; results never
// depend on the prediction.
// The recently-seen cache: "was this id evaluated lately?"  Lossy on purpose (a forgotten id is evaluated again and
// rejected by the beam's own membership test); never claims an id it has not stored.  HB_CACHE_WAYS = 2: sets of two
// entries, most recent first (one 64-bit read, one 64-bit write), same number of entries as the direct-mapped table.
#ifndef HB_CACHE_WAYS
#define HB_CACHE_WAYS 2
#endif
__device__ __forceinline__ bool cache_test_and_set(const WarpScratch& ws, uint32_t id) {
#if HB_CACHE_WAYS == 2
  uint2* set = reinterpret_cast<uint2*>(ws.cache) + ((id * 0x9E3779B1u) >> (ws.cache_shift + 1));
  const uint2 e = *set;
  if (e.x == id || e.y == id) return false;
  *set = make_uint2(id, e.x);
  return true;
#else
  const uint32_t h = (id * 0x9E3779B1u) >> ws.cache_shift;
  const bool isnew = ws.cache[h] != id;
  if (isnew) ws.cache[h] = id;
  return isnew;
#endif
}

template <class Beam>
__device__ __forceinline__ void beam_search(const GraphView& gv, uint4 qv, uint32_t ef, uint32_t ep, const WarpScratch& ws,
                                            uint32_t* spill, uint32_t spill_cap, Beam& b, SearchCounters& c,
                                            uint32_t& status) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t lt_mask = (1u << lane) - 1u;
  {
    const uint32_t d0 = hamming1(gv.sk, ep, qv);  // nsw.rs:140-141
    if (lane == 0) cache_test_and_set(ws, ep);  // nsw.rs:143-144
    b.reset(bm_pack(d0, ep));                                           // nsw.rs:145-146
    ++c.V;
    __syncwarp();
  }
  uint32_t pkey = kNoKey, pw0 = 0, pw1 = 0;  // predicted next candidate and the first two lines of its row
  for (;;) {
    // ---- c = cands.pop_min(); if c.distance > F break  (nsw.rs:148-154) ----
    uint32_t ckey;
    if (!b.pop_unexpanded(ckey)) {
      if (b.spill_n && b.spill_d == b.F() && b.size >= ef) {
        uint32_t kk = 0;
        if (lane == 0) kk = spill[b.spill_n - 1];
        ckey = __shfl_sync(kFullMask, kk, 0);
        --b.spill_n;
      } else {
        break;
      }
    }
#ifndef HB_DIAG_CLOCK
    ++c.X;
#endif
    HB_TICK(t_row0);
    // ---- for e in graph.neighborhood(c.key)  (nsw.rs:156) ----
    const uint32_t* row = gv.adj + (size_t)ckey * gv.stride;
    uint32_t w0, w1;
    if (ckey == pkey) {
      w0 = pw0;
      w1 = pw1;
#ifdef HB_DIAG_PRED
      c.A += 1u << 16;  // diagnostics build: prediction hits in the high half of A
#endif
    } else {
      w0 = __ldg(row + lane);
      w1 = __ldg(row + 32 + lane);
    }
#if HB_PREFETCH_ROW > 0
    {
      uint32_t n1 = kNoKey, n2 = kNoKey;
      b.peek_unexpanded2(n1, n2);
      pkey = n1;
      if (n1 != kNoKey) {
        const uint32_t* nrow = gv.adj + (size_t)n1 * gv.stride;
        pw0 = __ldg(nrow + lane);
        pw1 = __ldg(nrow + 32 + lane);
      }
#if HB_PREFETCH_ROW > 1
      if (n2 != kNoKey && lane < 2) prefetch_l2(gv.adj + (size_t)n2 * gv.stride + lane * 32);
#endif
    }
#endif
    const uint32_t deg = __shfl_sync(kFullMask, w0, 0);
    HB_TICK(t_row1);
    HB_ACC(c.X, t_row0, t_row1);
#ifndef HB_DIAG_CLOCK
    c.E += deg;
#endif
    bool sk_done = (pkey == kNoKey);
    for (uint32_t base = 0; base <= deg; base += 64) {
      // rows longer than one segment: fetch the next segment's words before working on this one
      uint32_t nw0 = 0, nw1 = 0;
      if (base + 64 <= deg) {
        if (base + 64 + lane <= deg) nw0 = __ldg(row + base + 64 + lane);
        if (base + 96 + lane <= deg) nw1 = __ldg(row + base + 96 + lane);
      }
      // ---- nsw.rs:157-161 through the recently-seen cache; order-preserving compaction of the unseen ids ----
      HB_TICK(t_chk0);
      uint32_t n_new;
      {
        const uint32_t wi = base + lane;
        const bool valid = (wi >= 1u) && (wi <= deg) && w0 != kNoKey;  // 0xFFFFFFFF: builder tombstone
        bool isnew = false;
        if (valid) {
          isnew = cache_test_and_set(ws, w0);
        }
        const uint32_t nm = __ballot_sync(kFullMask, isnew);
        if (isnew) ws.ids[__popc(nm & lt_mask)] = w0;
        n_new = __popc(nm);
      }
      if (base + 32 <= deg) {
        const uint32_t wi = base + 32 + lane;
        const bool valid = (wi <= deg) && w1 != kNoKey;
        bool isnew = false;
        if (valid) {
          isnew = cache_test_and_set(ws, w1);
        }
        const uint32_t nm = __ballot_sync(kFullMask, isnew);
        if (isnew) ws.ids[n_new + __popc(nm & lt_mask)] = w1;
        n_new += __popc(nm);
      }
#ifndef HB_DIAG_CLOCK
      c.V += n_new;
#endif
      __syncwarp();
      HB_TICK(t_chk1);
      HB_ACC(c.E, t_chk0, t_chk1);
#if HB_GATHER_CPASYNC && HB_OVERLAP_REPLAY
      if (n_new) hamming32_issue(gv.sk, ws.ids, min(32u, n_new), ws.stage);
#endif
      for (uint32_t g0 = 0; g0 < n_new; g0 += 32) {
        const uint32_t n = min(32u, n_new - g0);
        HB_TICK(t_g0);
        const uint32_t myid = ws.ids[g0 + (lane < n ? lane : 0u)];
#if HB_GATHER_CPASYNC && HB_OVERLAP_REPLAY
        const uint32_t myd = hamming32_finish(n, qv, ws.stage);  // nsw.rs:164-165
        if (g0 + 32 < n_new) hamming32_issue(gv.sk, ws.ids + g0 + 32, min(32u, n_new - g0 - 32), ws.stage);  // in flight during the replay
#else
        const uint32_t myd = hamming32(gv.sk, ws.ids + g0, n, qv, ws.stage);  // nsw.rs:164-165
#endif
        if (!sk_done) {
          prefetch_row_sketches(gv.sk, pw0, pw1);
          sk_done = true;
        }
#ifdef HB_DIAG_CLOCK
        const long long t_g1 = (myd != 0xFFFFFFFFu) ? clock64() : 0;  // depends on the distances: read after they arrived
        HB_ACC(c.V, t_g0, t_g1);
#endif
        // ---- sequential accept in neighbour order (nsw.rs:162-176) ----
        bool active = lane < n;
        for (;;) {
          const bool pass = active && (b.size < ef || myd < b.F());  // !(d >= F && |w| >= ef)
          const uint32_t pm = __ballot_sync(kFullMask, pass);
          if (!pm) break;
          const int first = __ffs(pm) - 1;
          active = active && ((int)lane > first);
          const uint64_t v = bm_pack(__shfl_sync(kFullMask, myd, first), __shfl_sync(kFullMask, myid, first));
          if (b.contains(v)) continue;  // a revisit the cache no longer remembered
#ifndef HB_DIAG_CLOCK
          ++c.A;
#endif
          const uint64_t ev = b.insert(v, ef);
          if (ev != kBeamInf && !(ev & 1ull) && bm_dist(ev) == b.F()) {
            // evicted before it was expanded and still tied with the worst distance: it stays a live candidate
            if (b.spill_n && b.spill_d != b.F()) b.spill_n = 0;
            b.spill_d = b.F();
            if (b.spill_n < spill_cap) {
              if (lane == 0) spill[b.spill_n] = bm_key(ev);
              ++b.spill_n;
            } else {
              status = 2;
            }
          }
        }
#ifdef HB_DIAG_CLOCK
        { HB_TICK(t_a1); HB_ACC(c.A, t_g1, t_a1); }
#endif
      }
      __syncwarp();  // ws.ids is rewritten by the next segment
      w0 = nw0;
      w1 = nw1;
    }
    if (!sk_done) prefetch_row_sketches(gv.sk, pw0, pw1);
    if (status) break;
  }
}

}  // namespace hb
