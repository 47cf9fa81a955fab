// beam.cuh -- warp-level building blocks of the batched beam search (sm_100a).
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
};

struct SearchCounters {
  uint32_t X = 0, E = 0, V = 0, A = 0;
};

// per-warp scratch in shared memory
struct WarpScratch {
  uint32_t* cache;       // 1 << cache_log2 recently seen ids (0xFFFFFFFF = empty)
  uint32_t cache_shift;  // 32 - cache_log2
  uint32_t* ids;         // 32 staged neighbour ids
};

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

// Distances of the n (1..32) sketches whose gids are staged in ids[0..n) to the query piece qv:
// lane i returns the distance of ids[i] (garbage for i >= n).  Lane = (g = lane / 8, s = lane % 8) owns the
// s-th 16-B piece of rows 4 * j + g, j = 0..7.
__device__ __forceinline__ uint32_t hamming32(const uint4* __restrict__ sk, const uint32_t* ids, uint32_t n, uint4 qv) {
  const uint32_t lane = threadIdx.x & 31, g = lane >> 3, s = lane & 7u;
  uint4 v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j)
    if (4u * j < n) v[j] = __ldg(sk + (size_t)ids[min(4u * j + g, n - 1u)] * 8 + s);
  uint32_t x[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) x[j] = (4u * j < n) ? popc_xor4(v[j], qv) : 0u;
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

// nsw::search(graph, q, ef, ep) (nsw.rs:133-184) on the graph that contains gid `ep`.  On return b holds w.
// status != 0: the boundary-tie stack overflowed (the caller reruns the query on the large-capacity path).
template <int R>
__device__ __forceinline__ void beam_search(const GraphView& gv, uint4 qv, uint32_t ef, uint32_t ep, const WarpScratch& ws,
                                            uint32_t* spill, uint32_t spill_cap, BeamRegs<R>& b, SearchCounters& c,
                                            uint32_t& status) {
  const uint32_t lane = threadIdx.x & 31;
  {
    const uint32_t d0 = hamming1(gv.sk, ep, qv);  // nsw.rs:140-141
    if (lane == 0) ws.cache[(ep * 0x9E3779B1u) >> ws.cache_shift] = ep;  // nsw.rs:143-144
    b.reset(bm_pack(d0, ep));                                           // nsw.rs:145-146
    ++c.V;
    __syncwarp();
  }
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
    ++c.X;
    // ---- for e in graph.neighborhood(c.key)  (nsw.rs:156) ----
    const uint32_t* row = gv.adj + (size_t)ckey * gv.stride;
    uint32_t word = __ldg(row + lane);
    const uint32_t deg = __shfl_sync(kFullMask, word, 0);
    c.E += deg;
    for (uint32_t base = 0; base <= deg; base += 32) {
      const uint32_t wi = base + lane;
      if (base) word = (wi <= deg) ? __ldg(row + wi) : 0u;
      const bool valid = (wi >= 1u) && (wi <= deg) && word != 0xFFFFFFFFu;  // 0xFFFFFFFF: builder tombstone
      bool isnew = false;
      if (valid) {  // nsw.rs:157-161 through the recently-seen cache
        const uint32_t h = (word * 0x9E3779B1u) >> ws.cache_shift;
        isnew = ws.cache[h] != word;
        if (isnew) ws.cache[h] = word;
      }
      const uint32_t nm = __ballot_sync(kFullMask, isnew);
      if (!nm) continue;
      const uint32_t n_new = __popc(nm);
      c.V += n_new;
      if (isnew) ws.ids[__popc(nm & ((1u << lane) - 1u))] = word;  // order-preserving compaction
      __syncwarp();
      const uint32_t myid = ws.ids[lane < n_new ? lane : 0];
      const uint32_t myd = hamming32(gv.sk, ws.ids, n_new, qv);  // nsw.rs:164-165
      __syncwarp();
      // ---- sequential accept in neighbour order (nsw.rs:162-176) ----
      bool active = lane < n_new;
      for (;;) {
        const bool pass = active && (b.size < ef || myd < b.F());  // !(d >= F && |w| >= ef)
        const uint32_t pm = __ballot_sync(kFullMask, pass);
        if (!pm) break;
        const int first = __ffs(pm) - 1;
        active = active && ((int)lane > first);
        const uint64_t v = bm_pack(__shfl_sync(kFullMask, myd, first), __shfl_sync(kFullMask, myid, first));
        if (b.contains(v)) continue;  // a revisit the cache no longer remembered
        ++c.A;
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
    }
    if (status) break;
  }
}

}  // namespace hb
