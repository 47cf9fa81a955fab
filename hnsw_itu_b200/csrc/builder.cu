// builder.cu -- HNSW index construction on the device.
//
// Follows HNSWBuilder::extend_parallel (src/index/hnsw.rs:44-169) and HNSWBuilder::add (hnsw.rs:183-244):
//   * point 0 through `add`; points 1 .. 49 999 one at a time through `add` (hnsw.rs:47-54), each with its own
//     random level (hnsw.rs:37-40); then chunks of 32 * rayon_threads points (hnsw.rs:56-59) with ONE level per
//     chunk (hnsw.rs:65);
//   * per chunk: nodes appended to base and layers[..level] without edges (hnsw.rs:74-87); ef = 1 descent from the
//     top layer to layer `level` through the stored down-links (hnsw.rs:94-108); for every layer below and the
//     base: search_select_neighbors = nsw::search(ef = efc) + select_neighbors heuristic (nsw.rs:12-58) against the
//     FROZEN graph, then insert_neighbors (nsw.rs:94-131);
//   * the reference's index-space quirks are kept: the descent result is reused unconverted as the entry index of
//     every lower layer (hnsw.rs:111-156), and `add` enters the next lower layer at the nearest selected
//     neighbour's index of the current one (hnsw.rs:222-232, nsw.rs:89-91).
//
// What runs where.  The searches and the heuristic are one warp per new point (beam.cuh); the reference applies a
// chunk's edges serially (hnsw.rs:131-139,159-167), here they are applied by atomic appends into the fixed-stride
// adjacency rows, followed by one CTA per over-full row for the M-pruning heuristic (nsw.rs:105-130) and a
// tombstone pass for the edges that pruning removed on the other side (collections/mod.rs:34-39).  Within a chunk the
// outcome of the reference depends on the serial order and on HashSet iteration order; here it depends on the
// order of atomics -- same graph family, not the same graph (the reference itself is not reproducible run to run:
// rayon's thread count is in the level seed, hnsw.rs:24-30).  The level stream is splitmix64 with the reference's
// seed formula and f32 construction, not rand's ChaCha12.
//
// The result is the same device layout index.cu produces from an .idx file, so the search kernel runs on it
// directly and hb_index_save_idx writes it in the reference's format.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <memory>

#include "beam.cuh"
#include "hb_internal.h"

namespace hb {
namespace {

constexpr uint32_t kTomb = 0xFFFFFFFFu;
constexpr int kMaxGraphs = 16;
constexpr int kBuildWarps = 4;
constexpr uint32_t kBuildCacheLog2 = 10;

struct BuildCtx {
  uint4* sk;
  uint32_t* adj;
  const uint32_t* down;  // gid of the copy one graph below, indexed by gid - n_base
  uint32_t stride;
  uint32_t n_base;
  uint32_t efc, m, M;
  uint32_t gbase[kMaxGraphs + 1];
  // per-step scratch
  uint32_t* eps;      // [chunk] entry index (local to the graph it is used in)
  uint32_t* sel;      // [chunk * m] selected neighbour gids
  uint32_t* sel_cnt;  // [chunk]
  uint32_t* spill;    // [chunk * spill_cap]
  uint32_t spill_cap;
  uint32_t* work;     // rows that went over M in this step
  uint32_t work_cap;
  uint32_t* req;      // (x, e) pairs: remove e from row x
  uint32_t req_cap;
  uint32_t* ctr;      // [0] |work| [1] |req| [2] work cursor; [8] dropped appends [9] error bits [10] prunes [11] removed edges
  uint32_t* stamp;    // [n_total] last step in which the row was pruned
};

enum : uint32_t { kErrSpill = 1u, kErrWork = 2u, kErrReq = 4u };

// ---- level stream (hnsw.rs:24-40) --------------------------------------------------------------
struct LevelRng {
  uint64_t s;
  explicit LevelRng(uint64_t seed) : s(seed) {}
  uint64_t next() {
    s += 0x9E3779B97F4A7C15ull;
    uint64_t z = s;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
  }
  uint32_t level(uint32_t connections) {
    float u;
    do { u = (float)(next() >> 40) * (1.0f / 16777216.0f); } while (u == 0.0f);  // ln(0) would give an unbounded level
    const uint32_t l = (uint32_t)(-std::log(u) * (1.0f / std::log((float)connections)));
    return std::min<uint32_t>(l, kMaxGraphs - 1);
  }
};

struct Step {
  uint64_t first;  // first base point
  uint32_t count;
  uint32_t level;
  bool sequential;  // through HNSWBuilder::add
};

// ---- kernels -------------------------------------------------------------------------------------

__device__ __forceinline__ WarpScratch warp_scratch(unsigned char* mine) { return make_warp_scratch(mine, kBuildCacheLog2); }

// ef = 1 descent through graphs g_top .. g_stop + 1 (hnsw.rs:94-108); eps[j] = entry index local to graph g_stop.
__global__ void __launch_bounds__(kBuildWarps * 32) build_descent_kernel(const BuildCtx c, uint32_t point_first, uint32_t count,
                                                                        uint32_t ep_gid, int g_top, int g_stop) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const uint32_t j = blockIdx.x * kBuildWarps + wib;
  if (j >= count) return;
  const size_t per_warp = warp_scratch_bytes(kBuildCacheLog2);
  const WarpScratch ws = warp_scratch(smem_raw + wib * per_warp);
  const GraphView gv{c.sk, c.adj, c.stride};
  const uint4 qv = __ldg(c.sk + (size_t)(point_first + j) * 8 + (lane & 7u));
  cache_clear(ws);
  uint32_t cur = ep_gid;
  SearchCounters sc;
  uint32_t status = 0;
  BeamRegs<1> b;
  for (int g = g_top; g > g_stop; --g) {
    beam_search(gv, qv, 1u, cur, ws, c.spill + (size_t)j * c.spill_cap, c.spill_cap, b, sc, status);
    cur = __ldg(c.down + (bm_key(b.at(0)) - c.n_base));
  }
  if (lane == 0) {
    c.eps[j] = cur - c.gbase[g_stop];
    if (status) atomicOr(c.ctr + 9, kErrSpill);
  }
}

// search_select_neighbors (nsw.rs:44-58) for the chunk's copies in graph g: node gids node_first .. node_first + count.
template <int R>
__global__ void __launch_bounds__(kBuildWarps * 32) build_search_select_kernel(const BuildCtx c, int g, uint32_t node_first,
                                                                              uint32_t count, int sequential) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5, grp = lane >> 3, s = lane & 7u;
  const uint32_t j = blockIdx.x * kBuildWarps + wib;
  if (j >= count) return;
  const size_t per_warp = warp_scratch_bytes(kBuildCacheLog2) + (size_t)c.m * 128;
  unsigned char* mine = smem_raw + wib * per_warp;
  const WarpScratch ws = warp_scratch(mine);
  uint4* sel_sk = reinterpret_cast<uint4*>(mine + warp_scratch_bytes(kBuildCacheLog2));  // [m][8]
  const GraphView gv{c.sk, c.adj, c.stride};
  const uint32_t self = node_first + j;
  const uint4 qv = __ldg(c.sk + (size_t)self * 8 + s);
  cache_clear(ws);
  SearchCounters sc;
  uint32_t status = 0;
  BeamRegs<R> b;
  beam_search(gv, qv, c.efc, c.gbase[g] + c.eps[j], ws, c.spill + (size_t)j * c.spill_cap, c.spill_cap, b, sc, status);

  // select_neighbors (nsw.rs:12-33): ascending (distance, key); keep e iff dist(e, r) > dist(e, q) for every kept r
  uint32_t nsel = 0;
  uint32_t* out = c.sel + (size_t)j * c.m;
  for (uint32_t i = 0; i < b.size && nsel < c.m; ++i) {
    const uint64_t e = b.at(i);
    const uint32_t ekey = bm_key(e), ed = bm_dist(e);
    const uint4 ev = __ldg(c.sk + (size_t)ekey * 8 + s);
    bool ok = true;
    for (uint32_t r0 = 0; r0 < nsel; r0 += 4) {
      const uint32_t r = r0 + grp;
      uint32_t d = (r < nsel) ? popc_xor4(ev, sel_sk[r * 8 + s]) : 0u;
      d += __shfl_xor_sync(kFullMask, d, 1);
      d += __shfl_xor_sync(kFullMask, d, 2);
      d += __shfl_xor_sync(kFullMask, d, 4);
      if (__any_sync(kFullMask, r < nsel && !(d > ed))) { ok = false; break; }
    }
    if (!ok) continue;
    if (grp == 0) sel_sk[nsel * 8 + s] = ev;
    if (lane == 0) out[nsel] = ekey;
    ++nsel;
    __syncwarp();
  }
  if (lane == 0) {
    c.sel_cnt[j] = nsel;
    if (sequential && nsel) c.eps[j] = out[0] - c.gbase[g];  // hnsw.rs:222-232: next layer enters at nb.front()'s index
    if (status) atomicOr(c.ctr + 9, kErrSpill);
  }
}

// insert_neighbors, first half (nsw.rs:101-103): add_edge(point, e) for every selected e.  One warp per new node;
// lane i owns selected[i].  Rows that end up longer than M are queued for the pruning pass (nsw.rs:105-111).
__global__ void __launch_bounds__(128) build_apply_kernel(const BuildCtx c, uint32_t node_first, uint32_t count) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t j = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (j >= count) return;
  const uint32_t p = node_first + j;
  const uint32_t n = c.sel_cnt[j];
  uint32_t* row_p = c.adj + (size_t)p * c.stride;
  uint32_t base = 0;
  if (lane == 0) base = atomicAdd(row_p, n);
  base = __shfl_sync(kFullMask, base, 0);
  for (uint32_t i = lane; i < n; i += 32) {
    const uint32_t e = c.sel[(size_t)j * c.m + i];
    if (base + i + 1 < c.stride) row_p[1 + base + i] = e;  // n <= m and the row was empty (or holds a few back edges): fits
    if (e == p) continue;                                   // self loop: one set entry (simplegraph.rs:93-100)
    uint32_t* row_e = c.adj + (size_t)e * c.stride;
    const uint32_t slot = atomicAdd(row_e, 1u);
    if (slot + 1 < c.stride) {
      row_e[1 + slot] = p;
      if (slot >= c.M) {  // degree is now > M
        const uint32_t w = atomicAdd(c.ctr + 0, 1u);
        if (w < c.work_cap) c.work[w] = e; else atomicOr(c.ctr + 9, kErrWork);
      }
    } else {
      // row e is physically full (a new top layer's star, hnsw.rs:89-92 + nsw.rs:101-103): the reference would have
      // pruned e before it got here; treat p as pruned by e and unlink the edge on p's side too
      atomicSub(row_e, 1u);
      atomicAdd(c.ctr + 8, 1u);
      const uint32_t q = atomicAdd(c.ctr + 1, 1u);
      if (q < c.req_cap) { c.req[2 * (size_t)q] = p; c.req[2 * (size_t)q + 1] = e; } else atomicOr(c.ctr + 9, kErrReq);
    }
  }
}

// insert_neighbors, second half (nsw.rs:105-130), one CTA per over-full row e: candidates = current neighbours,
// select_neighbors with m = M w.r.t. e, rewrite the row as the selection plus this step's new nodes (the reference
// re-adds the inserting point unconditionally, nsw.rs:129), and emit (x, e) for every dropped x so that e also
// leaves row x (clear_edges removes both directions, collections/mod.rs:34-39).
constexpr int kPruneThreads = 256;
constexpr int kPruneCap = 512;  // >= stride - 1

__global__ void __launch_bounds__(kPruneThreads) build_prune_kernel(const BuildCtx c, uint32_t new_lo, uint32_t new_hi, uint32_t step_id) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint4* csk = reinterpret_cast<uint4*>(smem_raw);                                   // [cap][8] candidate sketches
  const uint32_t cap = c.stride - 1;
  uint64_t* keys = reinterpret_cast<uint64_t*>(smem_raw + (size_t)cap * 128);        // [kPruneCap] d << 41 | gid << 9 | slot
  uint16_t* sel_slot = reinterpret_cast<uint16_t*>(keys + kPruneCap);                // [kPruneCap]
  uint8_t* chosen = reinterpret_cast<uint8_t*>(sel_slot + kPruneCap);                // [kPruneCap] by sorted rank
  __shared__ uint32_t s_e, s_skip;
  const uint32_t tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, grp = lane >> 3, s = lane & 7u;
  const uint32_t n_work = min(c.ctr[0], c.work_cap);

  for (;;) {
    __syncthreads();
    if (tid == 0) {
      const uint32_t w = atomicAdd(c.ctr + 2, 1u);
      s_skip = 0;
      if (w >= n_work) {
        s_e = kTomb;
      } else {
        s_e = c.work[w];
        if (atomicExch(c.stamp + s_e, step_id) == step_id) s_skip = 1;  // the row is queued once per crossing append
      }
    }
    __syncthreads();
    const uint32_t e = s_e;
    if (e == kTomb) break;
    if (s_skip) continue;
    uint32_t* row = c.adj + (size_t)e * c.stride;
    const uint32_t deg = min(row[0], cap);
    const uint4 ev = __ldg(c.sk + (size_t)e * 8 + s);
    for (uint32_t i = tid; i < kPruneCap; i += kPruneThreads) { keys[i] = ~0ull; chosen[i] = 0; }
    __syncthreads();
    // distances e <-> candidate (nsw.rs:113-119); 8 lanes per sketch row, 4 rows per warp step
    for (uint32_t i0 = warp * 4; i0 < deg; i0 += (kPruneThreads / 32) * 4) {
      const uint32_t i = i0 + grp;
      const uint32_t id = (i < deg) ? row[1 + i] : kTomb;
      uint32_t d = 0;
      if (id != kTomb) {
        const uint4 v = __ldg(c.sk + (size_t)id * 8 + s);
        csk[i * 8 + s] = v;
        d = popc_xor4(v, ev);
      }
      d += __shfl_xor_sync(kFullMask, d, 1);
      d += __shfl_xor_sync(kFullMask, d, 2);
      d += __shfl_xor_sync(kFullMask, d, 4);
      if (s == 0 && id != kTomb) keys[i] = ((uint64_t)d << 41) | ((uint64_t)id << 9) | i;
    }
    __syncthreads();
    // bitonic sort of kPruneCap keys, ascending (distance, key)
    for (uint32_t k2 = 2; k2 <= kPruneCap; k2 <<= 1)
      for (uint32_t j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
        for (uint32_t i = tid; i < kPruneCap; i += kPruneThreads) {
          const uint32_t l = i ^ j2;
          if (l > i) {
            const uint64_t a = keys[i], b = keys[l];
            const bool up = (i & k2) == 0;
            if ((a > b) == up) { keys[i] = b; keys[l] = a; }
          }
        }
        __syncthreads();
      }
    // valid keys sort first: n_valid = index of the first sentinel
    __shared__ uint32_t s_nv;
    if (tid == 0) s_nv = kPruneCap;
    __syncthreads();
    for (uint32_t i = tid; i < kPruneCap; i += kPruneThreads)
      if (keys[i] == ~0ull && (i == 0 || keys[i - 1] != ~0ull)) s_nv = i;
    __syncthreads();
    const uint32_t n_valid = s_nv;
    uint32_t ns = 0;
    if (n_valid <= c.M) {
      // only tombstones made the row look over-full: keep everything
      for (uint32_t i = tid; i < n_valid; i += kPruneThreads) chosen[i] = 1;
      ns = n_valid;
      __syncthreads();
    } else {
      // select_neighbors(candidates, M) (nsw.rs:12-33, 121): thread t compares the candidate with selected[t]
      for (uint32_t r = 0; r < n_valid && ns < c.M; ++r) {
        const uint64_t kv = keys[r];
        const uint32_t ci = (uint32_t)(kv & 511u), cd = (uint32_t)(kv >> 41);
        bool bad = false;
        for (uint32_t t = tid; t < ns; t += kPruneThreads) {
          const uint4* a = csk + ci * 8;
          const uint4* b = csk + (uint32_t)sel_slot[t] * 8;
          uint32_t d = 0;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
            const int piece = (q + (int)tid) & 7;  // rotate so that a warp's reads spread over all banks
            d += popc_xor4(a[piece], b[piece]);
          }
          bad |= !(d > cd);
        }
        if (!__syncthreads_or(bad)) {
          if (tid == 0) { sel_slot[ns] = (uint16_t)ci; chosen[r] = 1; }
          ++ns;
          __syncthreads();
        }
      }
      if (tid == 0) atomicAdd(c.ctr + 10, 1u);
    }
    // The reference re-adds the inserting point after every prune (nsw.rs:129), so this step's new nodes stay in the row
    // even when the heuristic did not pick them -- but it prunes at EVERY overflow, one insert at a time, so its row
    // never ends above M.  When a whole chunk lands on one row in a single step (the star of a new top layer,
    // hnsw.rs:89-92), selection + new nodes would stay above M here: keep only the latest new nodes (largest ids = last
    // in the reference's chunk order) up to M in total, and unlink the others like any dropped neighbour.
    __shared__ uint32_t s_len, s_cnt;
    uint32_t new_keep_from = new_lo;  // unselected new nodes with id >= this stay
    {
      const uint32_t budget = c.M > ns ? c.M - ns : 0u;
      uint32_t lo_t = new_lo, hi_t = new_hi;  // smallest T in [new_lo, new_hi] with count(unselected new id >= T) <= budget
      while (lo_t < hi_t) {
        const uint32_t mid = lo_t + ((hi_t - lo_t) >> 1);
        if (tid == 0) s_cnt = 0;
        __syncthreads();
        uint32_t mine = 0;
        for (uint32_t r = tid; r < n_valid; r += kPruneThreads) {
          const uint32_t id = (uint32_t)(keys[r] >> 9);
          mine += (!chosen[r] && id >= mid && id < new_hi) ? 1u : 0u;
        }
        if (mine) atomicAdd(&s_cnt, mine);
        __syncthreads();
        const uint32_t cnt = s_cnt;
        __syncthreads();
        if (cnt <= budget) hi_t = mid; else lo_t = mid + 1;
      }
      new_keep_from = lo_t;
    }
    if (tid == 0) s_len = 0;
    __syncthreads();
    if (warp == 0) {  // ordered compaction of the chosen ranks
      uint32_t len = 0;
      for (uint32_t r0 = 0; r0 < n_valid; r0 += 32) {
        const uint32_t r = r0 + lane;
        const bool take = r < n_valid && chosen[r];
        const uint32_t mk = __ballot_sync(kFullMask, take);
        if (take) row[1 + len + __popc(mk & ((1u << lane) - 1u))] = (uint32_t)(keys[r] >> 9);
        len += __popc(mk);
      }
      for (uint32_t r0 = 0; r0 < n_valid; r0 += 32) {
        const uint32_t r = r0 + lane;
        const uint32_t id = r < n_valid ? (uint32_t)(keys[r] >> 9) : 0u;
        const bool take = r < n_valid && !chosen[r] && id >= new_keep_from && id < new_hi;
        const uint32_t mk = __ballot_sync(kFullMask, take);
        if (take) row[1 + len + __popc(mk & ((1u << lane) - 1u))] = id;
        len += __popc(mk);
      }
      if (lane == 0) { row[0] = len; s_len = len; }
    } else {
      // dropped on this side: e must leave row x as well
      for (uint32_t r = tid - 32; r < n_valid; r += kPruneThreads - 32) {
        const uint32_t id = (uint32_t)(keys[r] >> 9);
        if (!chosen[r] && !(id >= new_keep_from && id < new_hi)) {
          const uint32_t q = atomicAdd(c.ctr + 1, 1u);
          if (q < c.req_cap) { c.req[2 * (size_t)q] = id; c.req[2 * (size_t)q + 1] = e; } else atomicOr(c.ctr + 9, kErrReq);
        }
      }
    }
  }
}

// remove_edge(x, e) on x's side: one warp per request; the slot becomes a tombstone (skipped by every reader,
// squeezed out when the row is next pruned or by the final compaction).
__global__ void __launch_bounds__(128) build_unlink_kernel(const BuildCtx c) {
  const uint32_t lane = threadIdx.x & 31;
  const uint32_t n_req = min(c.ctr[1], c.req_cap);
  for (uint32_t q = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; q < n_req; q += (gridDim.x * blockDim.x) >> 5) {
    const uint32_t x = c.req[2 * (size_t)q], e = c.req[2 * (size_t)q + 1];
    uint32_t* row = c.adj + (size_t)x * c.stride;
    const uint32_t deg = min(row[0], c.stride - 1);
    for (uint32_t i = lane; i < deg; i += 32)
      if (row[1 + i] == e) { row[1 + i] = kTomb; atomicAdd(c.ctr + 11, 1u); }
  }
}

// final pass: squeeze the tombstones out of every row (one warp per row), count directed edges, find the max degree
__global__ void __launch_bounds__(256) build_compact_kernel(uint32_t* adj, uint32_t stride, uint64_t n_rows, unsigned long long* edges,
                                                           uint32_t* max_deg) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n_rows) return;
  uint32_t* row = adj + r * stride;
  const uint32_t deg = min(row[0], stride - 1);
  uint32_t len = 0;
  for (uint32_t i0 = 0; i0 < deg; i0 += 32) {
    const uint32_t i = i0 + lane;
    const uint32_t id = i < deg ? row[1 + i] : kTomb;
    const uint32_t mk = __ballot_sync(kFullMask, id != kTomb);
    __syncwarp();
    if (id != kTomb) row[1 + len + __popc(mk & ((1u << lane) - 1u))] = id;
    len += __popc(mk);
    __syncwarp();
  }
  if (lane == 0) {
    row[0] = len;
    atomicAdd(edges, (unsigned long long)len);
    atomicMax(max_deg, len);
  }
}

__global__ void gather_degrees_kernel(const uint32_t* adj, uint32_t stride, uint64_t first, uint64_t n, uint32_t* deg) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) deg[i] = adj[(first + i) * stride];
}

// rows -> CSR ids (local to the graph): one warp per row
__global__ void rows_to_csr_kernel(const uint32_t* adj, uint32_t stride, uint64_t first, uint64_t n, uint32_t gid_base,
                                   const uint64_t* off, uint32_t* ids) {
  const uint32_t lane = threadIdx.x & 31;
  const uint64_t r = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (r >= n) return;
  const uint32_t* row = adj + (first + r) * stride;
  const uint32_t deg = row[0];
  for (uint32_t i = lane; i < deg; i += 32) ids[off[r] + i] = row[1 + i] - gid_base;
}

template <class T>
struct DevArr {
  T* p = nullptr;
  size_t n = 0;
  DevArr() = default;
  explicit DevArr(size_t count) { alloc(count); }
  void alloc(size_t count) {
    n = count;
    HB_CUDA(cudaMalloc(&p, std::max<size_t>(count, 1) * sizeof(T)));
  }
  ~DevArr() { cudaFree(p); }
  DevArr(const DevArr&) = delete;
  DevArr& operator=(const DevArr&) = delete;
};

using SelectKernel = void (*)(const BuildCtx, int, uint32_t, uint32_t, int);
SelectKernel select_kernel_for(uint32_t efc) {
  if (efc <= 32) return build_search_select_kernel<1>;
  if (efc <= 64) return build_search_select_kernel<2>;
  if (efc <= 128) return build_search_select_kernel<4>;
  return build_search_select_kernel<8>;
}

// device rows -> host CSR (graph 0 = base), ids local to each graph, canonical order = row order
void export_graphs(const Index& idx, const uint64_t* base_data, const std::vector<uint32_t>& up_point, std::vector<HostGraph>& out) {
  const size_t G = idx.g_nodes.size();
  out.assign(G, HostGraph{});
  for (size_t g = 0; g < G; ++g) {
    HostGraph& hg = out[g];
    const uint64_t n = idx.g_nodes[g], first = idx.g_base[g];
    hg.n = n;
    hg.sk.resize(n * 16);
    if (g == 0) {
      if (n) std::memcpy(hg.sk.data(), base_data, n * 128);
    } else {
      hg.down.resize(n);
      std::vector<uint32_t> dn(n);
      if (n) HB_CUDA(cudaMemcpy(dn.data(), idx.d_down + (first - idx.n_base), n * 4, cudaMemcpyDeviceToHost));
      for (uint64_t i = 0; i < n; ++i) {
        hg.down[i] = dn[i] - idx.g_base[g - 1];
        std::memcpy(hg.sk.data() + i * 16, base_data + (uint64_t)up_point[first - idx.n_base + i] * 16, 128);
      }
    }
    hg.off.assign(n + 1, 0);
    if (!n) continue;
    DevArr<uint32_t> d_deg(n);
    gather_degrees_kernel<<<(unsigned)((n + 255) / 256), 256>>>(idx.d_adj, idx.stride, first, n, d_deg.p);
    count_launches();
    HB_CUDA(cudaGetLastError());
    std::vector<uint32_t> deg(n);
    HB_CUDA(cudaMemcpy(deg.data(), d_deg.p, n * 4, cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < n; ++i) hg.off[i + 1] = hg.off[i] + deg[i];
    hg.ids.resize(hg.off[n]);
    if (hg.ids.empty()) continue;
    DevArr<uint64_t> d_off(n + 1);
    DevArr<uint32_t> d_ids(hg.ids.size());
    HB_CUDA(cudaMemcpy(d_off.p, hg.off.data(), (n + 1) * 8, cudaMemcpyHostToDevice));
    rows_to_csr_kernel<<<(unsigned)((n * 32 + 255) / 256), 256>>>(idx.d_adj, idx.stride, first, n, (uint32_t)first, d_off.p, d_ids.p);
    count_launches();
    HB_CUDA(cudaGetLastError());
    HB_CUDA(cudaMemcpy(hg.ids.data(), d_ids.p, hg.ids.size() * 4, cudaMemcpyDeviceToHost));
  }
}

}  // namespace

Index* build_hnsw(const uint64_t* data, uint64_t n, const hb_build_params& bp, int device, bool keep_host) {
  const uint32_t efc = bp.ef_construction, m = bp.connections, M = bp.max_connections;
  HB_REQUIRE(efc >= 1 && efc <= 256, HB_ERR_UNSUPPORTED, "ef_construction must be in 1..256");
  HB_REQUIRE(m >= 2 && m <= 64, HB_ERR_UNSUPPORTED, "connections must be in 2..64 (the level formula divides by ln m)");
  HB_REQUIRE(M >= m && M <= 416, HB_ERR_UNSUPPORTED, "max_connections must be in connections..416");
  HB_REQUIRE(bp.rayon_threads >= 1 && bp.rayon_threads <= 4096, HB_ERR_INVALID, "rayon_threads must be in 1..4096");
  HB_REQUIRE(n <= 0xF0000000ull, HB_ERR_UNSUPPORTED, "too many points");
  ensure_device(device);
  const auto t0 = std::chrono::steady_clock::now();
  if (n == 0) {
    std::vector<HostGraph> g(1);
    g[0].off.assign(1, 0);
    return index_from_host(bp.nsw ? HB_ALGO_NSW : HB_ALGO_HNSW, std::move(g), -1, device, keep_host);
  }

  // ---- plan: steps, levels, layer sizes (all host-side and deterministic) -----------------------------------
  const uint64_t prefix = std::min<uint64_t>(n, bp.serial_prefix ? bp.serial_prefix : 50000);  // hnsw.rs:54
  const uint32_t chunk = 32u * bp.rayon_threads;                                             // hnsw.rs:56
  LevelRng rng((uint64_t)bp.rayon_threads ^ n ^ efc ^ m ^ M);                               // hnsw.rs:24-30
  std::vector<Step> steps;
  steps.reserve(prefix + (n - prefix) / chunk + 2);
  steps.push_back(Step{0, 1, 0, true});  // first point: level = layers.len() = 0 (hnsw.rs:185-190)
  // NSWBuilder (nsw.rs:224-322) is the same insertion with every point on the base graph only and ep = the first point
  const bool nsw = bp.nsw != 0;
  for (uint64_t i = 1; i < prefix; ++i) steps.push_back(Step{i, 1, nsw ? 0u : rng.level(m), true});
  for (uint64_t i = prefix; i < n; i += chunk)
    steps.push_back(Step{i, (uint32_t)std::min<uint64_t>(chunk, n - i), nsw ? 0u : rng.level(m), false});
  uint32_t n_layers = 0;
  for (const Step& s : steps) n_layers = std::max(n_layers, s.level);
  const size_t G = (size_t)n_layers + 1;
  std::vector<uint64_t> gsize(G, 0), gbase(G + 1, 0);
  for (const Step& s : steps)
    for (uint32_t g = 0; g <= s.level; ++g) gsize[g] += s.count;
  for (size_t g = 0; g < G; ++g) gbase[g + 1] = gbase[g] + gsize[g];
  const uint64_t n_total = gbase[G];
  HB_REQUIRE(n_total <= 0xFFFFFFF0ull, HB_ERR_UNSUPPORTED, "more than 2^32 nodes over all graphs");
  const uint64_t n_upper = n_total - n;
  std::vector<uint32_t> up_point(n_upper), up_down(n_upper);
  {
    std::vector<uint64_t> cur(G, 0);
    for (const Step& s : steps) {
      for (uint32_t g = 1; g <= s.level; ++g)
        for (uint32_t j = 0; j < s.count; ++j) {
          const uint64_t at = gbase[g] - n + cur[g] + j;
          up_point[at] = (uint32_t)(s.first + j);
          up_down[at] = (uint32_t)(gbase[g - 1] + cur[g - 1] + j);
        }
      for (uint32_t g = 0; g <= s.level; ++g) cur[g] += s.count;
    }
  }

  // ---- device arrays ------------------------------------------------------------------------------------
  std::unique_ptr<Index> idx(new Index());
  idx->device = device;
  idx->algo = nsw ? HB_ALGO_NSW : HB_ALGO_HNSW;
  idx->n_total = n_total;
  idx->n_base = n;
  idx->stride = (M + 1 + 64 + 31) / 32 * 32;  // M + 1 after pruning (nsw.rs:129) plus room for one step's appends
  cudaDeviceProp prop;
  HB_CUDA(cudaGetDeviceProperties(&prop, device));
  idx->sm_count = prop.multiProcessorCount;
  HB_CUDA(cudaMalloc(&idx->d_sk, n_total * 128));
  HB_CUDA(cudaMalloc(&idx->d_adj, n_total * idx->stride * 4ull));
  HB_CUDA(cudaMalloc(&idx->d_down, std::max<uint64_t>(n_upper, 1) * 4));
  idx->device_bytes = n_total * 128 + n_total * idx->stride * 4ull + n_upper * 4;
  idx->adj_words = n_total * idx->stride;
  HB_CUDA(cudaMemset(idx->d_adj, 0, n_total * idx->stride * 4ull));
  HB_CUDA(cudaMemcpy(idx->d_sk, data, n * 128, cudaMemcpyHostToDevice));
  if (n_upper) {
    std::vector<uint64_t> up_sk(n_upper * 16);
    for (uint64_t i = 0; i < n_upper; ++i) std::memcpy(up_sk.data() + i * 16, data + (uint64_t)up_point[i] * 16, 128);
    HB_CUDA(cudaMemcpy((char*)idx->d_sk + n * 128, up_sk.data(), n_upper * 128, cudaMemcpyHostToDevice));
    HB_CUDA(cudaMemcpy(idx->d_down, up_down.data(), n_upper * 4, cudaMemcpyHostToDevice));
  }

  const uint32_t max_count = std::max<uint32_t>(1, std::min<uint64_t>(chunk, n));
  const uint32_t spill_cap = 1024;
  DevArr<uint32_t> d_eps(max_count), d_sel((size_t)max_count * m), d_cnt(max_count), d_spill((size_t)max_count * spill_cap);
  const uint32_t work_cap = max_count * m;
  const uint32_t req_cap = (uint32_t)std::min<uint64_t>((uint64_t)work_cap * (idx->stride - 1), 32ull << 20);
  DevArr<uint32_t> d_work(work_cap), d_req((size_t)req_cap * 2), d_ctr(16), d_stamp(n_total);
  HB_CUDA(cudaMemset(d_ctr.p, 0, 16 * 4));
  HB_CUDA(cudaMemset(d_stamp.p, 0, n_total * 4));

  BuildCtx c{};
  c.sk = idx->d_sk;
  c.adj = idx->d_adj;
  c.down = idx->d_down;
  c.stride = idx->stride;
  c.n_base = (uint32_t)n;
  c.efc = efc;
  c.m = m;
  c.M = M;
  HB_REQUIRE(G <= (size_t)kMaxGraphs, HB_ERR_INTERNAL, "too many layers");
  for (size_t g = 0; g <= G; ++g) c.gbase[g] = (uint32_t)gbase[g];
  c.eps = d_eps.p;
  c.sel = d_sel.p;
  c.sel_cnt = d_cnt.p;
  c.spill = d_spill.p;
  c.spill_cap = spill_cap;
  c.work = d_work.p;
  c.work_cap = work_cap;
  c.req = d_req.p;
  c.req_cap = req_cap;
  c.ctr = d_ctr.p;
  c.stamp = d_stamp.p;

  const SelectKernel select_kernel = select_kernel_for(efc);
  const size_t smem_descent = (size_t)kBuildWarps * warp_scratch_bytes(kBuildCacheLog2);
  const size_t smem_select = (size_t)kBuildWarps * (warp_scratch_bytes(kBuildCacheLog2) + (size_t)m * 128);
  const size_t smem_prune = (size_t)(idx->stride - 1) * 128 + kPruneCap * 8 + kPruneCap * 2 + kPruneCap;
  HB_CUDA(cudaFuncSetAttribute(select_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_select));
  HB_CUDA(cudaFuncSetAttribute(build_prune_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_prune));
  cudaStream_t stream = nullptr;

  // ---- the insertion loop: everything is enqueued without a host round trip -------------------------------
  std::vector<uint64_t> cur(G, 0);
  uint32_t cur_layers = 0, ep_gid = 0, step_id = 0;
  for (const Step& s : steps) {
    ++step_id;
    if (s.level > cur_layers || step_id == 1) {  // new top layers: ep = the first point's copy in the new top (hnsw.rs:89-92, 207-210)
      cur_layers = std::max(cur_layers, s.level);
      ep_gid = (uint32_t)(gbase[s.level] + cur[s.level]);
    }
    const uint32_t blocks = (s.count + kBuildWarps - 1) / kBuildWarps;
    build_descent_kernel<<<blocks, kBuildWarps * 32, smem_descent, stream>>>(c, (uint32_t)s.first, s.count, ep_gid, (int)cur_layers,
                                                                            (int)s.level);
    count_launches();
    for (int g = (int)s.level; g >= 0; --g) {
      const uint32_t node_first = (uint32_t)(gbase[g] + cur[g]);
      HB_CUDA(cudaMemsetAsync(d_ctr.p, 0, 3 * 4, stream));
      select_kernel<<<blocks, kBuildWarps * 32, smem_select, stream>>>(c, g, node_first, s.count, s.sequential ? 1 : 0);
    count_launches();
      build_apply_kernel<<<(s.count * 32 + 127) / 128, 128, 0, stream>>>(c, node_first, s.count);
    count_launches();
      const uint32_t prune_blocks = s.sequential ? 2u : (uint32_t)std::min<uint64_t>((uint64_t)idx->sm_count * 2, work_cap);
      build_prune_kernel<<<prune_blocks, kPruneThreads, smem_prune, stream>>>(c, node_first, node_first + s.count, step_id);
    count_launches();
      build_unlink_kernel<<<s.sequential ? 4u : (unsigned)idx->sm_count, 128, 0, stream>>>(c);
    count_launches();
    }
    for (uint32_t g = 0; g <= s.level; ++g) cur[g] += s.count;
    if ((step_id & 1023u) == 0) HB_CUDA(cudaGetLastError());
  }
  HB_CUDA(cudaGetLastError());

  // ---- finish: squeeze tombstones, statistics ----------------------------------------------------------------
  DevArr<unsigned long long> d_edges(G);
  DevArr<uint32_t> d_maxdeg(1);
  HB_CUDA(cudaMemsetAsync(d_edges.p, 0, G * 8, stream));
  HB_CUDA(cudaMemsetAsync(d_maxdeg.p, 0, 4, stream));
  for (size_t g = 0; g < G; ++g)
    if (gsize[g])
      build_compact_kernel<<<(unsigned)((gsize[g] * 32 + 255) / 256), 256, 0, stream>>>(idx->d_adj + gbase[g] * idx->stride, idx->stride,
                                                                                       gsize[g], d_edges.p + g, d_maxdeg.p);
    count_launches();
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaStreamSynchronize(stream));
  uint32_t h_ctr[16];
  HB_CUDA(cudaMemcpy(h_ctr, d_ctr.p, sizeof h_ctr, cudaMemcpyDeviceToHost));
  HB_REQUIRE(!(h_ctr[9] & kErrSpill), HB_ERR_INTERNAL, "index build: boundary-tie stack overflow");
  HB_REQUIRE(!(h_ctr[9] & kErrWork), HB_ERR_INTERNAL, "index build: prune worklist overflow");
  HB_REQUIRE(!(h_ctr[9] & kErrReq), HB_ERR_INTERNAL, "index build: unlink request list overflow");
  std::vector<unsigned long long> h_edges(G);
  HB_CUDA(cudaMemcpy(h_edges.data(), d_edges.p, G * 8, cudaMemcpyDeviceToHost));

  idx->g_nodes.assign(gsize.begin(), gsize.end());
  idx->g_edges.assign(h_edges.begin(), h_edges.end());
  idx->g_base.assign(gbase.begin(), gbase.begin() + G);
  idx->ep_gid = ep_gid;
  idx->ep_file = (int64_t)(ep_gid - gbase[cur_layers]);
  idx->build_dropped_appends = h_ctr[8];
  idx->build_prunes = h_ctr[10];
  idx->build_unlinked = h_ctr[11];
  idx->attrs.format_size = true;
  idx->attrs.data = "hamming";
  idx->attrs.size = n;
  idx->attrs.algo = nsw ? HB_ALGO_NSW : HB_ALGO_HNSW;
  idx->attrs.params = "index=(efc=" + std::to_string(efc) + ",m=" + std::to_string(m) + ",M=" + std::to_string(M) + ")";  // main.rs:171-174
  if (keep_host) {
    export_graphs(*idx, data, up_point, idx->host);
    idx->has_host = true;
  }
  idx->attrs.buildtime = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  if (std::getenv("HB_INDEX_LAYOUT") && std::string(std::getenv("HB_INDEX_LAYOUT")) == "compact") index_compact(*idx);
  return idx.release();
}

}  // namespace hb
