// search.cu -- batched HNSW / NSW beam search for sm_100a (north_star subsystem 2).
//
// Replaces, bit-exactly (ids + distances, neighbour order = .idx file order):
//   Index::knns            src/index/mod.rs:28-42   (one launch per batch; results in input order)
//   HNSW::search           src/index/hnsw.rs:278-300 (top-down ef=1 descent, base ef-beam, drain_asc().take(k))
//   NSW::search            src/index/nsw.rs:341-351  (same kernel with zero upper layers)
//   nsw::search            src/index/nsw.rs:133-184  (the hot loop)
//   Sketch::distance       cli/src/sketch.rs:18-29   (XOR + POPC over 16 u64)
//   Distance Ord           src/index/mod.rs:90-97    ((distance, key) packed into one u64 compare)
//   Indexes::search pad    cli/src/main.rs:467-474
//
// Mapping: ONE WARP PER QUERY, persistent warps pulling query numbers from an
// atomic counter.  Per expansion a warp
//   1. reads the fixed-stride adjacency row of the candidate (coalesced 128-B lines),
//   2. test-and-sets every neighbour in an EXACT open-addressing visited table
//      (one table per resident warp in global memory, sized to stay L2 resident),
//   3. compacts the unvisited ids in neighbour order (ballot + popc prefix),
//   4. gathers their 128-B sketches 8 lanes x LDG.128 per row, 4 rows per
//      instruction, up to 8 independent instructions in flight, XOR+POPC,
//      3 shuffle steps per 8-lane group,
//   5. replays the reference's sequential accept / evict rule IN NEIGHBOUR ORDER:
//      ballot(d < F or beam not full), take the first set lane, insert, refresh F,
//      re-ballot the lanes after it.
//
// The reference's unbounded `cands` heap is replaced by an exact bounded form
// (proved against the restatement in tests/test_bounded_model.py):
//   * the beam `w` is one sorted array of (distance<<33 | key<<1 | expanded) in
//     shared memory; "cands.pop_min()" = first element of w not yet expanded;
//   * an element evicted from a full beam can only ever be expanded again if its
//     distance equals the current worst distance F (break is strict, nsw.rs:152);
//     such elements are evicted in decreasing (distance, key) order, so a STACK of
//     keys sharing one distance reproduces pop_min over them; it dies when F drops.
#include <cstdlib>
#include <memory>

#include "beam.cuh"
#include "beam2.cuh"
#include "hb_internal.h"

namespace hb {
namespace {

constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr uint32_t kEmpty = 0xFFFFFFFFu;
#ifndef HB_WARPS
#define HB_WARPS 4
#endif
constexpr int kWarpsPerBlock = HB_WARPS;

#ifdef HB_DEBUG_BOUNDS
#define HB_CHK(cond, code, val)                                   \
  do {                                                            \
    if (!(cond)) {                                                \
      if (atomicCAS(p.ctl + 2, 0u, (uint32_t)(code)) == 0u) {     \
        p.ctl[3] = (uint32_t)(val);                               \
        p.ctl[4] = q;                                             \
      }                                                           \
      return;                                                     \
    }                                                             \
  } while (0)
#else
#define HB_CHK(cond, code, val) do {} while (0)
#endif

struct SearchParams {
  const uint4* sk;
  const uint32_t* adj;
  const uint32_t* rowoff;  // compact layout (nullptr: fixed stride)
  const uint32_t* down;
  const uint4* queries;
  const uint32_t* qmap;  // optional indirection for retries
  uint64_t* keys;
  uint64_t* dists;
  uint32_t* found;
  uint4* counters;  // optional
  uint32_t* vis;
  uint32_t* spill;
  uint32_t* ctl;      // [0] next query, [1] overflow flag
  uint32_t* qstatus;  // per query: 0 ok, 1 visited table full, 2 tie stack full
  uint32_t stride;
  uint32_t n_total;
  uint32_t n_base;
  uint32_t n_upper;  // number of upper graphs (0 for NSW)
  uint32_t ep_gid;
  uint32_t ef, k, flags, nq;
  uint32_t ef_cap;  // beam capacity in shared memory (>= ef)
  uint32_t vcap_log2, spill_cap;
};

__device__ __forceinline__ uint64_t pack(uint32_t d, uint32_t key) { return ((uint64_t)d << 33) | ((uint64_t)key << 1); }
__device__ __forceinline__ uint32_t pk_dist(uint64_t v) { return (uint32_t)(v >> 33); }
__device__ __forceinline__ uint32_t pk_key(uint64_t v) { return (uint32_t)(v >> 1); }

__device__ __forceinline__ uint32_t popc4(uint4 a, uint4 b) {
  return __popc(a.x ^ b.x) + __popc(a.y ^ b.y) + __popc(a.z ^ b.z) + __popc(a.w ^ b.w);
}

// exact visited set: returns true when `id` was not present (and inserts it).
// The caller guarantees load < 7/8, so the probe loop terminates.
__device__ __forceinline__ bool visit(uint32_t* tab, uint32_t mask, uint32_t shift, uint32_t id) {
  uint32_t h = (id * 0x9E3779B1u) >> shift;
  for (;;) {
    uint32_t old = atomicCAS(tab + h, kEmpty, id);
    if (old == kEmpty) return true;
    if (old == id) return false;
    h = (h + 1) & mask;
  }
}

struct Beam {
  uint32_t size;     // |w|
  uint32_t cursor;   // every element before it is expanded
  uint32_t F;        // distance of the worst element of w
  uint32_t spill_n;  // boundary-tie stack depth
  uint32_t spill_d;  // distance shared by the stack entries
};

// w.push(e); if |w| > ef: w.pop_max()   (nsw.rs:171-176) on the sorted array.
__device__ __forceinline__ void beam_insert(uint64_t* w, uint32_t ef, Beam& b, uint32_t d, uint32_t key, uint32_t* spill,
                                            uint32_t spill_cap, uint32_t& status, uint32_t lane) {
  const uint64_t v = pack(d, key);
  uint32_t pos = 0;  // number of elements below v
  for (uint32_t base = 0; base < b.size; base += 32) {
    const uint32_t i = base + lane;
    const bool lt = (i < b.size) && (w[i] < v);
    const uint32_t m = __ballot_sync(kFull, lt);
    pos += __popc(m);
    if (m != kFull) break;
  }
  const bool full = b.size >= ef;
  uint64_t ev = 0;
  if (full) ev = w[ef - 1];  // the element pop_max() removes (never the one being pushed: d < F)
  __syncwarp();              // every lane has read ev before lane 0 may overwrite w[ef - 1] (pos == ef - 1: no shift below)
  const int hi = full ? (int)ef - 2 : (int)b.size - 1;
  if (hi >= (int)pos) {
    for (int base = hi & ~31; base + 31 >= (int)pos; base -= 32) {
      const int i = base + (int)lane;
      const bool mv = i >= (int)pos && i <= hi;
      const uint64_t e = mv ? w[i] : 0ull;
      __syncwarp();
      if (mv) w[i + 1] = e;
      __syncwarp();
    }
  }
  if (lane == 0) w[pos] = v;
  __syncwarp();
  if (!full) b.size++;
  b.F = pk_dist(w[b.size - 1]);
  if (pos < b.cursor) b.cursor = pos;
  if (full && !(ev & 1ull) && pk_dist(ev) == b.F) {
    // evicted before it was expanded and still tied with the worst distance: it stays a live candidate
    if (b.spill_n && b.spill_d != b.F) b.spill_n = 0;
    b.spill_d = b.F;
    if (b.spill_n < spill_cap) {
      if (lane == 0) spill[b.spill_n] = pk_key(ev);
      b.spill_n++;
    } else {
      status = 2;
    }
  }
}

__global__ void __launch_bounds__(kWarpsPerBlock * 32, 6) hnsw_search_kernel(const SearchParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const uint32_t slot = blockIdx.x * kWarpsPerBlock + wib;
  uint64_t* const w = reinterpret_cast<uint64_t*>(smem_raw) + (size_t)wib * p.ef_cap;
  uint32_t* const ids_s = reinterpret_cast<uint32_t*>(reinterpret_cast<uint64_t*>(smem_raw) + (size_t)kWarpsPerBlock * p.ef_cap) + wib * 32;
  uint32_t* const vis = p.vis + ((size_t)slot << p.vcap_log2);
  uint32_t* const spill = p.spill + (size_t)slot * p.spill_cap;
  const uint32_t vmask = (1u << p.vcap_log2) - 1u, vshift = 32u - p.vcap_log2;
  const uint32_t vlimit = (7u << p.vcap_log2) >> 3;
  const uint32_t sub = lane & 7u;  // which 16-B piece of a 128-B sketch this lane owns

  for (;;) {
    uint32_t qi = 0;
    if (lane == 0) qi = atomicAdd(p.ctl, 1u);
    qi = __shfl_sync(kFull, qi, 0);
    if (qi >= p.nq) break;
    const uint32_t q = p.qmap ? p.qmap[qi] : qi;
    const uint4 qv = __ldg(p.queries + (size_t)q * 8 + sub);

    uint32_t X = 0, E = 0, V = 0, A = 0;
    uint32_t vcount = 0, status = 0;
    uint32_t cur = p.ep_gid;
    Beam b{0, 0, 0, 0, 0};

    for (int g = (int)p.n_upper; g >= 0 && status == 0; --g) {
      const uint32_t ef = (g == 0) ? p.ef : 1u;  // hnsw.rs:286 (ef = 1 above the base)
      // ---- nsw::search(graph g, q, ef, cur) ---------------------------------
      {
        HB_CHK(cur < p.n_total, 1, cur);
        uint32_t d0 = popc4(__ldg(p.sk + (size_t)cur * 8 + sub), qv);  // nsw.rs:140-141
        d0 += __shfl_xor_sync(kFull, d0, 1);
        d0 += __shfl_xor_sync(kFull, d0, 2);
        d0 += __shfl_xor_sync(kFull, d0, 4);
        if (lane == 0) {
          visit(vis, vmask, vshift, cur);  // nsw.rs:143-144
          w[0] = pack(d0, cur);            // nsw.rs:145-146
        }
        __syncwarp();
        ++vcount;
        ++V;
        b = Beam{1u, 0u, d0, 0u, 0u};
      }
      for (;;) {
        // ---- c = cands.pop_min(); break if c.distance > F  (nsw.rs:148-154) ----
        int ci = -1;
        for (uint32_t base = b.cursor; base < b.size; base += 32) {
          const uint32_t i = base + lane;
          const bool un = (i < b.size) && !(w[i] & 1ull);
          const uint32_t m = __ballot_sync(kFull, un);
          if (m) { ci = (int)base + __ffs(m) - 1; break; }
        }
        uint32_t ckey;
        if (ci >= 0) {
          const uint64_t c = w[ci];
          ckey = pk_key(c);
          __syncwarp();
          if (lane == 0) w[ci] = c | 1ull;
          __syncwarp();
          b.cursor = (uint32_t)ci + 1u;
        } else if (b.spill_n && b.spill_d == b.F && b.size >= ef) {
          b.cursor = b.size;
          uint32_t kk = 0;
          if (lane == 0) kk = spill[b.spill_n - 1];
          ckey = __shfl_sync(kFull, kk, 0);
          b.spill_n--;
        } else {
          break;
        }
        ++X;
        // ---- for e in graph.neighborhood(c.key)  (nsw.rs:156) ------------------
        HB_CHK(ckey < p.n_total, 2, ckey);
        const uint32_t* row = p.rowoff ? p.adj + (size_t)__ldg(p.rowoff + ckey) * 32 : p.adj + (size_t)ckey * p.stride;
        uint32_t word = __ldg(row + lane);
        const uint32_t deg = __shfl_sync(kFull, word, 0);
#ifdef HB_DEBUG_BOUNDS
        if (deg >= p.stride) { p.ctl[5] = ckey; p.ctl[6] = X; p.ctl[7] = (uint32_t)g | (ci < 0 ? 0x100u : 0u) | (b.size << 16); }
#endif
        HB_CHK(p.rowoff || deg < p.stride, 3, deg);
        E += deg;
        if (vcount + deg > vlimit) { status = 1; break; }
        for (uint32_t base = 0; base <= deg; base += 32) {
          const uint32_t wi = base + lane;
          if (base) word = (wi <= deg) ? __ldg(row + wi) : 0u;
          const bool valid = (wi >= 1u) && (wi <= deg);
          HB_CHK(__all_sync(kFull, !valid || word < p.n_total), 4, word);
          const bool isnew = valid && visit(vis, vmask, vshift, word);  // nsw.rs:157-161
          const uint32_t nm = __ballot_sync(kFull, isnew);
          if (!nm) continue;
          const uint32_t n_new = __popc(nm);
          vcount += n_new;
          V += n_new;
          if (isnew) ids_s[__popc(nm & ((1u << lane) - 1u))] = word;  // order-preserving compaction
          __syncwarp();
          const uint32_t myid = ids_s[lane < n_new ? lane : 0];
          uint32_t myd = 0;
          {
            // distance_fn(point, query) for all unvisited neighbours (nsw.rs:164-165)
            uint4 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (4u * j < n_new) {
                const uint32_t r = min(4u * j + (lane >> 3), n_new - 1u);
                v[j] = __ldg(p.sk + (size_t)ids_s[r] * 8 + sub);
              }
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              if (4u * j < n_new) {
                uint32_t x = popc4(v[j], qv);
                x += __shfl_xor_sync(kFull, x, 1);
                x += __shfl_xor_sync(kFull, x, 2);
                x += __shfl_xor_sync(kFull, x, 4);
                const uint32_t got = __shfl_sync(kFull, x, (lane & 3u) * 8u);
                if ((lane >> 2) == (uint32_t)j) myd = got;
              }
            }
          }
          __syncwarp();
          // ---- sequential accept in neighbour order (nsw.rs:162-176) ----------
          bool active = lane < n_new;
          for (;;) {
            const bool pass = active && (b.size < ef || myd < b.F);  // !(d >= F && |w| >= ef)
            const uint32_t pm = __ballot_sync(kFull, pass);
            if (!pm) break;
            const int first = __ffs(pm) - 1;
            const uint32_t di = __shfl_sync(kFull, myd, first), idi = __shfl_sync(kFull, myid, first);
            ++A;
            beam_insert(w, ef, b, di, idi, spill, p.spill_cap, status, lane);
            active = active && ((int)lane > first);
          }
        }
        if (status) break;
      }
      HB_CHK(!(g > 0 && status == 0) || (pk_key(w[0]) >= p.n_base && pk_key(w[0]) < p.n_total), 5, pk_key(w[0]));
      if (g > 0 && status == 0) cur = __ldg(p.down + (pk_key(w[0]) - p.n_base));  // hnsw.rs:288-292 (.point.1)
      __syncwarp();
    }

    // ---- drain_asc().take(k) (+ Indexes::search padding) -------------------------
    const uint32_t nres = status ? 0u : min(b.size, p.k);
    for (uint32_t j = lane; j < p.k; j += 32) {
      uint64_t key = 0, dist = 0;
      if (j < nres) { key = pk_key(w[j]); dist = pk_dist(w[j]); }
      else if ((p.flags & HB_PAD_LIKE_CLI) && nres) { key = pk_key(w[0]); dist = pk_dist(w[0]); }
      p.keys[(size_t)q * p.k + j] = key;
      p.dists[(size_t)q * p.k + j] = dist;
    }
    if (lane == 0) {
      p.found[q] = nres;
      p.qstatus[q] = status;
      if (status) atomicOr(p.ctl + 1, 1u);
      if (p.counters) p.counters[q] = make_uint4(X, E, V, A);
    }
    __syncwarp();
    // ---- leave the visited table empty for the next query of this warp ---------------
    uint4* v4 = reinterpret_cast<uint4*>(vis);
    const uint4 e4 = make_uint4(kEmpty, kEmpty, kEmpty, kEmpty);
    for (uint32_t i = lane; i < (1u << p.vcap_log2) / 4u; i += 32) v4[i] = e4;
    __syncwarp();
  }
}

struct FastLaunch {
  void (*kernel)(const SearchParams, uint32_t);
};

// ---------------------------------------------------------------------------------------------
// Fast path (ef <= 256; beam2.cuh): beam in split 32-bit registers, register gather with 4 lanes per sketch row,
// recently-seen cache with 16-bit tags.  Queries whose boundary-tie stack overflows are flagged and rerun on the
// large-capacity kernel above.
// ---------------------------------------------------------------------------------------------
// Resident CTAs per SM and entries (log2) of the recently-seen cache per warp, by beam size.  Measured on B200 (10 M
// points, 10 k queries; profiles/r02d_ab_gather_variants.log, r02e_ab_occupancy_rowahead_compact.log):
//   ef <= 64 (R <= 2, 56 registers): 9 CTAs with a 512-entry cache beat 8 CTAs with 1 024 entries (8.6-8.7 against
//   8.3-8.5 M queries/s) although the smaller cache re-evaluates more nodes (V 3 094 against 2 814; the reference's
//   exact visited set: 2 474) -- the kernel is bound by memory latency, and one more CTA of warps hides more of it;
//   ef > 64 (R >= 4, 72 registers -> 7 CTAs whatever the cache): the 1 024-entry cache (ef = 128: 4.8 against 4.2).
// Compile-time so that every shared-memory address folds.
#ifndef HB_MB2
#define HB_MB2 9
#endif
#define HB_MIN_BLOCKS2(R) ((R) <= 2 ? HB_MB2 : (HB_MB2 < 7 ? HB_MB2 : 7))
#ifndef HB_CACHE2_LOG2
#define HB_CACHE2_LOG2 9
#endif
#ifndef HB_CACHE2_LOG2_BIG
#define HB_CACHE2_LOG2_BIG 10
#endif
template <bool TAG16, int R>
__host__ __device__ constexpr uint32_t cache2_log2() {
  return (R <= 2 ? HB_CACHE2_LOG2 : HB_CACHE2_LOG2_BIG) - (TAG16 ? 0 : 1);  // full ids: half the entries in the same bytes
}

template <int R, bool TAG16, bool COUNT>
__global__ void __launch_bounds__(kWarpsPerBlock * 32, HB_MIN_BLOCKS2(R)) hnsw_search_v2_kernel(const SearchParams p, uint32_t) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr uint32_t cache_log2 = cache2_log2<TAG16, R>();
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  unsigned char* mine = smem_raw + wib * warp_scratch2_bytes<TAG16>(cache_log2);
  WarpScratch2<TAG16> ws;
  ws.stage = reinterpret_cast<uint4*>(mine);
  ws.ids = reinterpret_cast<uint32_t*>(mine + kStageBytes2);
  ws.spill = ws.ids + kIdsWords2;
  unsigned char* const cache_mem = mine + kStageBytes2 + (kIdsWords2 + kSpillCap2) * 4;
  ws.cache.init(cache_mem, cache_log2);
#ifdef HB_DEBUG_BOUNDS
  const GraphView gv{p.sk, p.adj, p.stride, p.rowoff, p.ctl + 2, p.n_total};
#else
  const GraphView gv{p.sk, p.adj, p.stride, p.rowoff};
#endif

  for (;;) {
    uint32_t qi = 0;
    if (lane == 0) qi = atomicAdd(p.ctl, 1u);
    qi = __shfl_sync(kFullMask, qi, 0);
    if (qi >= p.nq) break;
    const uint32_t q = p.qmap ? p.qmap[qi] : qi;
    const QueryRegs qr = load_query(p.queries, q);
#ifdef HB_DIAG_TIME
    unsigned long long t_start;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_start));
#endif
    fill_ff(cache_mem, SeenCache<TAG16>::bytes(cache_log2));

    SearchCounters c;
    uint32_t status = 0;
    uint32_t cur = p.ep_gid;
    BeamSplit<R> b;
    for (int g = (int)p.n_upper; g >= 0 && status == 0; --g) {
      const uint32_t ef = (g == 0) ? p.ef : 1u;  // hnsw.rs:286 (ef = 1 above the base)
      beam_search2<R, TAG16, COUNT>(gv, qr, ef, cur, ws, b, c, status);
      b.sort_prefix(g > 0 ? 1u : p.k);  // pop_min() above the base, drain_asc().take(k) at the base
      if (g > 0 && status == 0) {
        uint32_t d0, k0;
        b.at(0, d0, k0);
        cur = __ldg(p.down + (k0 - p.n_base));  // hnsw.rs:288-292 (.point.1)
      }
    }

    // ---- drain_asc().take(k) (+ Indexes::search padding) ----
    const uint32_t nres = status ? 0u : min(b.size, p.k);
    uint32_t fd, fk;
    b.at(0, fd, fk);
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const uint32_t j = b.sorted_index(r);
      if (j < nres) {
        p.keys[(size_t)q * p.k + j] = b.wk[r];
        p.dists[(size_t)q * p.k + j] = b.wd[r];
      }
    }
    for (uint32_t j = nres + lane; j < p.k; j += 32) {
      const bool pad = (p.flags & HB_PAD_LIKE_CLI) && nres;
      p.keys[(size_t)q * p.k + j] = pad ? fk : 0;
      p.dists[(size_t)q * p.k + j] = pad ? fd : 0;
    }
    if (lane == 0) {
      p.found[q] = nres;
      p.qstatus[q] = status;
      if (status) atomicOr(p.ctl + 1, 1u);
#ifdef HB_DIAG_TIME
      unsigned long long t_end;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_end));
      if (p.counters) p.counters[q] = make_uint4((uint32_t)t_start, (uint32_t)t_end, blockIdx.x * kWarpsPerBlock + wib, c.X);
#else
      if (COUNT) p.counters[q] = make_uint4(c.X, c.E, c.V, c.A);
#endif
    }
    __syncwarp();
  }
}

struct FastChoice {
  FastLaunch fl;
  size_t smem;
  uint32_t cache_log2;
};
template <int R, bool TAG16>
FastChoice v2_choice(bool count) {
  FastChoice c;
  c.fl.kernel = count ? hnsw_search_v2_kernel<R, TAG16, true> : hnsw_search_v2_kernel<R, TAG16, false>;
  c.cache_log2 = cache2_log2<TAG16, R>();
  c.smem = (size_t)kWarpsPerBlock * warp_scratch2_bytes<TAG16>(c.cache_log2);
  return c;
}
template <bool TAG16>
FastChoice v2_choice_for(uint32_t ef, bool count) {
  if (ef <= 32) return v2_choice<1, TAG16>(count);
  if (ef <= 64) return v2_choice<2, TAG16>(count);
  if (ef <= 128) return v2_choice<4, TAG16>(count);
  return v2_choice<8, TAG16>(count);
}

uint32_t ceil_log2(uint64_t x) {
  uint32_t l = 0;
  while ((1ull << l) < x) ++l;
  return l;
}

// ---- per-call workspace (pooled on the index) ---------------------------------------------------------------------
struct WorkspaceLease {
  const Index& idx;
  std::unique_ptr<SearchWorkspace> ws;
  explicit WorkspaceLease(const Index& i) : idx(i) {
    {
      std::lock_guard<std::mutex> lock(idx.pool_mu);
      if (!idx.ws_pool.empty()) {
        ws = std::move(idx.ws_pool.back());
        idx.ws_pool.pop_back();
      }
    }
    if (!ws) ws.reset(new SearchWorkspace());
  }
  ~WorkspaceLease() {
    std::lock_guard<std::mutex> lock(idx.pool_mu);
    idx.ws_pool.push_back(std::move(ws));
  }
};

void ensure_common(SearchWorkspace& ws, uint64_t nq) {
  if (!ws.ctl) HB_CUDA(cudaMalloc(&ws.ctl, 8 * sizeof(uint32_t)));
  if (!ws.ev0) HB_CUDA(cudaEventCreate(&ws.ev0));
  if (!ws.ev1) HB_CUDA(cudaEventCreate(&ws.ev1));
  if (ws.qcap < nq) {
    cudaFree(ws.qstatus);
    cudaFree(ws.qmap);
    ws.qstatus = ws.qmap = nullptr;
    ws.qcap = 0;
    HB_CUDA(cudaMalloc(&ws.qstatus, nq * 4));
    HB_CUDA(cudaMalloc(&ws.qmap, nq * 4));
    ws.qcap = nq;
  }
}

void ensure_tables(SearchWorkspace& ws, uint32_t vcap_log2, uint32_t slots, uint64_t nq, cudaStream_t stream) {
  if (ws.vcap_log2 != vcap_log2 || ws.slots < slots) {
    cudaFree(ws.vis);
    cudaFree(ws.spill);
    ws.vis = ws.spill = nullptr;
    ws.vcap_log2 = 0;
    ws.slots = 0;
    const size_t words = (size_t)slots << vcap_log2;
    HB_CUDA(cudaMalloc(&ws.vis, words * 4));
    // on the LAUNCH stream: the caller's stream may be non-blocking, i.e. not ordered after the NULL stream
    HB_CUDA(cudaMemsetAsync(ws.vis, 0xFF, words * 4, stream));
    ws.spill_cap = 1u << (vcap_log2 > 2 ? vcap_log2 - 2 : 0);
    HB_CUDA(cudaMalloc(&ws.spill, (size_t)slots * ws.spill_cap * 4));
    ws.vcap_log2 = vcap_log2;
    ws.slots = slots;
  }
  ensure_common(ws, nq);
}

// one launch on `stream`, timed with the workspace's events; returns after the stream has drained, h_ctl filled
template <class Launch>
double timed_launch(SearchWorkspace& ws, cudaStream_t stream, uint32_t* h_ctl, size_t n_ctl, Launch&& launch) {
  HB_CUDA(cudaMemsetAsync(ws.ctl, 0, 8 * sizeof(uint32_t), stream));
  HB_CUDA(cudaEventRecord(ws.ev0, stream));
  launch();
  count_launches();
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaEventRecord(ws.ev1, stream));
  HB_CUDA(cudaMemcpyAsync(h_ctl, ws.ctl, n_ctl * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  HB_CUDA(cudaStreamSynchronize(stream));
  float ms = 0.f;
  HB_CUDA(cudaEventElapsedTime(&ms, ws.ev0, ws.ev1));
  return ms;
}

// which ef <= 256 kernel serves this index: 16-bit cache tags while they are exact for n_total, full ids above
FastChoice choose_fast(const Index& idx, uint32_t ef, bool count) {
  const FastChoice tagged = v2_choice_for<true>(ef, count);
  if (idx.n_total <= SeenCache<true>::max_ids(tagged.cache_log2)) return tagged;
  return v2_choice_for<false>(ef, count);
}

}  // namespace

void knns_device(const Index& idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                 void* d_dists, void* d_found, void* d_counters, cudaStream_t stream) {
  HB_REQUIRE(idx.algo == HB_ALGO_HNSW || idx.algo == HB_ALGO_NSW, HB_ERR_INVALID, "graph search on a non-graph index");
  HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
  HB_REQUIRE(ef <= HB_MAX_EF, HB_ERR_UNSUPPORTED, "ef above HB_MAX_EF");
  HB_REQUIRE(nq <= 0xFFFFFFF0ull, HB_ERR_UNSUPPORTED, "too many queries in one call");
  if (nq == 0) return;
  HB_CUDA(cudaSetDevice(idx.device));
  if (idx.ep_file < 0 || idx.n_base == 0) {  // hnsw.rs:282 / nsw.rs:345: no entry point -> empty results
    HB_CUDA(cudaMemsetAsync(d_keys, 0, nq * k * 8, stream));
    HB_CUDA(cudaMemsetAsync(d_dists, 0, nq * k * 8, stream));
    HB_CUDA(cudaMemsetAsync(d_found, 0, nq * 4, stream));
    if (d_counters) HB_CUDA(cudaMemsetAsync(d_counters, 0, nq * 16, stream));
    HB_CUDA(cudaStreamSynchronize(stream));
    HB_REQUIRE(!(flags & HB_PAD_LIKE_CLI), HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here)");
    return;
  }
  if (ef == 0) ef = 1;  // nsw.rs:167-176 behaves identically for ef = 0 and ef = 1

  WorkspaceLease lease(idx);  // this call's own control words / tables: concurrent callers do not serialise
  SearchWorkspace& ws = *lease.ws;
  double search_ms = 0.0;
  SearchParams p{};
  p.sk = idx.d_sk;
  p.adj = idx.d_adj;
  p.rowoff = idx.d_rowoff;
  p.down = idx.d_down;
  p.queries = static_cast<const uint4*>(d_queries);
  p.keys = static_cast<uint64_t*>(d_keys);
  p.dists = static_cast<uint64_t*>(d_dists);
  p.found = static_cast<uint32_t*>(d_found);
  p.counters = static_cast<uint4*>(d_counters);
  p.stride = idx.stride;
  p.n_base = (uint32_t)idx.n_base;
  p.n_total = (uint32_t)idx.n_total;
  p.n_upper = (uint32_t)idx.g_nodes.size() - 1;
  p.ep_gid = idx.ep_gid;
  p.ef = ef;
  p.k = k;
  p.flags = flags;
  p.ef_cap = (ef + 31u) / 32u * 32u;

  uint64_t todo = nq;
  bool retry = false;
  std::vector<uint32_t> h_status, h_map;

  // ---- fast path: beam in registers (ef <= 256) ------------------------------------------------
  static const bool force_large = std::getenv("HB_FORCE_LARGE") != nullptr;
  if (ef <= 256 && !force_large && !(flags & HB_EXACT_VISITED)) {
    const FastChoice fc = choose_fast(idx, ef, d_counters != nullptr);
    HB_CUDA(cudaFuncSetAttribute(fc.fl.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fc.smem));
    int fbps = 0;
    HB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fbps, fc.fl.kernel, kWarpsPerBlock * 32, fc.smem));
    HB_REQUIRE(fbps >= 1, HB_ERR_INTERNAL, "fast search kernel does not fit on an SM");
    const uint32_t fblocks = (uint32_t)std::min<uint64_t>((uint64_t)idx.sm_count * fbps, (nq + kWarpsPerBlock - 1) / kWarpsPerBlock);
    ensure_common(ws, nq);
    p.ctl = ws.ctl;
    p.qstatus = ws.qstatus;
    p.qmap = nullptr;
    p.nq = (uint32_t)nq;
    uint32_t h_ctl[4] = {0, 0, 0, 0};
    search_ms += timed_launch(ws, stream, h_ctl, 4, [&] { fc.fl.kernel<<<fblocks, kWarpsPerBlock * 32, fc.smem, stream>>>(p, fc.cache_log2); });
#ifdef HB_DEBUG_BOUNDS
    HB_REQUIRE(h_ctl[2] == 0, HB_ERR_INTERNAL, "bounds check " + std::to_string(h_ctl[2]) + " failed in the ef <= 256 kernel, value " + std::to_string(h_ctl[3]));
#endif
    if (!h_ctl[1]) {
      idx.last_search_ms = search_ms;
      return;
    }
    // boundary-tie stack overflow (tie-heavy data): rerun exactly those queries on the large-capacity kernel
    h_status.resize(nq);
    HB_CUDA(cudaMemcpy(h_status.data(), ws.qstatus, nq * 4, cudaMemcpyDeviceToHost));
    for (uint64_t i = 0; i < nq; ++i) if (h_status[i]) h_map.push_back((uint32_t)i);
    HB_REQUIRE(!h_map.empty(), HB_ERR_INTERNAL, "overflow flag set without an overflowed query");
    HB_CUDA(cudaMemcpy(ws.qmap, h_map.data(), h_map.size() * 4, cudaMemcpyHostToDevice));
    todo = h_map.size();
    retry = true;
  }

  // ---- large-capacity kernel: beam in shared memory, exact visited tables in global memory -----------------------
  const size_t smem = (size_t)kWarpsPerBlock * p.ef_cap * 8 + (size_t)kWarpsPerBlock * 32 * 4;
  HB_CUDA(cudaFuncSetAttribute(hnsw_search_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int blocks_per_sm = 0;
  HB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, hnsw_search_kernel, kWarpsPerBlock * 32, smem));
  HB_REQUIRE(blocks_per_sm >= 1, HB_ERR_UNSUPPORTED, "beam does not fit in shared memory");
  const uint32_t max_blocks = (uint32_t)(idx.sm_count * blocks_per_sm);
  // visited-table capacity: V grows ~ 50 * ef on M=256 graphs; start at 128 * ef and quadruple on overflow
  uint32_t vcap_log2 = std::max(12u, ceil_log2((uint64_t)ef * 128u));
  const uint32_t vcap_max_log2 = std::min(31u, std::max(vcap_log2, ceil_log2(idx.n_total * 2 + 64)));
  vcap_log2 = std::min(vcap_log2, vcap_max_log2);

  for (;;) {
    uint32_t blocks = (uint32_t)std::min<uint64_t>(max_blocks, (todo + kWarpsPerBlock - 1) / kWarpsPerBlock);
    // bound scratch memory for the (rare) big-table retries
    while (blocks > 1 && ((size_t)blocks * kWarpsPerBlock << vcap_log2) * 4 > (size_t)2 << 30) blocks = (blocks + 1) / 2;
    ensure_tables(ws, vcap_log2, blocks * kWarpsPerBlock, nq, stream);
    p.vis = ws.vis;
    p.spill = ws.spill;
    p.ctl = ws.ctl;
    p.qstatus = ws.qstatus;
    p.qmap = retry ? ws.qmap : nullptr;
    p.nq = (uint32_t)todo;
    p.vcap_log2 = ws.vcap_log2;
    p.spill_cap = ws.spill_cap;
    uint32_t h_ctl[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    search_ms += timed_launch(ws, stream, h_ctl, 8, [&] { hnsw_search_kernel<<<blocks, kWarpsPerBlock * 32, smem, stream>>>(p); });
#ifdef HB_DEBUG_BOUNDS
    HB_REQUIRE(h_ctl[2] == 0, HB_ERR_INTERNAL,
               "bounds check " + std::to_string(h_ctl[2]) + " value " + std::to_string(h_ctl[3]) + " query " + std::to_string(h_ctl[4]) +
                   " vcap_log2 " + std::to_string(vcap_log2) + " retry " + std::to_string((int)retry) + " ckey " + std::to_string(h_ctl[5]) + " X " + std::to_string(h_ctl[6]) + " g/sp/size " + std::to_string(h_ctl[7]));
#endif
    if (!h_ctl[1]) break;
    // some queries outgrew their visited table / tie stack: rerun exactly those with 4x the capacity
    HB_REQUIRE(vcap_log2 < vcap_max_log2, HB_ERR_INTERNAL, "visited table overflow at maximum capacity");
    h_status.resize(nq);
    HB_CUDA(cudaMemcpy(h_status.data(), ws.qstatus, nq * 4, cudaMemcpyDeviceToHost));
    std::vector<uint32_t> next;
    if (!retry) {
      for (uint64_t i = 0; i < nq; ++i) if (h_status[i]) next.push_back((uint32_t)i);
    } else {
      for (uint32_t qidx : h_map) if (h_status[qidx]) next.push_back(qidx);
    }
    h_map.swap(next);
    HB_REQUIRE(!h_map.empty(), HB_ERR_INTERNAL, "overflow flag set without an overflowed query");
    HB_CUDA(cudaMemcpy(ws.qmap, h_map.data(), h_map.size() * 4, cudaMemcpyHostToDevice));
    todo = h_map.size();
    retry = true;
    vcap_log2 = std::min(vcap_log2 + 2, vcap_max_log2);
  }
  idx.last_search_ms = search_ms;
}

// ---------------------------------------------------------------------------------------------
// hb_knns_device_noalloc: the ef <= 256 search with its scratch (the work counter, the overflow flag and the per-query
// status words) in a file-scope device array instead of a pooled cudaMalloc'd workspace.  For callers that cannot
// allocate (CUDA graph capture around them, serving loops, harnesses that forbid it).  No counterpart in the reference;
// added at the request of the run's operators.  Every GPU that loads the library holds the array until exit; all
// calls on a device share it, so calls that overlap on the GPU race.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr uint32_t kNoAllocCtlWords = 64;  // control words, padded to one 256-byte block
static __device__ __align__(256) uint32_t g_noalloc_scratch[kNoAllocCtlWords + HB_NOALLOC_MAX_QUERIES];

// The array's device address on `device` (which must be current).  In host code the symbol's name is a host-side
// stand-in, not a device address: the runtime resolves it per device.
uint32_t* noalloc_scratch(int device) {
  constexpr int kMaxDevices = 64;
  static std::mutex mu;
  static uint32_t* addr[kMaxDevices] = {};
  HB_REQUIRE(device >= 0 && device < kMaxDevices, HB_ERR_UNSUPPORTED, "device ordinal above 63");
  std::lock_guard<std::mutex> lock(mu);
  if (!addr[device]) {
    uint32_t* ptr = nullptr;
    HB_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_noalloc_scratch));
    HB_REQUIRE(ptr != nullptr, HB_ERR_INTERNAL, "cudaGetSymbolAddress returned a null scratch address");
    addr[device] = ptr;
  }
  return addr[device];
}
}  // namespace

void knns_device_noalloc(const Index& idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                         void* d_dists, void* d_found, cudaStream_t stream) {
  HB_REQUIRE(idx.algo == HB_ALGO_HNSW || idx.algo == HB_ALGO_NSW, HB_ERR_INVALID, "graph search on a non-graph index");
  HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
  HB_REQUIRE(ef <= 256, HB_ERR_UNSUPPORTED, "hb_knns_device_noalloc: ef above 256 needs visited tables in allocated memory (use hb_knns_device)");
  HB_REQUIRE(!(flags & HB_EXACT_VISITED), HB_ERR_UNSUPPORTED, "hb_knns_device_noalloc: HB_EXACT_VISITED needs allocated visited tables");
  HB_REQUIRE(nq <= HB_NOALLOC_MAX_QUERIES, HB_ERR_UNSUPPORTED,
             "hb_knns_device_noalloc: " + std::to_string(nq) + " queries exceed the static scratch (HB_NOALLOC_MAX_QUERIES = " +
                 std::to_string((unsigned)HB_NOALLOC_MAX_QUERIES) + "); split the batch or use hb_knns_device");
  if (nq == 0) return;
  HB_CUDA(cudaSetDevice(idx.device));
  if (idx.ep_file < 0 || idx.n_base == 0) {  // hnsw.rs:282 / nsw.rs:345: no entry point -> empty results
    HB_CUDA(cudaMemsetAsync(d_keys, 0, nq * k * 8, stream));
    HB_CUDA(cudaMemsetAsync(d_dists, 0, nq * k * 8, stream));
    HB_CUDA(cudaMemsetAsync(d_found, 0, nq * 4, stream));
    HB_CUDA(cudaStreamSynchronize(stream));
    HB_REQUIRE(!(flags & HB_PAD_LIKE_CLI), HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here)");
    return;
  }
  if (ef == 0) ef = 1;
  uint32_t* scratch = noalloc_scratch(idx.device);

  SearchParams p{};
  p.sk = idx.d_sk;
  p.adj = idx.d_adj;
  p.rowoff = idx.d_rowoff;
  p.down = idx.d_down;
  p.queries = static_cast<const uint4*>(d_queries);
  p.keys = static_cast<uint64_t*>(d_keys);
  p.dists = static_cast<uint64_t*>(d_dists);
  p.found = static_cast<uint32_t*>(d_found);
  p.counters = nullptr;
  p.stride = idx.stride;
  p.n_base = (uint32_t)idx.n_base;
  p.n_total = (uint32_t)idx.n_total;
  p.n_upper = (uint32_t)idx.g_nodes.size() - 1;
  p.ep_gid = idx.ep_gid;
  p.ef = ef;
  p.k = k;
  p.flags = flags;
  p.ef_cap = (ef + 31u) / 32u * 32u;
  p.ctl = scratch;                        // [0] query counter, [1] overflow flag (+ the debug words)
  p.qstatus = scratch + kNoAllocCtlWords;  // every query's word is written by the kernel: no clearing needed
  p.qmap = nullptr;
  p.nq = (uint32_t)nq;

  const FastChoice fc = choose_fast(idx, ef, false);
  HB_CUDA(cudaFuncSetAttribute(fc.fl.kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fc.smem));
  int fbps = 0;
  HB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&fbps, fc.fl.kernel, kWarpsPerBlock * 32, fc.smem));
  HB_REQUIRE(fbps >= 1, HB_ERR_INTERNAL, "fast search kernel does not fit on an SM");
  const uint32_t fblocks = (uint32_t)std::min<uint64_t>((uint64_t)idx.sm_count * fbps, (nq + kWarpsPerBlock - 1) / kWarpsPerBlock);
  // the array keeps the last call's contents: the counter and the flag must start at zero
  HB_CUDA(cudaMemsetAsync(scratch, 0, kNoAllocCtlWords * sizeof(uint32_t), stream));
  fc.fl.kernel<<<fblocks, kWarpsPerBlock * 32, fc.smem, stream>>>(p, fc.cache_log2);
  count_launches();
  HB_CUDA(cudaGetLastError());
  uint32_t h_ctl[4] = {0, 0, 0, 0};
  HB_CUDA(cudaMemcpyAsync(h_ctl, scratch, sizeof(h_ctl), cudaMemcpyDeviceToHost, stream));
  HB_CUDA(cudaStreamSynchronize(stream));
  // tie-heavy data can overflow a query's boundary-tie stack; hb_knns_device reruns those queries with tables it
  // allocates, which this entry point may not do: their rows are empty (found = 0) and the call says so
  HB_REQUIRE(!h_ctl[1], HB_ERR_UNSUPPORTED,
             "hb_knns_device_noalloc: some queries overflowed the boundary-tie stack (their found[] is 0); rerun the batch with hb_knns_device");
}

}  // namespace hb
