// oracle/oracle_capi.cpp
//
// TEST INFRASTRUCTURE ONLY (see hnsw_oracle.hpp header).  C entry points over
// the CPU restatement so that tests/, bench.py's cpu_baseline leg and
// __graft_entry__.smoke() can drive it through ctypes.  Also holds:
//   * the bincode 1.3.3 `.idx` reader/writer (cli/src/main.rs:224-248, 480-484;
//     SURVEY Appendix A),
//   * the integer-exact synthetic SimHash sketch generator, and
//   * runners for the reference's own unit-test scenarios (golden vectors).
#include "hnsw_oracle.hpp"

#include <memory>
#include <variant>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace orc;

namespace {

thread_local std::string g_err;

// ---------------------------------------------------------------------------
// IndexFile<Sketch> -- cli/src/main.rs:311-320 (ResultAttrs), 441-446 (Indexes), 480-484
// ---------------------------------------------------------------------------
struct ResultAttrs {
  bool format_size = true;
  std::string data = "hamming";
  uint64_t size = 0;
  uint32_t algo = 0;  // 0 Bruteforce, 1 Nsw, 2 Hnsw  (main.rs:381-389)
  double buildtime = 0, querytime = 0;
  std::string params;
};

using BF = Bruteforce<Sketch, SketchDistance>;
using NSWc = NSW<Sketch, SketchDistance, CsrGraph>;
using HNSWc = HNSW<Sketch, SketchDistance, CsrGraph>;

struct Index {
  ResultAttrs attrs;
  std::variant<BF, NSWc, HNSWc> v;
  uint64_t size() const {
    return std::visit([](auto& x) { return (uint64_t)x.size(); }, v);
  }
};

// --- bincode 1.x legacy: LE fixed-width ints, u64 length prefixes, u32 enum tags
struct Reader {
  FILE* f;
  void raw(void* p, size_t n) {
    if (n && fread(p, 1, n, f) != n) throw std::runtime_error("unexpected end of .idx file");
  }
  uint64_t u64() { uint64_t x; raw(&x, 8); return x; }
  uint32_t u32() { uint32_t x; raw(&x, 4); return x; }
  uint8_t u8() { uint8_t x; raw(&x, 1); return x; }
  double f64() { double x; raw(&x, 8); return x; }
  std::string str() { uint64_t n = u64(); if (n > (1u << 20)) throw std::runtime_error("string too long"); std::string s(n, 0); raw(s.data(), n); return s; }
};
struct Writer {
  FILE* f;
  void raw(const void* p, size_t n) {
    if (n && fwrite(p, 1, n, f) != n) throw std::runtime_error("short write");
  }
  void u64(uint64_t x) { raw(&x, 8); }
  void u32(uint32_t x) { raw(&x, 4); }
  void u8(uint8_t x) { raw(&x, 1); }
  void f64(double x) { raw(&x, 8); }
  void str(const std::string& s) { u64(s.size()); raw(s.data(), s.size()); }
};

void read_adj(Reader& r, uint64_t n_nodes, std::vector<uint64_t>& off, std::vector<Idx>& ids) {
  uint64_t n = r.u64();
  if (n != n_nodes) throw std::runtime_error("adj_lists length != nodes length");
  off.assign(n + 1, 0);
  std::vector<uint64_t> buf;
  for (uint64_t i = 0; i < n; ++i) {
    uint64_t deg = r.u64();
    if (deg > (1ull << 32)) throw std::runtime_error("implausible degree");
    size_t at = ids.size();
    ids.resize(at + deg);
    r.raw(ids.data() + at, deg * 8);
    off[i + 1] = off[i] + deg;
  }
  uint64_t empty = r.u64();  // simplegraph.rs:13, always an empty set
  for (uint64_t i = 0; i < empty; ++i) r.u64();
}
void read_graph(Reader& r, CsrGraph<Sketch>& g) {
  uint64_t n = r.u64();
  g.nodes.resize(n);
  r.raw(g.nodes.data(), n * sizeof(Sketch));
  read_adj(r, n, g.off, g.ids);
}
void read_graph(Reader& r, CsrGraph<std::pair<Sketch, Idx>>& g) {
  uint64_t n = r.u64();
  g.nodes.resize(n);
  for (uint64_t i = 0; i < n; ++i) { r.raw(&g.nodes[i].first, sizeof(Sketch)); g.nodes[i].second = r.u64(); }
  read_adj(r, n, g.off, g.ids);
}
template <class Node>
void write_adj(Writer& w, const CsrGraph<Node>& g) {
  w.u64(g.size());
  for (size_t i = 0; i < g.size(); ++i) {
    uint64_t deg = g.off[i + 1] - g.off[i];
    w.u64(deg);
    w.raw(g.ids.data() + g.off[i], deg * 8);
  }
  w.u64(0);
}
void write_graph(Writer& w, const CsrGraph<Sketch>& g) {
  w.u64(g.size());
  w.raw(g.nodes.data(), g.size() * sizeof(Sketch));
  write_adj(w, g);
}
void write_graph(Writer& w, const CsrGraph<std::pair<Sketch, Idx>>& g) {
  w.u64(g.size());
  for (auto& n : g.nodes) { w.raw(&n.first, sizeof(Sketch)); w.u64(n.second); }
  write_adj(w, g);
}
std::optional<Idx> read_opt(Reader& r) {
  uint8_t t = r.u8();
  if (t == 0) return std::nullopt;
  if (t != 1) throw std::runtime_error("bad Option tag");
  return r.u64();
}
void write_opt(Writer& w, const std::optional<Idx>& o) {
  if (o) { w.u8(1); w.u64(*o); } else w.u8(0);
}

std::unique_ptr<Index> load_idx(const char* path) {
  FILE* f = fopen(path, "rb");
  if (!f) throw std::runtime_error(std::string("cannot open ") + path);
  std::vector<char> iobuf(1 << 22);
  setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
  auto idx = std::make_unique<Index>();
  try {
    Reader r{f};
    auto& a = idx->attrs;
    a.format_size = r.u8() != 0;
    a.data = r.str();
    a.size = r.u64();
    a.algo = r.u32();
    a.buildtime = r.f64();
    a.querytime = r.f64();
    a.params = r.str();
    uint32_t tag = r.u32();
    if (tag == 0) {
      BF bf;
      uint64_t n = r.u64();
      bf.points.resize(n);
      r.raw(bf.points.data(), n * sizeof(Sketch));
      idx->v = std::move(bf);
    } else if (tag == 1) {
      NSWc nsw;
      read_graph(r, nsw.graph);
      nsw.ep = read_opt(r);
      idx->v = std::move(nsw);
    } else if (tag == 2) {
      HNSWc h;
      uint64_t L = r.u64();
      if (L > 64) throw std::runtime_error("implausible layer count");
      h.layers.resize(L);
      for (auto& l : h.layers) read_graph(r, l);
      read_graph(r, h.base);
      h.ep = read_opt(r);
      idx->v = std::move(h);
    } else {
      throw std::runtime_error("bad Indexes tag");
    }
  } catch (...) {
    fclose(f);
    throw;
  }
  fclose(f);
  return idx;
}

void save_idx(const Index& idx, const char* path) {
  FILE* f = fopen(path, "wb");
  if (!f) throw std::runtime_error(std::string("cannot create ") + path);
  std::vector<char> iobuf(1 << 22);
  setvbuf(f, iobuf.data(), _IOFBF, iobuf.size());
  try {
    Writer w{f};
    auto& a = idx.attrs;
    w.u8(a.format_size ? 1 : 0);
    w.str(a.data);
    w.u64(a.size);
    w.u32(a.algo);
    w.f64(a.buildtime);
    w.f64(a.querytime);
    w.str(a.params);
    w.u32((uint32_t)idx.v.index());
    if (auto* bf = std::get_if<BF>(&idx.v)) {
      w.u64(bf->points.size());
      w.raw(bf->points.data(), bf->points.size() * sizeof(Sketch));
    } else if (auto* nsw = std::get_if<NSWc>(&idx.v)) {
      write_graph(w, nsw->graph);
      write_opt(w, nsw->ep);
    } else {
      auto& h = std::get<HNSWc>(idx.v);
      w.u64(h.layers.size());
      for (auto& l : h.layers) write_graph(w, l);
      write_graph(w, h.base);
      write_opt(w, h.ep);
    }
  } catch (...) {
    fclose(f);
    throw;
  }
  fclose(f);
}

std::string params_string(uint64_t efc, uint64_t m, uint64_t M) {  // main.rs:171-174
  char buf[128];
  snprintf(buf, sizeof buf, "index=(efc=%llu,m=%llu,M=%llu)", (unsigned long long)efc, (unsigned long long)m, (unsigned long long)M);
  return buf;
}

// Indexes::search -- cli/src/main.rs:457-477 (dispatch; padding optional)
std::vector<Dist> index_search(const Index& idx, const Sketch& q, size_t k, size_t ef, bool pad, Counters* c) {
  std::vector<Dist> res;
  if (auto* bf = std::get_if<BF>(&idx.v)) res = bf->search(q, k);
  else if (auto* nsw = std::get_if<NSWc>(&idx.v)) res = nsw->search(q, k, ef, c);
  else res = std::get<HNSWc>(idx.v).search(q, k, ef, c);
  if (pad) pad_like_cli(res, k);
  return res;
}

// ---------------------------------------------------------------------------
// Synthetic sketches: SimHash of a latent Gaussian-ish mixture in exact integer
// arithmetic, so that this C code, numpy and the CUDA generator agree bit for
// bit (SURVEY 8d explains why uniform random sketches are useless here).
// ---------------------------------------------------------------------------
inline uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
inline uint64_t stream_key(uint64_t seed, uint64_t stream) { return mix64(seed ^ mix64(stream + 0x632BE59BD9B4E019ull)); }
inline uint64_t hash_at(uint64_t key, uint64_t idx) { return mix64(key + (idx + 1) * 0x9E3779B97F4A7C15ull); }
inline int32_t gauss8(uint64_t r) {  // sum of 8 signed bytes (+4 to centre): ~N(0, 209^2), |x| <= 1028
  int32_t s = 4;
  for (int b = 0; b < 8; ++b) s += (int8_t)(r >> (8 * b));
  return s;
}

struct SynthParams {
  uint64_t seed_planes, seed_centres, seed_rows;
  uint32_t latent_dim, n_clusters, centre_scale, noise_scale, distinct_planes;
};

void synth_rows(const SynthParams& p, uint64_t start, uint64_t n, uint64_t* out) {
  const uint32_t L = p.latent_dim;
  const uint64_t kp = stream_key(p.seed_planes, 0), kc = stream_key(p.seed_centres, 0);
  const uint64_t kr0 = stream_key(p.seed_rows, 0), kr1 = stream_key(p.seed_rows, 1);
  std::vector<int32_t> planes((size_t)p.distinct_planes * L);
  for (size_t i = 0; i < planes.size(); ++i) planes[i] = gauss8(hash_at(kp, i));
#pragma omp parallel
  {
    std::vector<int32_t> z(L);
#pragma omp for schedule(static)
    for (long long ii = 0; ii < (long long)n; ++ii) {
      uint64_t i = start + (uint64_t)ii;
      uint64_t cl = hash_at(kr0, i) % p.n_clusters;
      for (uint32_t t = 0; t < L; ++t)
        z[t] = (int32_t)p.centre_scale * gauss8(hash_at(kc, cl * L + t)) + (int32_t)p.noise_scale * gauss8(hash_at(kr1, i * L + t));
      uint64_t* row = out + (size_t)ii * 16;
      for (int w = 0; w < 16; ++w) {
        uint64_t word = 0;
        for (int b = 0; b < 64; ++b) {
          uint32_t pj = (uint32_t)(w * 64 + b) % p.distinct_planes;
          const int32_t* a = &planes[(size_t)pj * L];
          int64_t s = 0;
          for (uint32_t t = 0; t < L; ++t) s += (int64_t)a[t] * z[t];
          if (s > 0) word |= 1ull << b;
        }
        row[w] = word;
      }
    }
  }
}


// ---------------------------------------------------------------------------
// Scalar model of the BOUNDED beam formulation the GPU kernel uses (DESIGN.md
// "bounded candidates"): w is one sorted array carrying an expanded flag per
// element; the reference's unbounded `cands` heap is replaced by "first
// unexpanded element of w" plus a stack of evicted-but-unexpanded elements
// whose distance equals the current worst distance F.  tests/ proves it equal
// to orc::search (the restatement of nsw.rs:133-184) on tie-heavy inputs
// before the CUDA kernel is trusted with it.
// ---------------------------------------------------------------------------
uint64_t g_spill_pops = 0;  // how often the boundary-tie stack supplied a candidate (test visibility)
// 0: exact visited set (nsw.rs:143).  c > 0: the kernel's formulation -- a direct-mapped cache of 2^c
// recently seen ids (a hit is certainly visited; a miss may be a revisit) plus "already in w" dedupe.
// A revisited node that is not in w was either rejected or evicted earlier, so its distance is >= the
// current F of a full beam and it is rejected again: results, X, E and A are unchanged, only V grows.
int g_cache_log2 = 0;
struct VisitCache {
  std::vector<uint32_t> tab;
  uint32_t shift = 0;
  explicit VisitCache(int log2) : tab((size_t)1 << log2, 0xFFFFFFFFu), shift(32u - (uint32_t)log2) {}
  bool test_and_set(uint32_t gid) {  // true when gid was cached (certainly visited)
    uint32_t h = (gid * 0x9E3779B1u) >> shift;
    if (tab[h] == gid) return true;
    tab[h] = gid;
    return false;
  }
};
template <class G, class Q, class DF>
std::vector<Dist> search_bounded(const G& graph, const Q& query, size_t ef, Idx ep, DF dist, Counters* ctr,
                                 VisitCache* cache = nullptr, uint64_t gid_base = 0) {
  if (ef == 0) ef = 1;
  struct El { Dist d; bool expanded; };
  std::vector<El> w;
  std::vector<uint64_t> spill;  // keys; all share distance spill_d
  uint64_t spill_d = 0;
  VisitedSet visited(2048);
  if (!cache) visited.insert(ep);
  else cache->test_and_set((uint32_t)(gid_base + ep));
  Counters c;
  c.V = 1;
  w.push_back(El{Dist{(uint64_t)dist(*graph.get(ep), query), ep}, false});
  size_t cursor = 0;
  for (;;) {
    uint64_t F = w.back().d.distance;
    size_t idx = cursor;
    while (idx < w.size() && w[idx].expanded) ++idx;
    uint64_t ckey;
    if (idx < w.size()) { w[idx].expanded = true; ckey = w[idx].d.key; cursor = idx + 1; }
    else if (!spill.empty() && spill_d == F && w.size() >= ef) { cursor = w.size(); ckey = spill.back(); spill.pop_back(); ++g_spill_pops; }
    else break;
    ++c.X;
    for (Idx e : graph.neighborhood(ckey)) {
      ++c.E;
      if (!cache) { if (!visited.insert(e)) continue; }
      else if (cache->test_and_set((uint32_t)(gid_base + e))) continue;
      ++c.V;
      Dist ed{(uint64_t)dist(*graph.get(e), query), e};
      F = w.back().d.distance;
      if (ed.distance >= F && w.size() >= ef) continue;
      size_t pos = 0;
      while (pos < w.size() && w[pos].d < ed) ++pos;
      if (cache && pos < w.size() && w[pos].d.key == e) continue;  // revisit of a node that is still in w
      ++c.A;
      bool full = w.size() >= ef;
      El ev{};
      if (full) { ev = w.back(); w.pop_back(); }
      w.insert(w.begin() + pos, El{ed, false});
      if (pos < cursor) cursor = pos;
      if (full && !ev.expanded && ev.d.distance == w.back().d.distance) {
        if (!spill.empty() && spill_d != ev.d.distance) spill.clear();
        spill_d = ev.d.distance;
        spill.push_back(ev.d.key);
      }
    }
  }
  if (ctr) ctr->add(c);
  std::vector<Dist> out;
  for (auto& e : w) out.push_back(e.d);
  return out;
}
std::vector<Dist> hnsw_search_bounded(const HNSWc& h, const Sketch& q, size_t k, size_t ef, Counters* ctr) {
  if (!h.ep) return {};
  Idx cur = *h.ep;
  auto ldist = [&](const std::pair<Sketch, Idx>& n, const Sketch& qq) { return SketchDistance{}(n.first, qq); };
  std::optional<VisitCache> cache;
  if (g_cache_log2 > 0) cache.emplace(g_cache_log2);
  VisitCache* cp = cache ? &*cache : nullptr;
  std::vector<uint64_t> gbase(h.layers.size() + 1, 0);  // one id space over all graphs, base first (as on the device)
  gbase[0] = 0;
  for (size_t l = 0; l < h.layers.size(); ++l) gbase[l + 1] = gbase[l] + (l == 0 ? h.base.size() : h.layers[l - 1].size());
  for (size_t l = h.layers.size(); l-- > 0;) {
    auto w = search_bounded(h.layers[l], q, 1, cur, ldist, ctr, cp, gbase[l + 1]);
    cur = h.layers[l].get(w.front().key)->second;
  }
  auto v = search_bounded(h.base, q, ef, cur, SketchDistance{}, ctr, cp, 0);
  if (v.size() > k) v.resize(k);
  return v;
}

int omp_threads_or_all(int t) {
#ifdef _OPENMP
  // all logical CPUs this process may run on, like rayon's default pool (std::thread::available_parallelism);
  // NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1, which would time the baseline on one core
  return t > 0 ? t : omp_get_num_procs();
#else
  (void)t;
  return 1;
#endif
}

template <class F>
int guarded(F f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    g_err = e.what();
    return -1;
  }
}

}  // namespace

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }
int orc_max_threads() { return omp_threads_or_all(0); }

uint64_t orc_hamming(const uint64_t* a, const uint64_t* b) {  // cli/src/sketch.rs:18-29
  return SketchDistance{}(*(const Sketch*)a, *(const Sketch*)b);
}

void* orc_index_load_idx(const char* path) {
  void* out = nullptr;
  guarded([&] { out = load_idx(path).release(); });
  return out;
}
int orc_index_save_idx(const void* h, const char* path) {
  return guarded([&] { save_idx(*(const Index*)h, path); });
}
void orc_index_free(void* h) { delete (Index*)h; }
uint64_t orc_index_size(const void* h) { return ((const Index*)h)->size(); }
int orc_index_algo(const void* h) { return (int)((const Index*)h)->v.index(); }
// number of graphs: 0 for bruteforce, 1 for NSW, 1 + L for HNSW (graph 0 = base, g = layers[g-1])
uint64_t orc_index_num_graphs(const void* h) {
  auto& v = ((const Index*)h)->v;
  if (std::holds_alternative<BF>(v)) return 0;
  if (std::holds_alternative<NSWc>(v)) return 1;
  return 1 + std::get<HNSWc>(v).layers.size();
}
int64_t orc_index_ep(const void* h) {
  auto& v = ((const Index*)h)->v;
  if (auto* n = std::get_if<NSWc>(&v)) return n->ep ? (int64_t)*n->ep : -1;
  if (auto* x = std::get_if<HNSWc>(&v)) return x->ep ? (int64_t)*x->ep : -1;
  return -1;
}
void orc_index_graph_info(const void* h, uint64_t g, uint64_t* n_nodes, uint64_t* n_edges) {
  auto& v = ((const Index*)h)->v;
  if (auto* n = std::get_if<NSWc>(&v)) { *n_nodes = n->graph.size(); *n_edges = n->graph.ids.size(); return; }
  if (auto* b = std::get_if<BF>(&v)) { *n_nodes = b->points.size(); *n_edges = 0; return; }
  auto& x = std::get<HNSWc>(v);
  if (g == 0) { *n_nodes = x.base.size(); *n_edges = x.base.ids.size(); }
  else { *n_nodes = x.layers[g - 1].size(); *n_edges = x.layers[g - 1].ids.size(); }
}
// sketches: n*16 u64; down: n u64 (ignored for graph 0 / NSW); off: n+1; ids: n_edges
void orc_index_graph_export(const void* h, uint64_t g, uint64_t* sketches, uint64_t* down, uint64_t* off, uint64_t* ids) {
  auto& v = ((const Index*)h)->v;
  auto copy_plain = [&](const CsrGraph<Sketch>& gr) {
    if (sketches) memcpy(sketches, gr.nodes.data(), gr.size() * sizeof(Sketch));
    if (off) memcpy(off, gr.off.data(), gr.off.size() * 8);
    if (ids) memcpy(ids, gr.ids.data(), gr.ids.size() * 8);
  };
  if (auto* b = std::get_if<BF>(&v)) { if (sketches) memcpy(sketches, b->points.data(), b->points.size() * sizeof(Sketch)); return; }
  if (auto* n = std::get_if<NSWc>(&v)) { copy_plain(n->graph); return; }
  auto& x = std::get<HNSWc>(v);
  if (g == 0) { copy_plain(x.base); return; }
  auto& gr = x.layers[g - 1];
  for (size_t i = 0; i < gr.size(); ++i) {
    if (sketches) memcpy(sketches + i * 16, &gr.nodes[i].first, sizeof(Sketch));
    if (down) down[i] = gr.nodes[i].second;
  }
  if (off) memcpy(off, gr.off.data(), gr.off.size() * 8);
  if (ids) memcpy(ids, gr.ids.data(), gr.ids.size() * 8);
}

// Assemble an index from arrays.  algo: 0 bruteforce, 1 NSW, 2 HNSW.  n_graphs as
// in orc_index_num_graphs; graph 0 = base.  ids are u32 or u64 (id_bytes = 4 | 8).
void* orc_index_from_parts(int algo, uint64_t n_graphs, const uint64_t* n_nodes, const uint64_t* const* sketches,
                           const uint64_t* const* down, const uint64_t* const* off, const void* const* ids, int id_bytes,
                           int64_t ep, uint64_t efc, uint64_t m, uint64_t M) {
  void* out = nullptr;
  guarded([&] {
    auto idx = std::make_unique<Index>();
    idx->attrs.algo = (uint32_t)algo;
    idx->attrs.size = n_nodes[0];
    idx->attrs.params = params_string(efc, m, M);
    auto fill_adj = [&](uint64_t g, std::vector<uint64_t>& o, std::vector<Idx>& d) {
      uint64_t n = n_nodes[g];
      o.assign(off[g], off[g] + n + 1);
      uint64_t ne = o.back();
      d.resize(ne);
      if (id_bytes == 8) memcpy(d.data(), ids[g], ne * 8);
      else for (uint64_t i = 0; i < ne; ++i) d[i] = ((const uint32_t*)ids[g])[i];
    };
    if (algo == 0) {
      BF bf;
      bf.points.resize(n_nodes[0]);
      memcpy(bf.points.data(), sketches[0], n_nodes[0] * sizeof(Sketch));
      idx->v = std::move(bf);
    } else if (algo == 1) {
      NSWc nsw;
      nsw.graph.nodes.resize(n_nodes[0]);
      memcpy(nsw.graph.nodes.data(), sketches[0], n_nodes[0] * sizeof(Sketch));
      fill_adj(0, nsw.graph.off, nsw.graph.ids);
      if (ep >= 0) nsw.ep = (Idx)ep;
      idx->v = std::move(nsw);
    } else if (algo == 2) {
      HNSWc h;
      h.base.nodes.resize(n_nodes[0]);
      memcpy(h.base.nodes.data(), sketches[0], n_nodes[0] * sizeof(Sketch));
      fill_adj(0, h.base.off, h.base.ids);
      h.layers.resize(n_graphs - 1);
      for (uint64_t g = 1; g < n_graphs; ++g) {
        auto& l = h.layers[g - 1];
        l.nodes.resize(n_nodes[g]);
        for (uint64_t i = 0; i < n_nodes[g]; ++i) {
          memcpy(&l.nodes[i].first, sketches[g] + i * 16, sizeof(Sketch));
          l.nodes[i].second = down[g][i];
        }
        fill_adj(g, l.off, l.ids);
      }
      if (ep >= 0) h.ep = (Idx)ep;
      idx->v = std::move(h);
    } else {
      throw std::runtime_error("bad algo");
    }
    out = idx.release();
  });
  return out;
}

// Build with the restated reference builders.  algo as above; parallel = 0 ->
// Extend/add one by one (single_threaded, main.rs:406-410,426-430), 1 ->
// extend_parallel with `rayon_threads` standing in for rayon::current_num_threads().
void* orc_build(int algo, const uint64_t* data, uint64_t n, uint64_t efc, uint64_t m, uint64_t M, int parallel,
                uint64_t rayon_threads, int omp_threads) {
  void* out = nullptr;
  guarded([&] {
#ifdef _OPENMP
    omp_set_num_threads(omp_threads_or_all(omp_threads));
#endif
    const Sketch* pts = (const Sketch*)data;
    auto idx = std::make_unique<Index>();
    idx->attrs.algo = (uint32_t)algo;
    idx->attrs.size = n;
    idx->attrs.params = params_string(efc, m, M);
    NSWOptions o{efc, m, M, n};  // main.rs:399-404,420-425: size = dataset size
    if (algo == 0) {
      BF bf;
      bf.points.assign(pts, pts + n);
      idx->v = std::move(bf);
    } else if (algo == 1) {
      NSWBuilder<Sketch, SketchDistance> b(o);
      if (parallel) b.extend_parallel(pts, n, rayon_threads); else b.extend(pts, pts + n);
      auto built = b.build();
      NSWc c;
      c.graph = to_csr(built.graph);
      c.ep = built.ep;
      idx->v = std::move(c);
    } else if (algo == 2) {
      HNSWBuilder<Sketch, SketchDistance> b(o, rayon_threads);
      if (parallel) b.extend_parallel(pts, n, rayon_threads); else b.extend(pts, pts + n);
      auto built = b.build();
      HNSWc c;
      c.base = to_csr(built.base);
      for (auto& l : built.layers) c.layers.push_back(to_csr(l));
      c.ep = built.ep;
      idx->v = std::move(c);
    } else {
      throw std::runtime_error("bad algo");
    }
    out = idx.release();
  });
  return out;
}

// Index::knns -- src/index/mod.rs:28-42 (rayon par_iter over queries == omp
// dynamic schedule; results in input order) through Indexes::search
// (main.rs:457-477).  keys/dists: nq*k (entries past found[i] are left 0 when
// pad == 0); counters: nq*4 (X,E,V,A) or NULL.
int orc_knns(const void* h, const uint64_t* queries, uint64_t nq, uint64_t k, uint64_t ef, int pad, int omp_threads,
             uint64_t* keys, uint64_t* dists, uint32_t* found, uint64_t* counters) {
  const Index& idx = *(const Index*)h;
  std::string err;
  int nt = omp_threads_or_all(omp_threads);
  (void)nt;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt)
  for (long long i = 0; i < (long long)nq; ++i) {
    try {
      Counters c;
      // found = result count before padding
      auto res = index_search(idx, *(const Sketch*)(queries + i * 16), k, ef, false, &c);
      if (found) found[i] = (uint32_t)res.size();
      if (pad) pad_like_cli(res, k);
      for (size_t j = 0; j < k; ++j) {
        keys[i * k + j] = j < res.size() ? res[j].key : 0;
        dists[i * k + j] = j < res.size() ? res[j].distance : 0;
      }
      if (counters) { counters[i * 4] = c.X; counters[i * 4 + 1] = c.E; counters[i * 4 + 2] = c.V; counters[i * 4 + 3] = c.A; }
    } catch (const std::exception& e) {
#pragma omp critical
      err = e.what();
    }
  }
  if (!err.empty()) { g_err = err; return -1; }
  return 0;
}

uint64_t orc_debug_spill_pops() { return g_spill_pops; }
void orc_debug_set_cache_log2(int c) { g_cache_log2 = c; }
// Same contract as orc_knns but through the bounded-beam model (HNSW / NSW only).
int orc_knns_bounded(const void* h, const uint64_t* queries, uint64_t nq, uint64_t k, uint64_t ef, uint64_t* keys,
                     uint64_t* dists, uint32_t* found, uint64_t* counters) {
  const Index& idx = *(const Index*)h;
  return guarded([&] {
    for (uint64_t i = 0; i < nq; ++i) {
      Counters c;
      const Sketch& q = *(const Sketch*)(queries + i * 16);
      std::vector<Dist> res;
      if (auto* x = std::get_if<HNSWc>(&idx.v)) res = hnsw_search_bounded(*x, q, k, ef, &c);
      else if (auto* n = std::get_if<NSWc>(&idx.v)) {
        if (n->ep) {
          std::optional<VisitCache> cache;
          if (g_cache_log2 > 0) cache.emplace(g_cache_log2);
          res = search_bounded(n->graph, q, ef, *n->ep, SketchDistance{}, &c, cache ? &*cache : nullptr, 0);
          if (res.size() > k) res.resize(k);
        }
      } else throw std::runtime_error("bounded model: graph indexes only");
      found[i] = (uint32_t)res.size();
      for (size_t j = 0; j < k; ++j) {
        keys[i * k + j] = j < res.size() ? res[j].key : 0;
        dists[i * k + j] = j < res.size() ? res[j].distance : 0;
      }
      if (counters) { counters[i * 4] = c.X; counters[i * 4 + 1] = c.E; counters[i * 4 + 2] = c.V; counters[i * 4 + 3] = c.A; }
    }
  });
}

// `ground-truth`: Bruteforce::search per query (no query blocking, as the
// reference does) followed by the CLI's sort (main.rs:738-741).  0-based keys.
int orc_ground_truth(const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint64_t k, int omp_threads,
                     uint64_t* keys, uint64_t* dists) {
  int nt = omp_threads_or_all(omp_threads);
  (void)nt;
  const Sketch* pts = (const Sketch*)data;
  SketchDistance sd;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nt)
  for (long long i = 0; i < (long long)nq; ++i) {
    const Sketch& q = *(const Sketch*)(queries + i * 16);
    struct GenIt {
      const Sketch* pts; const Sketch* q; size_t i;
      Dist operator*() const { return Dist{(uint64_t)SketchDistance{}(*q, pts[i]), i}; }
      GenIt& operator++() { ++i; return *this; }
      bool operator!=(const GenIt& o) const { return i != o.i; }
    };
    (void)sd;
    auto res = min_k<Dist>(GenIt{pts, &q, 0}, GenIt{pts, &q, (size_t)n}, k);
    std::sort(res.begin(), res.end());
    for (size_t j = 0; j < k; ++j) {
      keys[i * k + j] = j < res.size() ? res[j].key : 0;
      dists[i * k + j] = j < res.size() ? res[j].distance : 0;
    }
  }
  return 0;
}

// Bruteforce::search in library-API order (std BinaryHeap vector order, unsorted).
int orc_bruteforce_api_order(const uint64_t* data, uint64_t n, const uint64_t* query, uint64_t k, uint64_t* keys,
                             uint64_t* dists, uint64_t* n_out) {
  BF bf;
  bf.points.assign((const Sketch*)data, (const Sketch*)data + n);
  auto res = bf.search(*(const Sketch*)query, k);
  *n_out = res.size();
  for (size_t j = 0; j < res.size(); ++j) { keys[j] = res[j].key; dists[j] = res[j].distance; }
  return 0;
}

void orc_synth(uint64_t seed_planes, uint64_t seed_centres, uint64_t seed_rows, uint32_t latent_dim, uint32_t n_clusters,
               uint32_t centre_scale, uint32_t noise_scale, uint32_t distinct_planes, uint64_t start, uint64_t n, uint64_t* out) {
  synth_rows(SynthParams{seed_planes, seed_centres, seed_rows, latent_dim, n_clusters, centre_scale, noise_scale, distinct_planes},
             start, n, out);
}

// ---------------------------------------------------------------------------
// Golden-vector scenario runners (the reference's own unit tests, generic metric)
// ---------------------------------------------------------------------------
struct AbsI32 {  // src/index/nsw.rs:360-364
  size_t operator()(const int32_t& a, const int32_t& b) const { return (size_t)std::abs((long long)b - (long long)a); }
};
struct P3 { int32_t x, y, z; };
struct P3Dist {  // examples/main.rs:6-11
  size_t operator()(const P3& a, const P3& b) const {
    long long dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
    return (size_t)(dx * dx + dy * dy + dz * dz);
  }
};

// NSW over i32 points, sequential extend, one search.  Returns #results; out_points[i] = point value.
int64_t orc_golden_nsw_i32(const int32_t* pts, uint64_t n, uint64_t efc, uint64_t m, uint64_t M, int32_t query, uint64_t k,
                           uint64_t ef, int32_t* out_points, uint64_t* out_dists) {
  int64_t ret = -1;
  guarded([&] {
    NSWBuilder<int32_t, AbsI32> b(NSWOptions{efc, m, M, n});
    b.extend(pts, pts + n);
    auto nsw = b.build();
    auto res = nsw.search(query, k, ef);
    for (size_t i = 0; i < res.size(); ++i) { out_points[i] = nsw.graph.nodes[res[i].key]; out_dists[i] = res[i].distance; }
    ret = (int64_t)res.size();
  });
  return ret;
}
int64_t orc_golden_hnsw_i32(const int32_t* pts, uint64_t n, uint64_t efc, uint64_t m, uint64_t M, uint64_t threads, int32_t query,
                            uint64_t k, uint64_t ef, int32_t* out_points, uint64_t* out_dists) {
  int64_t ret = -1;
  guarded([&] {
    HNSWBuilder<int32_t, AbsI32> b(NSWOptions{efc, m, M, n}, threads);
    b.extend(pts, pts + n);
    auto h = b.build();
    auto res = h.search(query, k, ef);
    for (size_t i = 0; i < res.size(); ++i) { out_points[i] = h.base.nodes[res[i].key]; out_dists[i] = res[i].distance; }
    ret = (int64_t)res.size();
  });
  return ret;
}
// select_neighbors over explicit numbers w.r.t. q (nsw.rs:387-412): keys are positions.
int64_t orc_golden_heuristic_i32(const int32_t* numbers, uint64_t n, int32_t q, uint64_t m, int32_t* out_points) {
  int64_t ret = -1;
  guarded([&] {
    MinMaxHeap<Dist> heap;
    AbsI32 d;
    for (uint64_t i = 0; i < n; ++i) heap.push(Dist{(uint64_t)d(numbers[i], q), i});
    auto sel = select_neighbors(std::move(heap), m, [&](Idx k) -> const int32_t& { return numbers[k]; }, d);
    for (size_t i = 0; i < sel.size(); ++i) out_points[i] = numbers[sel[i].key];
    ret = (int64_t)sel.size();
  });
  return ret;
}
int64_t orc_golden_min_k_i32(const int32_t* v, uint64_t n, uint64_t k, int32_t* out) {
  auto r = min_k<int32_t>(v, v + n, k);
  for (size_t i = 0; i < r.size(); ++i) out[i] = r[i];
  return (int64_t)r.size();
}
// HNSW over 3-D integer points (README.md:12-74), sequential extend or extend_parallel.
int64_t orc_golden_hnsw_p3(const int32_t* xyz, uint64_t n, uint64_t efc, uint64_t m, uint64_t M, uint64_t threads, int parallel,
                           const int32_t* query, uint64_t k, uint64_t ef, uint64_t* out_keys, uint64_t* out_dists) {
  int64_t ret = -1;
  guarded([&] {
    std::vector<P3> pts(n);
    for (uint64_t i = 0; i < n; ++i) pts[i] = P3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]};
    HNSWBuilder<P3, P3Dist> b(NSWOptions{efc, m, M, n}, threads);
    if (parallel) b.extend_parallel(pts.data(), n, threads); else b.extend(pts.begin(), pts.end());
    auto h = b.build();
    auto res = h.search(P3{query[0], query[1], query[2]}, k, ef);
    for (size_t i = 0; i < res.size(); ++i) { out_keys[i] = res[i].key; out_dists[i] = res[i].distance; }
    ret = (int64_t)res.size();
  });
  return ret;
}
// Bruteforce over 3-D points + sort (order-free golden for the README vector).
int64_t orc_golden_bruteforce_p3(const int32_t* xyz, uint64_t n, const int32_t* query, uint64_t k, uint64_t* out_keys, uint64_t* out_dists) {
  Bruteforce<P3, P3Dist> bf;
  for (uint64_t i = 0; i < n; ++i) bf.points.push_back(P3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]});
  auto res = bf.search(P3{query[0], query[1], query[2]}, k);
  std::sort(res.begin(), res.end());
  for (size_t i = 0; i < res.size(); ++i) { out_keys[i] = res[i].key; out_dists[i] = res[i].distance; }
  return (int64_t)res.size();
}
// src/collections/simplegraph.rs:130-175 -- returns a bitmask of passed sub-tests (0xF == all)
int orc_golden_simplegraph() {
  int ok = 0;
  auto uneq = [](IdSpan s, std::vector<Idx> want) {
    std::vector<Idx> got(s.begin(), s.end());
    std::sort(got.begin(), got.end());
    std::sort(want.begin(), want.end());
    return got == want;
  };
  {  // test_add
    VecGraph<int> g;
    Idx ten = g.add(10), two = g.add(2);
    bool p = g.size() == 2;
    g.add_edge(ten, two);
    p = p && g.is_connected(two, ten);
    if (p) ok |= 1;
  }
  {  // test_is_connected
    VecGraph<int> g;
    g.add(0); g.add(1);
    bool p = !g.is_connected(0, 1);
    g.add_edge(0, 1);
    p = p && g.is_connected(0, 1);
    if (p) ok |= 2;
  }
  {  // test_neighborhood
    VecGraph<int> g;
    for (int i = 0; i < 10; ++i) g.add(i);
    for (Idx i = 1; i < 6; ++i) g.add_edge(0, i);
    if (uneq(g.neighborhood(0), {1, 2, 3, 4, 5})) ok |= 4;
  }
  {  // test_clear_edges
    VecGraph<int> g;
    for (int i = 0; i < 10; ++i) g.add(i);
    for (Idx i = 1; i < 6; ++i) g.add_edge(0, i);
    for (Idx i = 2; i < 6; ++i) g.add_edge(1, i);
    bool p = uneq(g.neighborhood(0), {1, 2, 3, 4, 5}) && uneq(g.neighborhood(1), {0, 2, 3, 4, 5});
    g.clear_edges(1);
    p = p && uneq(g.neighborhood(0), {2, 3, 4, 5}) && uneq(g.neighborhood(1), {});
    if (p) ok |= 8;
  }
  return ok;
}

}  // extern "C"
