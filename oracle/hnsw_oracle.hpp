// oracle/hnsw_oracle.hpp
//
// TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's algorithm for
// the binary-sketch query path (JonasUJ/hnsw-itu).  Only tests/, bench.py's
// cpu_baseline / --impl reference leg and __graft_entry__.smoke() may use it,
// and only as the checker / timed CPU baseline -- the product library
// (hnsw_itu_b200/csrc) never includes, links or calls anything in oracle/.
//
// Parity status: the reference cannot be compiled here (no cargo/rustc, crates
// un-vendored, no Cargo.lock), so this restatement is pinned against every
// golden vector the reference's own tests hold (tests/test_oracle_golden.py):
//   cli/src/sketch.rs:53-58, src/index/nsw.rs:366-412, src/index/hnsw.rs:309-360,
//   src/collections/mod.rs:100-114, src/collections/simplegraph.rs:130-175,
//   README.md:61-74 / examples/main.rs:53-65.
// What the reference itself leaves unpinned (std HashSet iteration order,
// SURVEY.md F4) is canonicalised here to ".idx file order".
//
// Each function cites the reference file:line it follows (paths relative to
// the reference checkout).
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <optional>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

namespace orc {

using Idx = uint64_t;  // src/collections/mod.rs:11  (usize)

// ---------------------------------------------------------------------------
// Distance: src/index/mod.rs:49-97.  Total order = (distance, key).
// The `point` reference of the original is recovered through `key`.
// ---------------------------------------------------------------------------
struct Dist {
  uint64_t distance;
  uint64_t key;
};
inline bool operator<(const Dist& a, const Dist& b) {  // mod.rs:90-97
  return a.distance < b.distance || (a.distance == b.distance && a.key < b.key);
}
inline bool operator>(const Dist& a, const Dist& b) { return b < a; }
inline bool operator<=(const Dist& a, const Dist& b) { return !(b < a); }
inline bool operator>=(const Dist& a, const Dist& b) { return !(a < b); }

// ---------------------------------------------------------------------------
// Sketch: cli/src/sketch.rs:7-29.  16 little-endian u64 words, Hamming.
// ---------------------------------------------------------------------------
struct Sketch {
  uint64_t data[16];
};
struct SketchDistance {  // sketch.rs:18-29
  size_t operator()(const Sketch& a, const Sketch& b) const {
    size_t acc = 0;
    for (int i = 0; i < 16; ++i) acc += (size_t)__builtin_popcountll(a.data[i] ^ b.data[i]);
    return acc;
  }
};

// Per-search counters.  V mirrors the reference's own "visited" trace
// (src/index/nsw.rs:180-181); X/E/A are added for the roofline byte model
// (SURVEY.md 8d).
struct Counters {
  uint64_t X = 0;  // expansions (candidates popped and expanded)
  uint64_t E = 0;  // neighbour ids iterated
  uint64_t V = 0;  // distance evaluations == |visited|
  uint64_t A = 0;  // accepted pushes into w/cands (excluding the entry point)
  void add(const Counters& o) { X += o.X; E += o.E; V += o.V; A += o.A; }
};

// ---------------------------------------------------------------------------
// Double-ended priority queue with the ordering contract of the
// `min-max-heap` crate 1.3.0 (third-party, not vendored in the reference):
// push / pop_min / pop_max / peek_min / peek_max / drain_asc.  Elements are
// totally ordered and unique on the search path, so only the
// contract matters; this is a textbook Atkinson min-max heap.
// ---------------------------------------------------------------------------
template <class T>
class MinMaxHeap {
 public:
  MinMaxHeap() = default;
  explicit MinMaxHeap(std::vector<T> init) {
    for (auto& x : init) push(x);
  }
  size_t size() const { return v_.size(); }
  bool empty() const { return v_.empty(); }
  void push(const T& x) {
    v_.push_back(x);
    bubble_up(v_.size() - 1);
  }
  const T& peek_min() const { return v_[0]; }
  const T& peek_max() const { return v_[max_index()]; }
  T pop_min() {
    T out = v_[0];
    v_[0] = v_.back();
    v_.pop_back();
    if (!v_.empty()) trickle_down(0);
    return out;
  }
  T pop_max() {
    size_t i = max_index();
    T out = v_[i];
    v_[i] = v_.back();
    v_.pop_back();
    if (i < v_.size()) trickle_down(i);
    return out;
  }
  std::vector<T> drain_asc() {
    std::vector<T> out;
    out.reserve(v_.size());
    while (!v_.empty()) out.push_back(pop_min());
    return out;
  }
  const std::vector<T>& raw() const { return v_; }

 private:
  std::vector<T> v_;
  static bool min_level(size_t i) { return ((63 - __builtin_clzll((unsigned long long)(i + 1))) & 1) == 0; }
  size_t max_index() const {
    if (v_.size() <= 1) return 0;
    if (v_.size() == 2) return 1;
    return (v_[1] < v_[2]) ? 2 : 1;
  }
  void bubble_up(size_t i) {
    if (i == 0) return;
    size_t p = (i - 1) / 2;
    if (min_level(i)) {
      if (v_[p] < v_[i]) { std::swap(v_[p], v_[i]); bubble_up_dir<false>(p); }
      else bubble_up_dir<true>(i);
    } else {
      if (v_[i] < v_[p]) { std::swap(v_[p], v_[i]); bubble_up_dir<true>(p); }
      else bubble_up_dir<false>(i);
    }
  }
  template <bool kMin>
  void bubble_up_dir(size_t i) {
    while (i > 2) {
      size_t g = ((i - 1) / 2 - 1) / 2;
      bool mv = kMin ? (v_[i] < v_[g]) : (v_[g] < v_[i]);
      if (!mv) break;
      std::swap(v_[i], v_[g]);
      i = g;
    }
  }
  void trickle_down(size_t i) {
    if (min_level(i)) trickle_dir<true>(i); else trickle_dir<false>(i);
  }
  template <bool kMin>
  void trickle_dir(size_t i) {
    auto better = [&](size_t a, size_t b) { return kMin ? (v_[a] < v_[b]) : (v_[b] < v_[a]); };
    const size_t n = v_.size();
    for (;;) {
      size_t c = 2 * i + 1;
      if (c >= n) return;
      size_t m = c;
      if (c + 1 < n && better(c + 1, m)) m = c + 1;
      for (size_t g = 4 * i + 3; g < n && g <= 4 * i + 6; ++g)
        if (better(g, m)) m = g;
      if (m >= 4 * i + 3) {
        if (!better(m, i)) return;
        std::swap(v_[m], v_[i]);
        size_t p = (m - 1) / 2;
        if (better(p, m)) std::swap(v_[m], v_[p]);
        i = m;
      } else {
        if (better(m, i)) std::swap(v_[m], v_[i]);
        return;
      }
    }
  }
};

// ---------------------------------------------------------------------------
// Visited set: the reference uses std HashSet<usize>::with_capacity(2048)
// (src/index/nsw.rs:143).  Same class of structure (open-addressing hash set,
// 8-byte keys); a cheap multiplicative hash instead of SipHash only makes the
// CPU baseline faster, i.e. the GPU/CPU ratio conservative.
// ---------------------------------------------------------------------------
class VisitedSet {
 public:
  explicit VisitedSet(size_t capacity_hint = 2048) {
    size_t cap = 16;
    while (cap < capacity_hint * 2) cap <<= 1;
    tab_.assign(cap, kEmpty);
  }
  // returns true when newly inserted
  bool insert(uint64_t k) {
    if ((n_ + 1) * 8 > tab_.size() * 7) grow();
    return raw_insert(k);
  }
  bool contains(uint64_t k) const {
    size_t mask = tab_.size() - 1, h = hash(k) & mask;
    for (;;) {
      uint64_t s = tab_[h];
      if (s == kEmpty) return false;
      if (s == k) return true;
      h = (h + 1) & mask;
    }
  }
  size_t size() const { return n_; }

 private:
  static constexpr uint64_t kEmpty = ~0ull;
  std::vector<uint64_t> tab_;
  size_t n_ = 0;
  static size_t hash(uint64_t k) { return (size_t)((k * 0x9E3779B97F4A7C15ull) >> 20); }
  bool raw_insert(uint64_t k) {
    size_t mask = tab_.size() - 1, h = hash(k) & mask;
    for (;;) {
      uint64_t s = tab_[h];
      if (s == kEmpty) { tab_[h] = k; ++n_; return true; }
      if (s == k) return false;
      h = (h + 1) & mask;
    }
  }
  void grow() {
    std::vector<uint64_t> old;
    old.swap(tab_);
    tab_.assign(old.size() * 2, kEmpty);
    n_ = 0;
    for (uint64_t s : old) if (s != kEmpty) raw_insert(s);
  }
};

// ---------------------------------------------------------------------------
// Graph containers.  src/collections/simplegraph.rs:10-122 + Graph trait
// src/collections/mod.rs:13-50.  The reference keeps one std HashSet<Idx> per
// node, whose iteration order is unpinnable; the canonical order
// here is: insertion order inside the builder, file order after loading.
// Set semantics (no duplicate edges) are kept.
// ---------------------------------------------------------------------------
struct IdSpan {
  const Idx* b;
  const Idx* e;
  const Idx* begin() const { return b; }
  const Idx* end() const { return e; }
  size_t size() const { return (size_t)(e - b); }
};

template <class Node>
struct VecGraph {
  std::vector<Node> nodes;
  std::vector<std::vector<Idx>> adj;

  Idx add(const Node& t) {  // simplegraph.rs:82-87
    nodes.push_back(t);
    adj.emplace_back();
    return nodes.size() - 1;
  }
  const Node* get(Idx v) const { return v < nodes.size() ? &nodes[v] : nullptr; }  // :89-91
  bool in_bounds(Idx v, Idx w) const { return v < adj.size() && w < adj.size(); }  // :31-34
  void connect_directed(Idx s, Idx t) {                                            // :36-40
    auto& a = adj[s];
    if (std::find(a.begin(), a.end(), t) == a.end()) a.push_back(t);
  }
  void disconnect_directed(Idx s, Idx t) {  // :42-46
    auto& a = adj[s];
    auto it = std::find(a.begin(), a.end(), t);
    if (it != a.end()) a.erase(it);
  }
  void add_edge(Idx v, Idx w) {  // :93-100
    if (!in_bounds(v, w)) return;
    connect_directed(v, w);
    connect_directed(w, v);
  }
  void remove_edge(Idx v, Idx w) {  // :102-109
    if (!in_bounds(v, w)) return;
    disconnect_directed(v, w);
    disconnect_directed(w, v);
  }
  IdSpan neighborhood(Idx v) const {  // :111-117
    if (v >= adj.size()) return {nullptr, nullptr};
    return {adj[v].data(), adj[v].data() + adj[v].size()};
  }
  size_t size() const { return nodes.size(); }                         // :119-121
  bool is_connected(Idx v, Idx w) const {                              // collections/mod.rs:26-28
    for (Idx i : neighborhood(v)) if (i == w) return true;
    return false;
  }
  size_t degree(Idx v) const { return neighborhood(v).size(); }        // mod.rs:30-32
  void clear_edges(Idx v) {                                            // mod.rs:34-39
    std::vector<Idx> nb(neighborhood(v).begin(), neighborhood(v).end());
    for (Idx w : nb) remove_edge(v, w);
  }
  template <class It>
  void add_neighbors(Idx v, It b, It e) {  // mod.rs:41-49
    for (; b != e; ++b) add_edge(v, *b);
  }
};

// Immutable CSR form used after loading an .idx (neighbour order = file order).
template <class Node>
struct CsrGraph {
  std::vector<Node> nodes;
  std::vector<uint64_t> off;  // size()+1
  std::vector<Idx> ids;       // 8-byte ids as in the reference
  const Node* get(Idx v) const { return v < nodes.size() ? &nodes[v] : nullptr; }
  IdSpan neighborhood(Idx v) const {
    if (v >= nodes.size()) return {nullptr, nullptr};
    return {ids.data() + off[v], ids.data() + off[v + 1]};
  }
  size_t size() const { return nodes.size(); }
};

template <class Node>
CsrGraph<Node> to_csr(const VecGraph<Node>& g) {
  CsrGraph<Node> c;
  c.nodes = g.nodes;
  c.off.resize(g.size() + 1, 0);
  for (size_t i = 0; i < g.size(); ++i) c.off[i + 1] = c.off[i] + g.adj[i].size();
  c.ids.reserve(c.off.back());
  for (auto& a : g.adj) c.ids.insert(c.ids.end(), a.begin(), a.end());
  return c;
}

// ---------------------------------------------------------------------------
// nsw::search  -- src/index/nsw.rs:133-184.
// `dist(node, query)`.
// ---------------------------------------------------------------------------
template <class G, class Q, class DF>
MinMaxHeap<Dist> search(const G& graph, const Q& query, size_t ef, Idx ep, DF dist, Counters* ctr = nullptr) {
  auto* ep_elem = graph.get(ep);  // :140
  if (!ep_elem) throw std::runtime_error("entry point was not in graph");
  Dist d0{(uint64_t)dist(*ep_elem, query), ep};  // :141
  VisitedSet visited(2048);                      // :143
  visited.insert(ep);                            // :144
  MinMaxHeap<Dist> w, cands;                     // :145-146
  w.push(d0);
  cands.push(d0);
  Counters c;
  c.V = 1;
  while (!cands.empty()) {                 // :148
    Dist cur = cands.pop_min();            // :149
    Dist f = w.peek_max();                 // :150
    if (cur.distance > f.distance) break;  // :152-154
    ++c.X;
    for (Idx e : graph.neighborhood(cur.key)) {  // :156
      ++c.E;
      if (!visited.insert(e)) continue;  // :157-161
      Dist f2 = w.peek_max();            // :162
      Dist ed{(uint64_t)dist(*graph.get(e), query), e};  // :164-165
      ++c.V;
      if (ed.distance >= f2.distance && w.size() >= ef) continue;  // :167-169
      cands.push(ed);                                              // :171
      w.push(ed);                                                  // :172
      ++c.A;
      if (w.size() > ef) w.pop_max();  // :174-176
    }
  }
  if (ctr) ctr->add(c);
  return w;  // :183
}

// select_neighbors heuristic -- src/index/nsw.rs:12-33
// `pdist(a,b)` is the point-to-point metric, `point_of(key)` fetches a point.
template <class PointOf, class PD>
std::vector<Dist> select_neighbors(MinMaxHeap<Dist> candidates, size_t m, PointOf point_of, PD pdist) {
  std::vector<Dist> ret;
  while (!candidates.empty()) {  // :19
    Dist e = candidates.pop_min();
    if (ret.size() >= m) break;  // :20-22
    bool all = true;             // :24-26
    for (auto& r : ret)
      if (!((uint64_t)pdist(point_of(e.key), point_of(r.key)) > e.distance)) { all = false; break; }
    if (all) ret.push_back(e);  // :27-29
  }
  return ret;
}

// search_select_neighbors -- src/index/nsw.rs:44-58
template <class G, class Q, class DFq, class PointOf, class PD>
std::vector<Idx> search_select_neighbors(const G& graph, const Q& point, size_t m, size_t ef, Idx ep, DFq dist_q,
                                         PointOf point_of, PD pdist, Counters* ctr = nullptr) {
  auto w = search(graph, point, ef, ep, dist_q, ctr);
  auto sel = select_neighbors(std::move(w), m, point_of, pdist);
  std::vector<Idx> out;
  out.reserve(sel.size());
  for (auto& d : sel) out.push_back(d.key);
  return out;
}

// insert_neighbors -- src/index/nsw.rs:94-131.  `node_dist(a,b)` is the metric
// between two graph nodes.
template <class Node, class ND>
void insert_neighbors(VecGraph<Node>& graph, Idx point_idx, const std::vector<Idx>& neighbors, size_t m_max, ND node_dist) {
  for (Idx e : neighbors) graph.add_edge(point_idx, e);  // :101-103
  for (Idx e : neighbors) {                              // :105
    const Node& e_elem = *graph.get(e);
    std::vector<Idx> e_conn(graph.neighborhood(e).begin(), graph.neighborhood(e).end());  // :107
    if (e_conn.size() <= m_max) continue;                                                  // :109-111
    MinMaxHeap<Dist> cands;                                                                // :113-119
    for (Idx idx : e_conn) cands.push(Dist{(uint64_t)node_dist(*graph.get(idx), e_elem), idx});
    auto sel = select_neighbors(std::move(cands), m_max, [&](Idx k) -> const Node& { return *graph.get(k); }, node_dist);  // :121
    std::vector<Idx> keys;
    for (auto& d : sel) keys.push_back(d.key);  // :123-126
    graph.clear_edges(e);                       // :127
    graph.add_neighbors(e, keys.begin(), keys.end());  // :128
    graph.add_edge(point_idx, e);                      // :129
  }
}

// insert_idx -- src/index/nsw.rs:73-92
template <class Node, class ND>
Idx insert_idx(VecGraph<Node>& graph, Idx point_idx, size_t m, size_t m_max, size_t ef, Idx ep, ND node_dist) {
  const Node point = *graph.get(point_idx);  // copy: the graph is mutated below
  auto nb = search_select_neighbors(graph, point, m, ef, ep, node_dist,
                                    [&](Idx k) -> const Node& { return *graph.get(k); }, node_dist);
  insert_neighbors(graph, point_idx, nb, m_max, node_dist);
  if (nb.empty()) throw std::runtime_error("there should at least be the element we inserted");
  return nb.front();  // :89-91
}

// NSWOptions -- src/index/nsw.rs:186-202
struct NSWOptions {
  size_t ef_construction = 100;
  size_t connections = 16;
  size_t max_connections = 32;
  size_t size = 0;
};

// Level RNG.  The reference uses rand 0.8.5 StdRng seeded with
// threads ^ size ^ efc ^ m ^ M (src/index/hnsw.rs:24-30) and draws an f32 in
// [0,1) (hnsw.rs:37-40).  StdRng (ChaCha12) is not reproduced: the seed mixes
// rayon's thread count, so the stream is machine-dependent anyway.  Same seed
// formula, splitmix64 stream, same 24-bit f32 construction and level formula.
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
  float gen_f32() { return (float)(next() >> 40) * (1.0f / 16777216.0f); }
};
inline size_t level_from_uniform(float u, size_t connections) {  // hnsw.rs:37-40
  return (size_t)(-std::log(u) * (1.0f / std::log((float)connections)));
}

// ---------------------------------------------------------------------------
// NSW -- src/index/nsw.rs:204-352
// ---------------------------------------------------------------------------
template <class P, class PD, template <class> class G = VecGraph>
struct NSW {
  G<P> graph;
  std::optional<Idx> ep;
  PD pdist{};
  size_t size() const { return graph.size(); }
  std::vector<Dist> search(const P& q, size_t k, size_t ef, Counters* ctr = nullptr) const {  // :341-351
    if (!ep) return {};
    auto w = orc::search(graph, q, ef, *ep, pdist, ctr);
    auto v = w.drain_asc();
    if (v.size() > k) v.resize(k);
    return v;
  }
};

template <class P, class PD>
struct NSWBuilder {
  VecGraph<P> graph;
  std::optional<Idx> ep;
  size_t efc, m, M;
  PD pdist{};
  explicit NSWBuilder(const NSWOptions& o) : efc(o.ef_construction), m(o.connections), M(o.max_connections) {}
  void add(const P& point) {  // :290-314
    if (ep) {
      Idx idx = graph.add(point);  // insert_point :60-71
      insert_idx(graph, idx, m, M, efc, *ep, pdist);
    } else {
      Idx e = graph.add(point);
      ep = e;
      insert_idx(graph, e, m, M, efc, e, pdist);
    }
  }
  template <class It>
  void extend(It b, It e) { for (; b != e; ++b) add(*b); }  // :279-285
  // :225-276; `threads` stands in for rayon::current_num_threads()
  void extend_parallel(const P* pts, size_t n, size_t threads) {
    size_t i = 0;
    if (!ep && i < n) add(pts[i++]);
    while (i < n && graph.size() < 50000) add(pts[i++]);  // :235
    const size_t chunk = threads * 32;                    // :237
    while (i < n) {
      size_t c = std::min(chunk, n - i);
      std::vector<Idx> idxs(c);
      for (size_t j = 0; j < c; ++j) idxs[j] = graph.add(pts[i + j]);  // :246-249
      std::vector<std::vector<Idx>> nbs(c);
#pragma omp parallel for schedule(dynamic, 1)
      for (long j = 0; j < (long)c; ++j) {  // :250-265 (frozen graph, read-only)
        const P& p = graph.nodes[idxs[j]];
        nbs[j] = search_select_neighbors(graph, p, m, efc, *ep, pdist,
                                         [&](Idx k) -> const P& { return graph.nodes[k]; }, pdist);
      }
      for (size_t j = 0; j < c; ++j) insert_neighbors(graph, idxs[j], nbs[j], M, pdist);  // :266-274
      i += c;
    }
  }
  NSW<P, PD> build() { return NSW<P, PD>{std::move(graph), ep, pdist}; }  // :316-321
};

// ---------------------------------------------------------------------------
// HNSW -- src/index/hnsw.rs
// ---------------------------------------------------------------------------
template <class P, class PD, template <class> class G = VecGraph>
struct HNSW {  // hnsw.rs:255-261
  using LNode = std::pair<P, Idx>;
  std::vector<G<LNode>> layers;
  G<P> base;
  std::optional<Idx> ep;
  PD pdist{};
  size_t size() const { return base.size(); }  // :274-276

  // hnsw.rs:278-300.  per_layer (optional) receives counters split by layer:
  // index 0 = base, 1.. = layers[0]..
  std::vector<Dist> search(const P& q, size_t k, size_t ef, Counters* ctr = nullptr) const {
    if (!ep) return {};  // :282
    Idx cur = *ep;
    auto ldist = [&](const LNode& n, const P& qq) { return pdist(n.first, qq); };
    for (size_t l = layers.size(); l-- > 0;) {  // :285-293
      auto w = orc::search(layers[l], q, 1, cur, ldist, ctr);
      cur = layers[l].get(w.pop_min().key)->second;
    }
    auto w = orc::search(base, q, ef, cur, pdist, ctr);  // :296
    auto v = w.drain_asc();                              // :297
    if (v.size() > k) v.resize(k);                       // :298
    return v;
  }
};

template <class P, class PD>
struct HNSWBuilder {  // hnsw.rs:8-16
  using LNode = std::pair<P, Idx>;
  std::vector<VecGraph<LNode>> layers;
  VecGraph<P> base;
  std::optional<Idx> ep;
  LevelRng rng;
  size_t efc, m, M;
  PD pdist{};

  // hnsw.rs:19-35; `threads` stands in for rayon::current_num_threads()
  HNSWBuilder(const NSWOptions& o, size_t threads)
      : rng((uint64_t)(threads ^ o.size ^ o.ef_construction ^ o.connections ^ o.max_connections)),
        efc(o.ef_construction), m(o.connections), M(o.max_connections) {}

  size_t random_level() {  // hnsw.rs:37-40 (u == 0 would give an unbounded level: redraw)
    float u;
    do { u = rng.gen_f32(); } while (u == 0.0f);
    return level_from_uniform(u, m);
  }

  struct LDist {
    PD pd;
    size_t operator()(const LNode& a, const LNode& b) const { return pd(a.first, b.first); }
  };

  void add(const P& point) {  // hnsw.rs:183-244
    Idx base_idx = base.add(point);
    size_t level;
    if (ep) level = random_level();
    else { ep = base_idx; level = layers.size(); }
    bool new_ep = false;
    while (layers.size() < level) { layers.emplace_back(); new_ep = true; }  // :193-197
    std::vector<Idx> idxs;  // idxs[l] = index of the point in layers[l]   (:199-205)
    Idx below = base_idx;
    for (size_t l = 0; l < level; ++l) { below = layers[l].add(LNode{point, below}); idxs.push_back(below); }
    if (new_ep) ep = idxs.back();  // :207-210
    Idx cur = *ep;
    auto ldq = [&](const LNode& n, const P& q) { return pdist(n.first, q); };
    for (size_t l = layers.size(); l-- > level;) {  // :215-219
      auto w = orc::search(layers[l], point, 1, cur, ldq);
      cur = layers[l].get(w.peek_min().key)->second;
    }
    for (size_t l = level; l-- > 0;)  // :222-232 (ep = nearest selected index in *this* layer -- reference quirk)
      cur = insert_idx(layers[l], idxs[l], m, M, efc, cur, LDist{pdist});
    insert_idx(base, base_idx, m, M, efc, cur, pdist);  // :235-243
  }
  template <class It>
  void extend(It b, It e) { for (; b != e; ++b) add(*b); }  // :172-178

  // One chunk of extend_parallel (hnsw.rs:58-168) given its level.
  void parallel_chunk(const P* pts, size_t c, size_t level) {
    bool new_ep = false;
    while (layers.size() < level) { layers.emplace_back(); new_ep = true; }  // :67-71
    std::vector<std::vector<Idx>> cidx(c);                                   // :74-87
    for (size_t j = 0; j < c; ++j) {
      Idx below = base.add(pts[j]);
      cidx[j].push_back(below);
      for (size_t l = 0; l < level; ++l) { below = layers[l].add(LNode{pts[j], below}); cidx[j].push_back(below); }
    }
    if (new_ep) ep = cidx[0].back();  // :89-92
    std::vector<Idx> eps(c);
    auto ldq = [&](const LNode& n, const P& q) { return pdist(n.first, q); };
#pragma omp parallel for schedule(dynamic, 1)
    for (long j = 0; j < (long)c; ++j) {  // :94-108
      Idx cur = *ep;
      for (size_t l = layers.size(); l-- > level;) {
        auto w = orc::search(layers[l], pts[j], 1, cur, ldq);
        cur = layers[l].get(w.peek_min().key)->second;
      }
      eps[j] = cur;
    }
    std::vector<std::vector<Idx>> nbs(c);
    for (size_t l = level; l-- > 0;) {  // :111-139 (same ep for every layer -- reference quirk)
      auto& layer = layers[l];
#pragma omp parallel for schedule(dynamic, 1)
      for (long j = 0; j < (long)c; ++j) {
        LNode probe{pts[j], 0};
        nbs[j] = search_select_neighbors(layer, probe, m, efc, eps[j], LDist{pdist},
                                         [&](Idx k) -> const LNode& { return layer.nodes[k]; }, LDist{pdist});
      }
      for (size_t j = 0; j < c; ++j) insert_neighbors(layer, cidx[j][l + 1], nbs[j], M, LDist{pdist});
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (long j = 0; j < (long)c; ++j)  // :142-156
      nbs[j] = search_select_neighbors(base, pts[j], m, efc, eps[j], pdist,
                                       [&](Idx k) -> const P& { return base.nodes[k]; }, pdist);
    for (size_t j = 0; j < c; ++j) insert_neighbors(base, cidx[j][0], nbs[j], M, pdist);  // :159-167
  }

  // hnsw.rs:44-169
  void extend_parallel(const P* pts, size_t n, size_t threads) {
    size_t i = 0;
    if (!ep && i < n) add(pts[i++]);                     // :47-51
    while (i < n && base.size() < 50000) add(pts[i++]);  // :54
    const size_t chunk = threads * 32;                   // :56
    while (i < n) {
      size_t c = std::min(chunk, n - i);
      size_t level = random_level();  // :65 -- ONE level per chunk
      parallel_chunk(pts + i, c, level);
      i += c;
    }
  }

  HNSW<P, PD> build() { return HNSW<P, PD>{std::move(layers), std::move(base), ep, pdist}; }  // :246-252
};

// ---------------------------------------------------------------------------
// MinK / Bruteforce -- src/collections/mod.rs:52-79, src/index/bruteforce.rs:37-47.
// Restates Rust std BinaryHeap (max-heap): collect() == heapify; PeekMut
// assignment == write root + sift_down(0).  Returns heap.into_vec() (heap
// order, unsorted), which is what the library API exposes; the CLI sorts.
// ---------------------------------------------------------------------------
template <class T>
struct RustBinaryHeap {
  std::vector<T> d;
  void sift_down_range(size_t pos, size_t end) {
    T elt = d[pos];
    size_t hole = pos, child = 2 * pos + 1;
    while (end >= 2 && child <= end - 2) {
      if (d[child] <= d[child + 1]) child += 1;
      if (elt >= d[child]) { d[hole] = elt; return; }
      d[hole] = d[child];
      hole = child;
      child = 2 * hole + 1;
    }
    if (child == end - 1 && elt < d[child]) { d[hole] = d[child]; hole = child; }
    d[hole] = elt;
  }
  void rebuild() {
    size_t n = d.size() / 2;
    while (n > 0) { --n; sift_down_range(n, d.size()); }
  }
};

template <class T, class It>
std::vector<T> min_k(It b, It e, size_t k) {  // collections/mod.rs:53-77
  if (k == 0) return {};
  RustBinaryHeap<T> heap;
  for (; b != e && heap.d.size() < k; ++b) heap.d.push_back(*b);  // :63
  heap.rebuild();
  for (; b != e; ++b) {  // :65-73
    T i = *b;
    if (heap.d[0] > i) { heap.d[0] = i; heap.sift_down_range(0, heap.d.size()); }
  }
  return heap.d;  // :75
}

template <class P, class PD>
struct Bruteforce {  // bruteforce.rs:8-47
  std::vector<P> points;
  PD pdist{};
  size_t size() const { return points.size(); }
  std::vector<Dist> search(const P& q, size_t k) const {
    struct GenIt {
      const Bruteforce* bf; const P* q; size_t i;
      Dist operator*() const { return Dist{(uint64_t)bf->pdist(*q, bf->points[i]), i}; }
      GenIt& operator++() { ++i; return *this; }
      bool operator!=(const GenIt& o) const { return i != o.i; }
    };
    return min_k<Dist>(GenIt{this, &q, 0}, GenIt{this, &q, points.size()}, k);
  }
};

// Indexes::search padding -- cli/src/main.rs:457-477.  Throws when empty
// (the reference panics on `res.first().unwrap()`).
inline void pad_like_cli(std::vector<Dist>& res, size_t k) {
  if (res.size() < k) {
    if (res.empty()) throw std::runtime_error("search returned no elements (reference panics here)");
    Dist fst = res.front();
    res.resize(k, fst);
  }
}

}  // namespace orc
