// idx_format.cpp -- the reference's `.idx` wire format (bincode 1.3.3 legacy
// `serialize_into` / `deserialize_from` of IndexFile<Sketch>), replaces
// read_index / write_index (cli/src/main.rs:224-248) and the serde derives at
// hnsw.rs:255-261, nsw.rs:324-328, bruteforce.rs:8-11, simplegraph.rs:8-14,
// sketch.rs:7-10, main.rs:311-320,381-389,441-446,480-484.
//
// Layout: little-endian fixed-width ints, usize = u64,
// sequence / string lengths = u64 prefix, enum tag = u32, Option = u8 tag,
// [u64;16] and tuples without prefix, struct fields in declaration order.
//
// Single streaming pass; 8-byte neighbour ids are narrowed to u32 on the fly
// (node counts must stay below 2^32 - 1) in file order, which is the canonical
// neighbour order of this engine.
#include <cstdio>
#include <cstring>
#include <memory>

#include "hb_internal.h"

namespace hb {
namespace {

struct FileCloser {
  void operator()(FILE* f) const { if (f) fclose(f); }
};
using FilePtr = std::unique_ptr<FILE, FileCloser>;

class In {
 public:
  explicit In(const char* path) : f_(fopen(path, "rb")), buf_(1 << 22) {
    HB_REQUIRE(f_ != nullptr, HB_ERR_IO, std::string("cannot open index file ") + path);
    setvbuf(f_.get(), buf_.data(), _IOFBF, buf_.size());
  }
  void bytes(void* dst, size_t n) {
    if (n == 0) return;
    HB_REQUIRE(fread(dst, 1, n, f_.get()) == n, HB_ERR_FORMAT, "index file ends early");
  }
  template <class T>
  T get() {
    T v;
    bytes(&v, sizeof v);
    return v;
  }
  std::string text() {
    uint64_t n = get<uint64_t>();
    HB_REQUIRE(n <= (1u << 20), HB_ERR_FORMAT, "string length in index file is implausible");
    std::string s((size_t)n, '\0');
    bytes(s.data(), (size_t)n);
    return s;
  }
  bool at_end() {
    int c = fgetc(f_.get());
    if (c == EOF) return true;
    ungetc(c, f_.get());
    return false;
  }

 private:
  FilePtr f_;
  std::vector<char> buf_;
};

class Out {
 public:
  explicit Out(const char* path) : f_(fopen(path, "wb")), buf_(1 << 22) {
    HB_REQUIRE(f_ != nullptr, HB_ERR_IO, std::string("cannot create index file ") + path);
    setvbuf(f_.get(), buf_.data(), _IOFBF, buf_.size());
  }
  void bytes(const void* src, size_t n) {
    if (n == 0) return;
    HB_REQUIRE(fwrite(src, 1, n, f_.get()) == n, HB_ERR_IO, "short write to index file");
  }
  template <class T>
  void put(T v) { bytes(&v, sizeof v); }
  void text(const std::string& s) {
    put<uint64_t>(s.size());
    bytes(s.data(), s.size());
  }
  void finish() {
    HB_REQUIRE(fflush(f_.get()) == 0, HB_ERR_IO, "flush of index file failed");
  }

 private:
  FilePtr f_;
  std::vector<char> buf_;
};

constexpr uint64_t kMaxNodes = 0xFFFFFFFEull;  // 0xFFFFFFFF is the empty marker of the visited table

// Vec<HashSet<usize>> followed by the always-empty `empty` set (simplegraph.rs:12-13)
void read_adjacency(In& in, HostGraph& g) {
  uint64_t lists = in.get<uint64_t>();
  HB_REQUIRE(lists == g.n, HB_ERR_FORMAT, "adj_lists length differs from nodes length");
  g.off.resize(g.n + 1);
  g.off[0] = 0;
  std::vector<uint64_t> wide;
  for (uint64_t v = 0; v < g.n; ++v) {
    uint64_t deg = in.get<uint64_t>();
    HB_REQUIRE(deg <= g.n, HB_ERR_FORMAT, "degree larger than the graph");
    wide.resize((size_t)deg);
    in.bytes(wide.data(), (size_t)deg * 8);
    size_t at = g.ids.size();
    g.ids.resize(at + (size_t)deg);
    for (uint64_t j = 0; j < deg; ++j) {
      HB_REQUIRE(wide[j] < g.n, HB_ERR_FORMAT, "neighbour id out of range");
      g.ids[at + j] = (uint32_t)wide[j];
    }
    g.off[v + 1] = g.off[v] + deg;
  }
  uint64_t empty_len = in.get<uint64_t>();
  for (uint64_t j = 0; j < empty_len; ++j) (void)in.get<uint64_t>();
}

// SimpleGraph<Sketch>: nodes are bare 128-byte sketches
void read_plain_graph(In& in, HostGraph& g) {
  g.n = in.get<uint64_t>();
  HB_REQUIRE(g.n <= kMaxNodes, HB_ERR_UNSUPPORTED, "more than 2^32-2 nodes in one graph");
  g.sk.resize((size_t)g.n * 16);
  in.bytes(g.sk.data(), (size_t)g.n * 128);
  read_adjacency(in, g);
}

// SimpleGraph<(Sketch, usize)>: 136 bytes per node, the usize indexes the graph below
void read_layer_graph(In& in, HostGraph& g) {
  g.n = in.get<uint64_t>();
  HB_REQUIRE(g.n <= kMaxNodes, HB_ERR_UNSUPPORTED, "more than 2^32-2 nodes in one graph");
  g.sk.resize((size_t)g.n * 16);
  g.down.resize((size_t)g.n);
  for (uint64_t v = 0; v < g.n; ++v) {
    in.bytes(&g.sk[(size_t)v * 16], 128);
    g.down[(size_t)v] = in.get<uint64_t>();
  }
  read_adjacency(in, g);
}

void write_adjacency(Out& out, const HostGraph& g) {
  out.put<uint64_t>(g.n);
  std::vector<uint64_t> wide;
  for (uint64_t v = 0; v < g.n; ++v) {
    uint64_t b = g.off[v], e = g.off[v + 1];
    out.put<uint64_t>(e - b);
    wide.resize((size_t)(e - b));
    for (uint64_t j = b; j < e; ++j) wide[(size_t)(j - b)] = g.ids[(size_t)j];
    out.bytes(wide.data(), wide.size() * 8);
  }
  out.put<uint64_t>(0);  // `empty`
}

int64_t read_option_usize(In& in) {
  uint8_t tag = in.get<uint8_t>();
  if (tag == 0) return -1;
  HB_REQUIRE(tag == 1, HB_ERR_FORMAT, "bad Option tag for the entry point");
  uint64_t v = in.get<uint64_t>();
  HB_REQUIRE(v <= kMaxNodes, HB_ERR_FORMAT, "entry point out of range");
  return (int64_t)v;
}

void write_option_usize(Out& out, int64_t v) {
  if (v < 0) {
    out.put<uint8_t>(0);
  } else {
    out.put<uint8_t>(1);
    out.put<uint64_t>((uint64_t)v);
  }
}

}  // namespace

void read_idx_file(const char* path, Attrs& a, int& algo, std::vector<HostGraph>& graphs, int64_t& ep) {
  In in(path);
  a.format_size = in.get<uint8_t>() != 0;
  a.data = in.text();
  a.size = in.get<uint64_t>();
  a.algo = in.get<uint32_t>();
  a.buildtime = in.get<double>();
  a.querytime = in.get<double>();
  a.params = in.text();
  uint32_t tag = in.get<uint32_t>();
  graphs.clear();
  ep = -1;
  switch (tag) {
    case HB_ALGO_BRUTEFORCE: {  // Bruteforce<Sketch>{points}
      graphs.resize(1);
      HostGraph& g = graphs[0];
      g.n = in.get<uint64_t>();
      HB_REQUIRE(g.n <= kMaxNodes, HB_ERR_UNSUPPORTED, "more than 2^32-2 points");
      g.sk.resize((size_t)g.n * 16);
      in.bytes(g.sk.data(), (size_t)g.n * 128);
      g.off.assign((size_t)g.n + 1, 0);
      break;
    }
    case HB_ALGO_NSW: {  // NSW<Sketch>{graph, ep}
      graphs.resize(1);
      read_plain_graph(in, graphs[0]);
      ep = read_option_usize(in);
      break;
    }
    case HB_ALGO_HNSW: {  // HNSW<Sketch>{layers, base, ep}
      uint64_t n_layers = in.get<uint64_t>();
      HB_REQUIRE(n_layers < 64, HB_ERR_FORMAT, "implausible number of layers");
      graphs.resize((size_t)n_layers + 1);
      for (uint64_t l = 0; l < n_layers; ++l) read_layer_graph(in, graphs[(size_t)l + 1]);
      read_plain_graph(in, graphs[0]);
      ep = read_option_usize(in);
      break;
    }
    default:
      throw Error(HB_ERR_FORMAT, "unknown Indexes variant tag");
  }
  algo = (int)tag;
  HB_REQUIRE(in.at_end(), HB_ERR_FORMAT, "trailing bytes after the index");
}

void write_idx_file(const char* path, const Attrs& a, int algo, const std::vector<HostGraph>& graphs, int64_t ep) {
  HB_REQUIRE(!graphs.empty(), HB_ERR_INVALID, "no host copy of the index to write");
  Out out(path);
  out.put<uint8_t>(a.format_size ? 1 : 0);
  out.text(a.data);
  out.put<uint64_t>(a.size);
  out.put<uint32_t>(a.algo);
  out.put<double>(a.buildtime);
  out.put<double>(a.querytime);
  out.text(a.params);
  out.put<uint32_t>((uint32_t)algo);
  auto plain = [&](const HostGraph& g) {
    out.put<uint64_t>(g.n);
    out.bytes(g.sk.data(), (size_t)g.n * 128);
    write_adjacency(out, g);
  };
  if (algo == HB_ALGO_BRUTEFORCE) {
    out.put<uint64_t>(graphs[0].n);
    out.bytes(graphs[0].sk.data(), (size_t)graphs[0].n * 128);
  } else if (algo == HB_ALGO_NSW) {
    plain(graphs[0]);
    write_option_usize(out, ep);
  } else {
    out.put<uint64_t>(graphs.size() - 1);
    for (size_t l = 1; l < graphs.size(); ++l) {
      const HostGraph& g = graphs[l];
      out.put<uint64_t>(g.n);
      for (uint64_t v = 0; v < g.n; ++v) {
        out.bytes(&g.sk[(size_t)v * 16], 128);
        out.put<uint64_t>(g.down[(size_t)v]);
      }
      write_adjacency(out, g);
    }
    plain(graphs[0]);
    write_option_usize(out, ep);
  }
  out.finish();
}

}  // namespace hb
