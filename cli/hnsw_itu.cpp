// hnsw-itu (C++ stand-in) -- the reference CLI's subcommands, driving libhnswb200.so through the C ABI.
//
// The reference CLI is Rust + clap + libhdf5 (cli/src/main.rs); this image has neither a Rust toolchain nor libhdf5, so
// this binary keeps the subcommand names, flags, defaults and result conventions:
//
//   hnsw-itu query        -d DATA -Q QUERIES [-o result.h5] [-i IDX] [-k 10] [-e 96] [-c 96] [-m 24] [-M 256]
//                         [-a hnsw|nsw|bruteforce] [-s] [-S]                                       (main.rs:486-571)
//   hnsw-itu index        -d DATA [-o index.idx] [-c 96] [-m 24] [-M 256] [-b START] [-l LEN] [-a ALGO] [-S]  (main.rs:573-632)
//   hnsw-itu query-index  -i IDX -Q QUERIES [-o result.h5] [-k 10] [-e 96] [-s] [-S]               (main.rs:634-681)
//   hnsw-itu ground-truth -d DATA -Q QUERIES [-o groundtruth.h5] [-b START] [-l LEN] [-k 100] [-s true|false]  (main.rs:683-754)
//   hnsw-itu inspect      IDX                                                                       (main.rs:756-821)
//   extra (not in the reference): --devices 0,1,..  spread query-index (replicated index, batch split) and ground-truth
//                                 (rows sharded, all-gather + merge) over several GPUs of the host (hb_multi_*)
//
// Files.  DATA / QUERIES: an HDF5 file with the 2-D u64 dataset "hamming" (cli/src/dataset.rs:15-79) read by cli/h5min.hpp,
// or raw little-endian u64 (16 per row, `.u64`), or a NumPy `.npy` of shape (n, 16) `<u8`.  Results: when the output
// name ends in .h5 / .hdf5, an HDF5 file with the dataset "knns" (nq x k, keys + 1 as the reference writes them,
// main.rs:277,745), "dists" for ground-truth, and for query / query-index the six root attributes of write_result
// (main.rs:287-306); otherwise the raw side files `<out>.knns.u64`, `<out>.dists.u64`, `<out>.attrs.json`.
// Errors print the library's message and exit non-zero (anyhow in the reference, main.rs:31,63).
#include <algorithm>
#include <chrono>
#include <cinttypes>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <queue>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/hnsw_b200.h"
#include "h5min.hpp"

namespace {

[[noreturn]] void die(const std::string& m) {
  std::fprintf(stderr, "hnsw-itu: %s\n", m.c_str());
  std::exit(1);
}
void check(int rc, const char* what) {
  if (rc != HB_OK) die(std::string(what) + ": " + hb_last_error());
}
bool ends_with(const std::string& s, const std::string& suf) {
  return s.size() >= suf.size() && !s.compare(s.size() - suf.size(), suf.size(), suf);
}
bool is_h5_name(const std::string& p) { return ends_with(p, ".h5") || ends_with(p, ".hdf5"); }

// ---- sketch files ------------------------------------------------------------------------------------------------
// rows [skip, skip + take) of the file (BufferedDataset + .skip().take(), main.rs:118-140); rows = rows in the file
std::vector<uint64_t> read_sketches(const std::string& path, uint64_t& rows, uint64_t skip = 0, uint64_t take = ~0ull) {
  std::ifstream f(path, std::ios::binary);
  if (!f) die("cannot open " + path);
  char magic[8] = {0};
  f.read(magic, 8);
  static const unsigned char h5sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
  if (is_h5_name(path) || !std::memcmp(magic, h5sig, 8)) {
    try {
      h5min::Reader r(path);
      const h5min::DatasetInfo d = r.dataset("hamming");
      if (d.dims.size() != 2 || d.dims[1] != 16) die(path + ": dataset \"hamming\" must have shape (n, 16)");
      rows = d.dims[0];
      skip = std::min(skip, rows);
      take = std::min(take, rows - skip);
      std::vector<uint64_t> v(take * 16);
      constexpr uint64_t kSlice = 50000;  // dataset.rs:117: the reference reads slices of 50 000 rows
      for (uint64_t at = 0; at < take; at += kSlice) r.read_rows_u64(d, skip + at, std::min(kSlice, take - at), v.data() + at * 16);
      return v;
    } catch (const h5min::Error& e) {
      die(e.what());
    }
  }
  f.clear();
  f.seekg(0, std::ios::end);
  const uint64_t bytes = (uint64_t)f.tellg();
  uint64_t hdr_bytes = 0;
  if (bytes >= 10 && std::memcmp(magic, "\x93NUMPY", 6) == 0) {
    f.seekg(6);
    unsigned char ver[2], l2[2], l4[4];
    f.read(reinterpret_cast<char*>(ver), 2);
    uint32_t hlen;
    if (ver[0] == 1) {
      f.read(reinterpret_cast<char*>(l2), 2);
      hlen = l2[0] | (l2[1] << 8);
      hdr_bytes = 10 + hlen;
    } else {
      f.read(reinterpret_cast<char*>(l4), 4);
      hlen = l4[0] | (l4[1] << 8) | (l4[2] << 16) | ((uint32_t)l4[3] << 24);
      hdr_bytes = 12 + hlen;
    }
    std::string hdr(hlen, '\0');
    f.read(hdr.data(), hlen);
    if (hdr.find("<u8") == std::string::npos && hdr.find("<i8") == std::string::npos) die(path + ": .npy must hold little-endian 64-bit integers");
    if (hdr.find("'fortran_order': True") != std::string::npos) die(path + ": .npy must be C-ordered");
    if (hdr.find(", 16)") == std::string::npos && hdr.find(",16)") == std::string::npos) die(path + ": .npy shape must be (n, 16)");
  }
  if ((bytes - hdr_bytes) % 128 != 0) die(path + ": size is not a multiple of 128 bytes (16 x u64 per sketch)");
  rows = (bytes - hdr_bytes) / 128;
  skip = std::min(skip, rows);
  take = std::min(take, rows - skip);
  std::vector<uint64_t> v(take * 16);
  f.seekg((std::streamoff)(hdr_bytes + skip * 128));
  f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(take * 128));
  if (!f) die("short read on " + path);
  return v;
}

void write_u64(const std::string& path, const std::vector<uint64_t>& v) {
  std::ofstream f(path, std::ios::binary);
  if (!f) die("cannot write " + path);
  f.write(reinterpret_cast<const char*>(v.data()), (std::streamsize)(v.size() * 8));
}

// ---- flags ----------------------------------------------------------------------------------------------------------
struct Args {
  std::map<std::string, std::string> kv;
  std::vector<std::string> positional;
  bool has(const std::string& k) const { return kv.count(k) != 0; }
  std::string str(const std::string& k, const std::string& dflt = "") const { return has(k) ? kv.at(k) : dflt; }
  uint64_t num(const std::string& k, uint64_t dflt) const { return has(k) ? std::strtoull(kv.at(k).c_str(), nullptr, 10) : dflt; }
  std::string need(const std::string& k) const {
    if (!has(k)) die("missing required option --" + k);
    return kv.at(k);
  }
};

// long name for every short flag of the subcommands (main.rs:486-549,573-602,634-664,683-712)
const std::map<std::string, std::string> kShort = {
    {"-i", "indexfile"}, {"-Q", "queryfile"}, {"-o", "outfile"}, {"-k", "k"}, {"-e", "ef"}, {"-s", "sort"}, {"-S", "single-threaded"},
    {"-d", "datafile"},  {"-b", "start"},     {"-l", "len"},     {"-c", "ef-construction"}, {"-m", "connections"}, {"-M", "max-connections"},
    {"-a", "algorithm"}};

// `sort_takes_value`: ground-truth declares `--sort <bool>` (default true, main.rs:710-712), the others a plain flag
Args parse(int argc, char** argv, int from, bool sort_takes_value) {
  Args a;
  for (int i = from; i < argc; ++i) {
    std::string t = argv[i], name;
    if (t.rfind("--", 0) == 0) name = t.substr(2);
    else if (kShort.count(t)) name = kShort.at(t);
    else if (!t.empty() && t[0] != '-') {
      a.positional.push_back(t);
      continue;
    } else die("unknown option " + t);
    const bool next_is_value = i + 1 < argc && argv[i + 1][0] != '-';
    const bool is_bool = name == "single-threaded" || (name == "sort" && !(sort_takes_value && next_is_value));
    if (is_bool) {
      a.kv[name] = "true";
    } else {
      if (!next_is_value) die("option --" + name + " needs a value");
      a.kv[name] = argv[++i];
    }
  }
  return a;
}

std::vector<int> device_list(const Args& a) {
  std::vector<int> d;
  std::string s = a.str("devices", "0");
  for (size_t p = 0; p <= s.size();) {
    const size_t q = std::min(s.find(',', p), s.size());
    if (q > p) d.push_back(std::atoi(s.substr(p, q - p).c_str()));
    p = q + 1;
  }
  if (d.empty()) d.push_back(0);
  return d;
}

// format_size_string (main.rs:250-259)
std::string size_string(uint64_t n, bool format) {
  if (format) {
    if (n >= 90000 && n <= 110000) return "100K";
    if (n >= 270000 && n <= 330000) return "300K";
    if (n >= 9000000 && n <= 11000000) return "10M";
    if (n >= 27000000 && n <= 33000000) return "30M";
    if (n >= 90000000 && n <= 110000000) return "100M";
  }
  return std::to_string(n);
}

std::string json_escape(const std::string& s) {
  std::string o;
  for (char c : s) {
    if (c == '"' || c == '\\') o += '\\';
    o += c;
  }
  return o;
}

const char* algo_name(int a) { return a == HB_ALGO_BRUTEFORCE ? "Bruteforce" : a == HB_ALGO_NSW ? "Nsw" : "Hnsw"; }  // {:?} of Algorithm
int algo_from(const std::string& s) {
  if (s == "hnsw") return HB_ALGO_HNSW;
  if (s == "nsw") return HB_ALGO_NSW;
  if (s == "bruteforce") return HB_ALGO_BRUTEFORCE;
  die("unknown --algorithm " + s + " (bruteforce, nsw, hnsw)");
}

// ResultAttrs (main.rs:311-320) as the six attributes write_result attaches (main.rs:287-306)
struct ResultAttrs {
  bool present = false;
  std::string data = "hamming", size, algo, params;
  double buildtime = 0, querytime = 0;
};

// write_result (main.rs:261-309) / the output loop of ground-truth (main.rs:733-750): keys are written + 1
void write_results(const std::string& out, uint64_t nq, uint32_t k, std::vector<uint64_t> keys, const std::vector<uint64_t>* dists,
                   const ResultAttrs& at) {
  for (auto& x : keys) x += 1;  // main.rs:277,745
  if (is_h5_name(out)) {
    std::vector<h5min::OutDataset> ds{{"knns", nq, k, keys.data()}};
    if (dists) ds.push_back({"dists", nq, k, dists->data()});
    std::vector<std::pair<std::string, h5min::Attr>> attrs;
    if (at.present) {
      auto s = [](const std::string& v) { h5min::Attr a; a.kind = h5min::Attr::STR; a.s = v; return a; };
      auto f = [](double v) { h5min::Attr a; a.f = v; return a; };
      attrs = {{"data", s(at.data)}, {"size", s(at.size)}, {"algo", s(at.algo)}, {"buildtime", f(at.buildtime)},
               {"querytime", f(at.querytime)}, {"params", s(at.params)}};
    }
    try {
      h5min::Writer::write(out, ds, attrs);
    } catch (const h5min::Error& e) {
      die(e.what());
    }
    return;
  }
  write_u64(out + ".knns.u64", keys);
  if (dists) write_u64(out + ".dists.u64", *dists);
  if (at.present) {
    std::ofstream f(out + ".attrs.json");
    if (!f) die("cannot write " + out + ".attrs.json");
    char nums[128];
    std::snprintf(nums, sizeof nums, "\"buildtime\": %.9g, \"querytime\": %.9g", at.buildtime, at.querytime);
    f << "{\"data\": \"" << json_escape(at.data) << "\", \"size\": \"" << at.size << "\", \"algo\": \"" << at.algo << "\", " << nums
      << ", \"params\": \"" << json_escape(at.params) << "\"}\n";
  }
}

// ---- build_index (main.rs:111-179) -------------------------------------------------------------------------------------
struct Built {
  hb_index* idx = nullptr;
  std::vector<uint64_t> data;  // kept for Bruteforce handles and for `query`
};

Built build_index(const Args& a, bool sub_range_flags) {
  const std::string path = a.need("datafile");
  const int algo = algo_from(a.str("algorithm", "hnsw"));
  const bool has_start = sub_range_flags && a.has("start"), has_len = sub_range_flags && a.has("len");
  const uint64_t skip = has_start ? a.num("start", 0) : 0, take = has_len ? a.num("len", 0) : ~0ull;
  uint64_t rows = 0;
  Built b;
  b.data = read_sketches(path, rows, skip, take);
  const uint64_t size = b.data.size() / 16;
  if (has_len && take != size)  // main.rs:125-133
    std::fprintf(stderr, "warn: Dataset range will be smaller than specified `len` size=%" PRIu64 " len=%" PRIu64 "\n", size, take);
  hb_build_params bp{};
  bp.ef_construction = (uint32_t)a.num("ef-construction", 96);
  bp.connections = (uint32_t)a.num("connections", 24);
  bp.max_connections = (uint32_t)a.num("max-connections", 256);
  bp.rayon_threads = 16;
  // -S: `builder.extend(...)` adds every point one at a time; otherwise extend_parallel: 50 000 sequential, then chunks
  bp.serial_prefix = a.has("single-threaded") ? (uint32_t)std::min<uint64_t>(size, 0xFFFFFFFFu) : 0;
  bp.nsw = algo == HB_ALGO_NSW;
  std::fprintf(stderr, "Building index size=%" PRIu64 " algorithm=%s single_threaded=%s\n", size, algo_name(algo), a.has("single-threaded") ? "true" : "false");
  const auto t0 = std::chrono::steady_clock::now();
  if (algo == HB_ALGO_BRUTEFORCE) {
    const uint64_t n = size;
    const uint64_t* sk = b.data.data();
    check(hb_index_from_parts(HB_ALGO_BRUTEFORCE, 1, &n, &sk, nullptr, nullptr, nullptr, 8, -1, 0, 1, &b.idx), "build");
  } else {
    check(hb_build_hnsw(b.data.data(), size, &bp, 0, 1, &b.idx), "build");
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::fprintf(stderr, "Total build time: %.3f s, per element: %.3f us\n", secs, size ? secs / size * 1e6 : 0.0);
  char params[128];  // main.rs:171-174
  std::snprintf(params, sizeof params, "index=(efc=%u,m=%u,M=%u)", bp.ef_construction, bp.connections, bp.max_connections);
  check(hb_index_set_attrs(b.idx, secs, params), "attrs");
  check(hb_index_set_format_size(b.idx, !(has_start || has_len)), "attrs");  // main.rs:122
  return b;
}

// ---- query_index (main.rs:181-222) + write_result --------------------------------------------------------------------------
void query_and_write(const Args& a, hb_index* idx) {
  const uint32_t k = (uint32_t)a.num("k", 10), ef = (uint32_t)a.num("ef", 96);
  if (k > ef) std::fprintf(stderr, "error: `k` (%u) is greater than `ef` (%u), this can have adverse effects\n", k, ef);  // main.rs:190-195
  uint64_t nq = 0;
  const std::vector<uint64_t> q = read_sketches(a.need("queryfile"), nq);
  nq = q.size() / 16;
  std::vector<uint64_t> keys(nq * k), dists(nq * k);
  std::vector<uint32_t> found(nq);
  const std::vector<int> devs = device_list(a);
  hb_multi* multi = nullptr;
  if (devs.size() > 1) {  // replicated index, batch split contiguously over the devices
    check(hb_multi_create(devs.data(), (int)devs.size(), &multi), "multi");
    check(hb_multi_replicate(multi, idx), "replicate");
  }
  const auto t0 = std::chrono::steady_clock::now();
  if (multi) {
    check(hb_multi_knns(multi, q.data(), nq, k, ef, HB_PAD_LIKE_CLI, keys.data(), dists.data(), found.data()), "knns");
  } else if (a.has("single-threaded")) {  // main.rs:203-207: one search per query
    for (uint64_t i = 0; i < nq; ++i)
      check(hb_search(idx, q.data() + i * 16, k, ef, HB_PAD_LIKE_CLI, keys.data() + i * k, dists.data() + i * k, &found[i]), "search");
  } else {
    check(hb_knns(idx, q.data(), nq, k, ef, HB_PAD_LIKE_CLI, keys.data(), dists.data(), found.data(), nullptr), "knns");
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::fprintf(stderr, "Total query time: %.6f s, per query: %.3f us\n", secs, nq ? secs / nq * 1e6 : 0.0);
  for (uint64_t i = 0; i < nq; ++i)
    if (found[i] < k) std::fprintf(stderr, "warn: query %" PRIu64 " found %u < k results; padded with the first\n", i, found[i]);  // main.rs:467-474
  // results are ascending (distance, key) already, so -s/--sort (main.rs:273-275) changes nothing except for padded rows,
  // where the reference's sort moves the repeated first element to the front -- which is where it already is
  ResultAttrs at;
  at.present = true;
  int fmt = 1;
  uint64_t size = 0;
  double qt = 0;
  char data[256], params[1024];
  check(hb_index_get_attrs(idx, &fmt, &size, &at.buildtime, &qt, data, sizeof data, params, sizeof params), "attrs");
  hb_index_info info{};
  check(hb_index_get_info(idx, &info), "info");
  at.data = data;
  at.size = size_string(size, fmt != 0);
  at.algo = algo_name(info.algo);
  at.querytime = secs;
  at.params = std::string(params) + ",query=(ef=" + std::to_string(ef) + ")";  // main.rs:218-219
  // the reference writes only `knns` (main.rs:261-309); the raw side-file form also keeps the distances
  const std::string out = a.str("outfile", "result.h5");
  write_results(out, nq, k, std::move(keys), is_h5_name(out) ? nullptr : &dists, at);
  if (multi) hb_multi_free(multi);
}

int cmd_query(const Args& a) {  // main.rs:550-571
  Built b = build_index(a, false);
  if (a.has("indexfile")) check(hb_index_save_idx(b.idx, a.str("indexfile").c_str()), "write_index");
  query_and_write(a, b.idx);
  hb_index_free(b.idx);
  return 0;
}

int cmd_index(const Args& a) {  // main.rs:624-631
  Built b = build_index(a, true);
  const std::string out = a.str("outfile", "index.idx");
  check(hb_index_save_idx(b.idx, out.c_str()), "write_index");
  std::fprintf(stderr, "wrote %s\n", out.c_str());
  hb_index_free(b.idx);
  return 0;
}

int cmd_query_index(const Args& a) {  // main.rs:666-681
  hb_index* idx = nullptr;
  check(hb_index_load_idx(a.need("indexfile").c_str(), device_list(a)[0], 0, &idx), "read_index");
  query_and_write(a, idx);
  hb_index_free(idx);
  return 0;
}

int cmd_ground_truth(const Args& a) {  // main.rs:715-754
  const uint32_t k = (uint32_t)a.num("k", 100);
  const bool has_len = a.has("len");
  const uint64_t skip = a.num("start", 0), take = has_len ? a.num("len", 0) : ~0ull;
  uint64_t rows = 0, nq = 0;
  const std::vector<uint64_t> data = read_sketches(a.need("datafile"), rows, skip, take);
  const uint64_t len = data.size() / 16;
  if (has_len && take != len) std::fprintf(stderr, "warn: Dataset range will be smaller than specified `len` size=%" PRIu64 " len=%" PRIu64 "\n", len, take);
  const std::vector<uint64_t> q = read_sketches(a.need("queryfile"), nq);
  nq = q.size() / 16;
  std::vector<uint64_t> keys(nq * k), dists(nq * k);
  const std::vector<int> devs = device_list(a);
  const auto t0 = std::chrono::steady_clock::now();
  if (devs.size() > 1) {  // rows sharded over the devices, ncclAllGather + exact merge inside the library
    hb_multi* multi = nullptr;
    check(hb_multi_create(devs.data(), (int)devs.size(), &multi), "multi");
    check(hb_multi_ground_truth(multi, data.data(), len, q.data(), nq, k, keys.data(), dists.data()), "ground_truth");
    hb_multi_free(multi);
  } else {
    check(hb_ground_truth(data.data(), len, q.data(), nq, k, devs[0], keys.data(), dists.data()), "ground_truth");
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::fprintf(stderr, "ground truth of %" PRIu64 " queries over %" PRIu64 " sketches in %.3f s\n", nq, len, secs);
  if (a.str("sort", "true") == "false")
    std::fprintf(stderr, "note: --sort false asks for Bruteforce's heap order (bruteforce.rs:37-47); rows are written sorted\n");
  if (len == 0 && nq) die("no rows to search: the reference panics on an empty result (main.rs:472)");
  if (len < k) {  // Indexes::search pads with the first result (main.rs:467-474)
    std::fprintf(stderr, "warn: only %" PRIu64 " < k rows; every result row is padded with its first element\n", len);
    for (uint64_t i = 0; i < nq; ++i)
      for (uint32_t j = (uint32_t)len; j < k; ++j) {
        keys[i * k + j] = keys[i * k];
        dists[i * k + j] = dists[i * k];
      }
  }
  // keys are relative to --start like the reference's sub-range Bruteforce index
  write_results(a.str("outfile", "groundtruth.h5"), nq, k, std::move(keys), &dists, ResultAttrs{});
  return 0;
}

int cmd_inspect(const Args& a) {  // main.rs:763-821
  const std::string path = !a.positional.empty() ? a.positional[0] : a.need("indexfile");
  hb_index* idx = nullptr;
  check(hb_index_load_idx(path.c_str(), 0, 1, &idx), "read_index");
  hb_index_info info{};
  check(hb_index_get_info(idx, &info), "info");
  int fmt = 1;
  uint64_t size = 0;
  double bt = 0, qt = 0;
  char data[256], params[1024];
  check(hb_index_get_attrs(idx, &fmt, &size, &bt, &qt, data, sizeof data, params, sizeof params), "attrs");
  std::printf("ResultAttrs { format_size: %s, data: \"%s\", size: %" PRIu64 ", algo: %s, buildtime: %g, querytime: %g, params: \"%s\" }\n",
              fmt ? "true" : "false", data, size, algo_name(info.algo), bt, qt, params);
  if (info.algo == HB_ALGO_BRUTEFORCE) {
    hb_index_free(idx);
    return 0;
  }
  std::vector<uint64_t> base_off, base_ids;
  for (int g = (int)info.n_graphs - 1; g >= 0; --g) {  // layers top-down, then the base (main.rs:799-803)
    uint64_t n = 0, e = 0;
    check(hb_index_graph_dims(idx, (uint32_t)g, &n, &e), "dims");
    std::vector<uint64_t> off(n + 1), ids(e);
    check(hb_index_graph_export(idx, (uint32_t)g, nullptr, nullptr, off.data(), ids.data()), "export");
    std::vector<uint64_t> conn(n);
    for (uint64_t v = 0; v < n; ++v) conn[v] = off[v + 1] - off[v];
    std::sort(conn.begin(), conn.end());
    const std::string name = g == 0 ? "base" : "layer" + std::to_string(g - 1);
    std::printf("\n%s has %" PRIu64 " nodes, %" PRIu64 " total connections, and %" PRIu64 " average connections\n", name.c_str(), n, e, n ? e / n : 0);
    std::printf("connection distribution:\n");
    for (uint64_t i = 0; i < 11 && n; ++i) std::printf("p%" PRIu64 " %" PRIu64 "\n", i * 10, conn[std::min(n - 1, n / 10 * i)]);  // main.rs:783-785
    if (g == 0) {
      base_off.swap(off);
      base_ids.swap(ids);
    }
  }
  // `search(base.get(0), size, size)`: with ef = size the beam never fills, so the result is every node reachable from
  // the base-layer entry; run on the device while ef fits (HB_MAX_EF), counted on the exported base graph above that
  const uint64_t n = info.size;
  uint64_t reached = 0;
  std::vector<uint64_t> q0(16), sk(n ? 16 * n : 16);
  if (n) {
    check(hb_index_graph_export(idx, 0, sk.data(), nullptr, nullptr, nullptr), "export");
    std::copy(sk.begin(), sk.begin() + 16, q0.begin());
    if (n <= HB_MAX_EF) {
      std::vector<uint64_t> keys(n), dists(n);
      uint32_t found = 0;
      check(hb_search(idx, q0.data(), (uint32_t)n, (uint32_t)n, 0, keys.data(), dists.data(), &found), "search");
      reached = found;
    } else {
      uint64_t key = 0, dist = 0;
      uint32_t found = 0;
      check(hb_search(idx, q0.data(), 1, 1, 0, &key, &dist, &found), "search");  // the descent's landing point in the base graph
      std::vector<char> seen(n, 0);
      std::queue<uint64_t> fifo;
      if (found) {
        fifo.push(key);
        seen[key] = 1;
      }
      while (!fifo.empty()) {
        const uint64_t v = fifo.front();
        fifo.pop();
        ++reached;
        for (uint64_t j = base_off[v]; j < base_off[v + 1]; ++j)
          if (!seen[base_ids[j]]) {
            seen[base_ids[j]] = 1;
            fifo.push(base_ids[j]);
          }
      }
    }
  }
  std::printf("\nquery on whole index returned %" PRIu64 "/%" PRIu64 " elements\n", reached, n);
  hb_index_free(idx);
  return 0;
}

void usage() {
  std::fprintf(stderr,
               "usage: hnsw-itu <query|index|query-index|ground-truth|inspect> [options]\n"
               "  query        -d DATA -Q QUERIES [-o result.h5] [-i IDX] [-k 10] [-e 96] [-c 96] [-m 24] [-M 256] [-a hnsw|nsw|bruteforce] [-s] [-S]\n"
               "  index        -d DATA [-o index.idx] [-c 96] [-m 24] [-M 256] [-b START] [-l LEN] [-a hnsw|nsw|bruteforce] [-S]\n"
               "  query-index  -i IDX -Q QUERIES [-o result.h5] [-k 10] [-e 96] [-s] [-S] [--devices 0,1,..]\n"
               "  ground-truth -d DATA -Q QUERIES [-o groundtruth.h5] [-b START] [-l LEN] [-k 100] [-s true|false] [--devices 0,1,..]\n"
               "  inspect      IDX\n"
               "DATA/QUERIES: HDF5 with the (n, 16) u64 dataset \"hamming\", raw little-endian u64 (.u64), or .npy of shape (n, 16) <u8\n"
               "outputs ending in .h5/.hdf5 are HDF5 (datasets knns[, dists] + result attributes); any other name writes raw side files\n");
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2 || !std::strcmp(argv[1], "-h") || !std::strcmp(argv[1], "--help")) {
    usage();
    return argc < 2 ? 2 : 0;
  }
  const std::string cmd = argv[1];
  const Args a = parse(argc, argv, 2, cmd == "ground-truth");
  if (cmd == "query") return cmd_query(a);
  if (cmd == "index") return cmd_index(a);
  if (cmd == "query-index") return cmd_query_index(a);
  if (cmd == "ground-truth") return cmd_ground_truth(a);
  if (cmd == "inspect") return cmd_inspect(a);
  usage();
  return 2;
}
