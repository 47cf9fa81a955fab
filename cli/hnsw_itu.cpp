// hnsw-itu (C++ stand-in) -- the reference CLI's subcommands on the hot path, driving libhnswb200.so through the C ABI.
//
// The reference CLI is Rust + clap + HDF5 (cli/src/main.rs); this image has neither a Rust toolchain nor libhdf5, so
// this binary keeps the subcommand names, flags and result conventions and swaps the container format:
//   input  sketches: raw little-endian u64, 16 per row (`.u64`), or a NumPy `.npy` holding a (n, 16) `<u8` array --
//          the `"hamming"` dataset of the reference's HDF5 files (cli/src/dataset.rs:15-79) without the container
//   output `<out>.knns.u64` (nq x k, keys + 1 as the reference writes them, main.rs:277,745), `<out>.dists.u64`
//          (nq x k) and `<out>.attrs.json` (ResultAttrs, main.rs:311-320, written by main.rs:287-306)
//
//   hnsw-itu index        -d DATA -o OUT.idx [-c 96] [-m 24] [-M 256] [-b START] [-l LEN]        (main.rs:583-632)
//   hnsw-itu query-index  -i IDX -Q QUERIES -o OUT [-k 10] [-e 96] [-s] [-S]                      (main.rs:634-681)
//   hnsw-itu ground-truth -d DATA -Q QUERIES -o OUT [-b START] [-l LEN] [-k 100] [-s true|false]  (main.rs:683-754)
//   hnsw-itu inspect      -i IDX                                                                    (main.rs:756-821, sizes only)
// Errors print the library's message and exit non-zero (anyhow in the reference, main.rs:31,63).
#include <chrono>
#include <cinttypes>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

#include "../include/hnsw_b200.h"

namespace {

[[noreturn]] void die(const std::string& m) {
  std::fprintf(stderr, "hnsw-itu: %s\n", m.c_str());
  std::exit(1);
}
void check(int rc, const char* what) {
  if (rc != HB_OK) die(std::string(what) + ": " + hb_last_error());
}

// ---- sketch files ------------------------------------------------------------------------------------------------
std::vector<uint64_t> read_sketches(const std::string& path, uint64_t& rows) {
  std::ifstream f(path, std::ios::binary);
  if (!f) die("cannot open " + path);
  f.seekg(0, std::ios::end);
  const uint64_t bytes = (uint64_t)f.tellg();
  f.seekg(0);
  uint64_t skip = 0;
  char magic[6] = {0};
  f.read(magic, 6);
  if (bytes >= 10 && std::memcmp(magic, "\x93NUMPY", 6) == 0) {
    unsigned char ver[2], l2[2], l4[4];
    f.read(reinterpret_cast<char*>(ver), 2);
    uint32_t hlen;
    if (ver[0] == 1) {
      f.read(reinterpret_cast<char*>(l2), 2);
      hlen = l2[0] | (l2[1] << 8);
      skip = 10 + hlen;
    } else {
      f.read(reinterpret_cast<char*>(l4), 4);
      hlen = l4[0] | (l4[1] << 8) | (l4[2] << 16) | ((uint32_t)l4[3] << 24);
      skip = 12 + hlen;
    }
    std::string hdr(hlen, '\0');
    f.read(hdr.data(), hlen);
    if (hdr.find("<u8") == std::string::npos && hdr.find("<i8") == std::string::npos) die(path + ": .npy must hold little-endian 64-bit integers");
    if (hdr.find("'fortran_order': True") != std::string::npos) die(path + ": .npy must be C-ordered");
    if (hdr.find(", 16)") == std::string::npos && hdr.find(",16)") == std::string::npos) die(path + ": .npy shape must be (n, 16)");
  }
  if ((bytes - skip) % 128 != 0) die(path + ": size is not a multiple of 128 bytes (16 x u64 per sketch)");
  rows = (bytes - skip) / 128;
  std::vector<uint64_t> v(rows * 16);
  f.seekg((std::streamoff)skip);
  f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(rows * 128));
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
  bool has(const std::string& k) const { return kv.count(k) != 0; }
  std::string str(const std::string& k, const std::string& dflt = "") const { return has(k) ? kv.at(k) : dflt; }
  uint64_t num(const std::string& k, uint64_t dflt) const { return has(k) ? std::strtoull(kv.at(k).c_str(), nullptr, 10) : dflt; }
  std::string need(const std::string& k) const {
    if (!has(k)) die("missing required option --" + k);
    return kv.at(k);
  }
};

// long name for every short flag of the three subcommands (main.rs:486-549,583-602,634-664,683-712)
const std::map<std::string, std::string> kShort = {
    {"-i", "indexfile"}, {"-Q", "queryfile"}, {"-o", "outfile"}, {"-k", "k"}, {"-e", "ef"}, {"-s", "sort"}, {"-S", "single-threaded"},
    {"-d", "datafile"},  {"-b", "start"},     {"-l", "len"},     {"-c", "ef-construction"}, {"-m", "connections"}, {"-M", "max-connections"}};
const char* kBoolFlags[] = {"single-threaded", "sort"};

Args parse(int argc, char** argv, int from) {
  Args a;
  for (int i = from; i < argc; ++i) {
    std::string t = argv[i], name;
    if (t.rfind("--", 0) == 0) name = t.substr(2);
    else if (kShort.count(t)) name = kShort.at(t);
    else die("unknown option " + t);
    bool is_bool = false;
    for (const char* b : kBoolFlags) is_bool |= name == b;
    const bool next_is_value = i + 1 < argc && argv[i + 1][0] != '-';
    if (is_bool && !(name == "sort" && next_is_value)) {
      a.kv[name] = "true";
    } else {
      if (!next_is_value) die("option --" + name + " needs a value");
      a.kv[name] = argv[++i];
    }
  }
  return a;
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

// the six attributes write_result attaches to `knns` (main.rs:287-306)
void write_attrs(const std::string& path, const std::string& data, const std::string& size, const std::string& algo, double buildtime,
                 double querytime, const std::string& params) {
  std::ofstream f(path);
  if (!f) die("cannot write " + path);
  char nums[128];
  std::snprintf(nums, sizeof nums, "\"buildtime\": %.9g, \"querytime\": %.9g", buildtime, querytime);
  f << "{\"data\": \"" << json_escape(data) << "\", \"size\": \"" << size << "\", \"algo\": \"" << algo << "\", " << nums
    << ", \"params\": \"" << json_escape(params) << "\"}\n";
}

const char* algo_name(int a) { return a == HB_ALGO_BRUTEFORCE ? "Bruteforce" : a == HB_ALGO_NSW ? "Nsw" : "Hnsw"; }

// rows [start, start + len) of the data file (main.rs:136-140)
void sub_range(const Args& a, uint64_t rows, uint64_t& start, uint64_t& len) {
  start = a.num("start", 0);
  if (start > rows) die("--start is past the end of the data file");
  len = a.has("len") ? a.num("len", 0) : rows - start;
  if (start + len > rows) die("--start + --len is past the end of the data file");
}

int cmd_index(const Args& a) {
  uint64_t rows = 0, start = 0, len = 0;
  const std::vector<uint64_t> data = read_sketches(a.need("datafile"), rows);
  sub_range(a, rows, start, len);
  if (a.has("algorithm") && a.str("algorithm") != "hnsw") die("only --algorithm hnsw is built on the device");
  hb_build_params bp{};
  bp.ef_construction = (uint32_t)a.num("ef-construction", 96);
  bp.connections = (uint32_t)a.num("connections", 24);
  bp.max_connections = (uint32_t)a.num("max-connections", 256);
  bp.rayon_threads = a.has("single-threaded") ? 1 : 16;
  hb_index* idx = nullptr;
  const auto t0 = std::chrono::steady_clock::now();
  check(hb_build_hnsw(data.data() + start * 16, len, &bp, 0, 1, &idx), "build");
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  char params[128];  // main.rs:420-436
  std::snprintf(params, sizeof params, "index=(efc=%u,m=%u,M=%u)", bp.ef_construction, bp.connections, bp.max_connections);
  check(hb_index_set_attrs(idx, secs, params), "attrs");
  check(hb_index_save_idx(idx, a.str("outfile", "index.idx").c_str()), "write_index");
  std::fprintf(stderr, "indexed %" PRIu64 " sketches in %.2f s -> %s\n", len, secs, a.str("outfile", "index.idx").c_str());
  hb_index_free(idx);
  return 0;
}

int cmd_query_index(const Args& a) {
  const uint32_t k = (uint32_t)a.num("k", 10), ef = (uint32_t)a.num("ef", 96);
  if (k > ef) std::fprintf(stderr, "error: `k` (%u) is greater than `ef` (%u), this can have adverse effects\n", k, ef);  // main.rs:190-195
  hb_index* idx = nullptr;
  check(hb_index_load_idx(a.need("indexfile").c_str(), 0, 0, &idx), "read_index");
  uint64_t nq = 0;
  const std::vector<uint64_t> q = read_sketches(a.need("queryfile"), nq);
  std::vector<uint64_t> keys(nq * k), dists(nq * k);
  std::vector<uint32_t> found(nq);
  const auto t0 = std::chrono::steady_clock::now();
  if (a.has("single-threaded")) {  // main.rs:203-207: one search per query
    for (uint64_t i = 0; i < nq; ++i)
      check(hb_search(idx, q.data() + i * 16, k, ef, HB_PAD_LIKE_CLI, keys.data() + i * k, dists.data() + i * k, &found[i]), "search");
  } else {
    check(hb_knns(idx, q.data(), nq, k, ef, HB_PAD_LIKE_CLI, keys.data(), dists.data(), found.data(), nullptr), "knns");
  }
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::fprintf(stderr, "Total query time: %.6f s, per query: %.3f us\n", secs, nq ? secs / nq * 1e6 : 0.0);
  for (uint64_t i = 0; i < nq; ++i) {
    if (found[i] == 0) die("query " + std::to_string(i) + " returned no result (the reference panics here, main.rs:472)");
    if (found[i] < k) std::fprintf(stderr, "warn: query %" PRIu64 " found %u < k results; padded with the first\n", i, found[i]);
  }
  // results are ascending (distance, key) already, so -s/--sort (main.rs:273-275) changes nothing except for padded rows,
  // where the reference's sort moves the repeated first element to the front -- which is where it already is
  for (auto& x : keys) x += 1;  // main.rs:277
  const std::string out = a.str("outfile", "result");
  write_u64(out + ".knns.u64", keys);
  write_u64(out + ".dists.u64", dists);
  int fmt = 1;
  uint64_t size = 0;
  double buildtime = 0, qt = 0;
  char data[256], params[1024];
  check(hb_index_get_attrs(idx, &fmt, &size, &buildtime, &qt, data, sizeof data, params, sizeof params), "attrs");
  hb_index_info info{};
  check(hb_index_get_info(idx, &info), "info");
  write_attrs(out + ".attrs.json", data, size_string(size, fmt != 0), algo_name(info.algo), buildtime, secs,
              std::string(params) + ",query=(ef=" + std::to_string(ef) + ")");  // main.rs:218-219
  hb_index_free(idx);
  return 0;
}

int cmd_ground_truth(const Args& a) {
  const uint32_t k = (uint32_t)a.num("k", 100);
  uint64_t rows = 0, start = 0, len = 0, nq = 0;
  const std::vector<uint64_t> data = read_sketches(a.need("datafile"), rows);
  sub_range(a, rows, start, len);
  const std::vector<uint64_t> q = read_sketches(a.need("queryfile"), nq);
  std::vector<uint64_t> keys(nq * k), dists(nq * k);
  const auto t0 = std::chrono::steady_clock::now();
  check(hb_ground_truth(data.data() + start * 16, len, q.data(), nq, k, 0, keys.data(), dists.data()), "ground_truth");
  const double secs = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  std::fprintf(stderr, "ground truth of %" PRIu64 " queries over %" PRIu64 " sketches in %.3f s\n", nq, len, secs);
  if (a.str("sort", "true") == "false")
    std::fprintf(stderr, "note: --sort false asks for Bruteforce's heap order (bruteforce.rs:37-47); rows are written sorted\n");
  if (len < k) die("fewer than k rows: the reference panics when a query has no k results");
  for (auto& x : keys) x += 1;  // main.rs:745; keys are relative to --start like the reference's sub-range index
  const std::string out = a.str("outfile", "groundtruth");
  write_u64(out + ".knns.u64", keys);
  write_u64(out + ".dists.u64", dists);
  return 0;
}

int cmd_inspect(const Args& a) {
  hb_index* idx = nullptr;
  check(hb_index_load_idx(a.need("indexfile").c_str(), 0, 1, &idx), "read_index");
  hb_index_info info{};
  check(hb_index_get_info(idx, &info), "info");
  std::printf("algo %s size %" PRIu64 " graphs %u entry_point %" PRId64 "\n", algo_name(info.algo), info.size, info.n_graphs, info.entry_point);
  for (uint32_t g = 0; g < info.n_graphs; ++g) {
    uint64_t n = 0, e = 0;
    check(hb_index_graph_dims(idx, g, &n, &e), "dims");
    std::printf("%s %u: %" PRIu64 " nodes, %" PRIu64 " directed edges, mean degree %.2f\n", g ? "layer" : "base", g, n, e, n ? (double)e / n : 0.0);
  }
  hb_index_free(idx);
  return 0;
}

void usage() {
  std::fprintf(stderr,
               "usage: hnsw-itu <index|query-index|ground-truth|inspect> [options]\n"
               "  index        -d DATA -o OUT.idx [-c 96] [-m 24] [-M 256] [-b START] [-l LEN] [-S]\n"
               "  query-index  -i IDX -Q QUERIES [-o result] [-k 10] [-e 96] [-s] [-S]\n"
               "  ground-truth -d DATA -Q QUERIES [-o groundtruth] [-b START] [-l LEN] [-k 100] [-s true|false]\n"
               "  inspect      -i IDX\n"
               "DATA/QUERIES: raw little-endian u64, 16 per sketch (.u64), or .npy of shape (n, 16) <u8\n");
}

}  // namespace

int main(int argc, char** argv) {
  if (argc < 2 || !std::strcmp(argv[1], "-h") || !std::strcmp(argv[1], "--help")) {
    usage();
    return argc < 2 ? 2 : 0;
  }
  const std::string cmd = argv[1];
  const Args a = parse(argc, argv, 2);
  if (cmd == "index") return cmd_index(a);
  if (cmd == "query-index") return cmd_query_index(a);
  if (cmd == "ground-truth") return cmd_ground_truth(a);
  if (cmd == "inspect") return cmd_inspect(a);
  usage();
  return 2;
}
