// h5tool -- exercises cli/h5min.hpp from the command line (used by tests/test_h5min.py; no GPU, no libhdf5):
//   h5tool pack OUT.h5 NAME COLS IN.u64 [NAME2 COLS2 IN2.u64 ...] [--attr-str K V]... [--attr-f64 K V]...
//   h5tool dump IN.h5 NAME OUT.u64        (prints "rows cols")
//   h5tool attrs IN.h5                    (prints one "name<TAB>kind<TAB>value" line per root attribute)
#include <cstdio>
#include <cstdlib>
#include <iostream>

#include "h5min.hpp"

static std::vector<uint64_t> slurp(const std::string& p) {
  std::ifstream f(p, std::ios::binary);
  if (!f) throw h5min::Error("cannot open " + p);
  f.seekg(0, std::ios::end);
  std::vector<uint64_t> v((size_t)f.tellg() / 8);
  f.seekg(0);
  f.read(reinterpret_cast<char*>(v.data()), (std::streamsize)(v.size() * 8));
  return v;
}

int main(int argc, char** argv) {
  try {
    const std::string cmd = argc > 1 ? argv[1] : "";
    if (cmd == "pack" && argc >= 6) {
      std::vector<std::vector<uint64_t>> keep;
      std::vector<h5min::OutDataset> ds;
      std::vector<std::pair<std::string, h5min::Attr>> attrs;
      int i = 3;
      while (i < argc) {
        const std::string t = argv[i];
        if (t == "--attr-str" && i + 2 < argc) {
          h5min::Attr a;
          a.kind = h5min::Attr::STR;
          a.s = argv[i + 2];
          attrs.push_back({argv[i + 1], a});
          i += 3;
        } else if (t == "--attr-f64" && i + 2 < argc) {
          h5min::Attr a;
          a.f = std::atof(argv[i + 2]);
          attrs.push_back({argv[i + 1], a});
          i += 3;
        } else if (i + 2 < argc) {
          keep.push_back(slurp(argv[i + 2]));
          const uint64_t cols = std::strtoull(argv[i + 1], nullptr, 10);
          ds.push_back({t, cols ? keep.back().size() / cols : 0, cols, nullptr});
          i += 3;
        } else {
          throw h5min::Error("bad arguments");
        }
      }
      for (size_t j = 0; j < ds.size(); ++j) ds[j].data = keep[j].data();
      h5min::Writer::write(argv[2], ds, attrs);
      return 0;
    }
    if (cmd == "dump" && argc == 5) {
      h5min::Reader r(argv[2]);
      const h5min::DatasetInfo d = r.dataset(argv[3]);
      std::vector<uint64_t> v(d.dims[0] * d.dims[1]);
      r.read_rows_u64(d, 0, d.dims[0], v.data());
      std::ofstream f(argv[4], std::ios::binary);
      f.write(reinterpret_cast<const char*>(v.data()), (std::streamsize)(v.size() * 8));
      std::printf("%llu %llu\n", (unsigned long long)d.dims[0], (unsigned long long)d.dims[1]);
      return 0;
    }
    if (cmd == "attrs" && argc == 3) {
      h5min::Reader r(argv[2]);
      for (const auto& kv : r.root_attrs()) {
        if (kv.second.kind == h5min::Attr::F64) std::printf("%s\tf64\t%.17g\n", kv.first.c_str(), kv.second.f);
        else std::printf("%s\tstr\t%s\n", kv.first.c_str(), kv.second.s.c_str());
      }
      return 0;
    }
    std::fprintf(stderr, "usage: h5tool pack|dump|attrs ...\n");
    return 2;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "h5tool: %s\n", e.what());
    return 1;
  }
}
