// h5min.hpp -- a minimal HDF5 reader/writer for the CLI stand-in, header only, no libhdf5.
//
// The reference CLI does all its I/O through libhdf5 (cli/src/dataset.rs:15-79,116-143; attributes main.rs:287-306):
//   read   2-D u64 dataset "hamming" (n x 16) of the data / query files, in slices of 50 000 rows
//   write  2-D u64 datasets "knns" (and "dists" for ground-truth) plus six ROOT-GROUP attributes: data, size, algo,
//          params as variable-length UTF-8 strings (hdf5::types::VarLenUnicode) and buildtime, querytime as f64 scalars
//          (`add_attr` goes through `self.file.new_attr`, dataset.rs:56-62).
// libhdf5 is not in this image, so this file implements the subset of the HDF5 File Format Specification (version 3.0)
// those calls touch, written from the specification:
//   reader  superblock v0-v3; object headers v1 and v2 (with continuation blocks); root-group lookup through a v1
//           symbol table (local heap + v1 B-tree + symbol nodes) or compact link messages; dataspace v1/v2; fixed-point
//           8-byte datatypes; data layout v3 contiguous and chunked (v1 B-tree chunk index, no filters); root attributes
//           (f64 scalars, fixed and variable-length strings through the global heap)
//   writer  what libhdf5 emits with default (earliest) format bounds: superblock v0, v1 object headers, symbol-table
//           root group, contiguous datasets, variable-length strings in one global heap collection
// It CANNOT be checked against libhdf5/h5py here (neither exists in this environment); tests/test_h5min.py checks it
// against an independent Python restatement of the same specification sections and by round trip.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace h5min {

constexpr uint64_t kUndef = ~0ull;
struct Error : std::runtime_error {
  using std::runtime_error::runtime_error;
};

struct Attr {
  enum Kind { F64, STR } kind = F64;
  double f = 0.0;
  std::string s;
};

struct DatasetInfo {
  std::vector<uint64_t> dims;
  uint32_t elem_size = 0;
  bool chunked = false;
  uint64_t data_addr = kUndef, data_size = 0;  // contiguous
  uint64_t btree_addr = kUndef;                // chunked
  std::vector<uint64_t> chunk_dims;            // chunked (elements per dimension)
};

// ------------------------------------------------------------------------------------------------------------ reader
class Reader {
 public:
  explicit Reader(const std::string& path) : f_(path, std::ios::binary), path_(path) {
    if (!f_) throw Error("cannot open " + path);
    f_.seekg(0, std::ios::end);
    size_ = (uint64_t)f_.tellg();
    parse_superblock();
  }

  bool has_dataset(const std::string& name) { return links().count(name) != 0; }

  DatasetInfo dataset(const std::string& name) {
    auto it = links().find(name);
    if (it == links().end()) throw Error(path_ + ": no dataset \"" + name + "\" in the root group");
    DatasetInfo d;
    bool have_space = false, have_type = false, have_layout = false;
    for (const Msg& m : messages(it->second)) {
      const uint8_t* p = m.data.data();
      if (m.type == 0x0001) {  // dataspace
        const uint8_t ver = p[0], rank = p[1];
        const size_t off = ver == 1 ? 8 : 4;
        if (ver != 1 && ver != 2) throw Error("dataspace version " + std::to_string(ver) + " not supported");
        for (int i = 0; i < rank; ++i) d.dims.push_back(rd(p + off + 8 * i, len_size_));
        have_space = true;
      } else if (m.type == 0x0003) {  // datatype
        const uint8_t cls = p[0] & 0x0F;
        d.elem_size = (uint32_t)rd(p + 4, 4);
        if (cls != 0) throw Error("dataset \"" + name + "\": only fixed-point element types are supported");
        if (p[1] & 1) throw Error("dataset \"" + name + "\": big-endian integers are not supported");
        have_type = true;
      } else if (m.type == 0x0008) {  // data layout
        const uint8_t ver = p[0];
        if (ver != 3) throw Error("data layout version " + std::to_string(ver) + " not supported (libver=latest files use v4)");
        const uint8_t cls = p[1];
        if (cls == 1) {
          d.data_addr = rd(p + 2, off_size_);
          d.data_size = rd(p + 2 + off_size_, len_size_);
        } else if (cls == 2) {
          d.chunked = true;
          const uint8_t nd = p[2];  // dataset rank + 1 (the element size is the last "dimension")
          d.btree_addr = rd(p + 3, off_size_);
          for (int i = 0; i + 1 < nd; ++i) d.chunk_dims.push_back(rd(p + 3 + off_size_ + 4 * i, 4));
        } else if (cls == 0) {
          throw Error("compact dataset layout not supported");
        } else {
          throw Error("unknown data layout class");
        }
        have_layout = true;
      } else if (m.type == 0x000B) {  // filter pipeline
        throw Error("dataset \"" + name + "\" uses filters (compression); only unfiltered data is supported");
      }
    }
    if (!have_space || !have_type || !have_layout) throw Error("dataset \"" + name + "\": incomplete object header");
    return d;
  }

  // rows [row0, row0 + nrows) of a 2-D dataset of 8-byte integers into out (nrows * dims[1] values)
  void read_rows_u64(const DatasetInfo& d, uint64_t row0, uint64_t nrows, uint64_t* out) {
    if (d.dims.size() != 2 || d.elem_size != 8) throw Error("expected a 2-D dataset of 8-byte integers");
    if (row0 + nrows > d.dims[0]) throw Error("row range past the end of the dataset");
    const uint64_t cols = d.dims[1];
    if (!d.chunked) {
      if (d.data_addr == kUndef) {  // never written: fill value 0
        std::memset(out, 0, nrows * cols * 8);
        return;
      }
      read_at(base_ + d.data_addr + row0 * cols * 8, out, nrows * cols * 8);
      return;
    }
    std::memset(out, 0, nrows * cols * 8);
    read_chunks(d, d.btree_addr, row0, nrows, out);
  }

  const std::map<std::string, Attr>& root_attrs() {
    if (!attrs_loaded_) {
      for (const Msg& m : messages(root_header_))
        if (m.type == 0x000C) parse_attr(m);
      attrs_loaded_ = true;
    }
    return attrs_;
  }

 private:
  struct Msg {
    uint16_t type;
    uint8_t flags;
    std::vector<uint8_t> data;
  };

  static uint64_t rd(const uint8_t* p, int n) {
    uint64_t v = 0;
    for (int i = 0; i < n; ++i) v |= (uint64_t)p[i] << (8 * i);
    if (n < 8 && v == ((1ull << (8 * n)) - 1)) return kUndef;  // undefined address in a narrower field
    return v;
  }
  void read_at(uint64_t pos, void* dst, uint64_t n) {
    if (pos + n > size_) throw Error(path_ + ": read past the end of the file (corrupt or truncated)");
    f_.seekg((std::streamoff)pos);
    f_.read(reinterpret_cast<char*>(dst), (std::streamsize)n);
    if (!f_) throw Error(path_ + ": short read");
  }
  std::vector<uint8_t> bytes_at(uint64_t pos, uint64_t n) {
    std::vector<uint8_t> b(n);
    read_at(pos, b.data(), n);
    return b;
  }

  void parse_superblock() {
    static const uint8_t sig[8] = {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n'};
    uint64_t at = 0;
    for (;; at = at ? at * 2 : 512) {  // the superblock may sit at 0, 512, 1024, ...
      if (at + 8 > size_) throw Error(path_ + ": not an HDF5 file (no superblock signature)");
      uint8_t s[8];
      read_at(at, s, 8);
      if (!std::memcmp(s, sig, 8)) break;
    }
    const std::vector<uint8_t> b = bytes_at(at, std::min<uint64_t>(size_ - at, 128));
    const uint8_t ver = b[8];
    if (ver <= 1) {
      off_size_ = b[13];
      len_size_ = b[14];
      size_t p = 24 + (ver == 1 ? 4 : 0);
      base_ = rd(&b[p], off_size_);
      p += 4 * (size_t)off_size_;  // base, free-space, end-of-file, driver-info addresses
      // root group symbol table entry: link name offset, object header address, cache type, reserved, scratch
      root_header_ = rd(&b[p + off_size_], off_size_);
      const uint32_t cache = (uint32_t)rd(&b[p + 2 * off_size_], 4);
      if (cache == 1) {
        root_btree_ = rd(&b[p + 2 * off_size_ + 8], off_size_);
        root_heap_ = rd(&b[p + 2 * off_size_ + 8 + off_size_], off_size_);
      }
    } else if (ver == 2 || ver == 3) {
      off_size_ = b[9];
      len_size_ = b[10];
      base_ = rd(&b[12], off_size_);
      root_header_ = rd(&b[12 + 3 * (size_t)off_size_], off_size_);
    } else {
      throw Error(path_ + ": superblock version " + std::to_string(ver) + " not supported");
    }
    if (base_ == kUndef) base_ = 0;
    if (off_size_ != 8 && off_size_ != 4) throw Error("unsupported size of offsets");
  }

  // every message of an object header (v1 or v2), continuation blocks followed
  std::vector<Msg> messages(uint64_t addr) {
    std::vector<Msg> out;
    uint8_t head[16];
    read_at(base_ + addr, head, 16);
    if (!std::memcmp(head, "OHDR", 4)) {
      const uint8_t flags = head[5];
      size_t p = 6;
      if (flags & 0x20) p += 16;  // times
      if (flags & 0x10) p += 4;   // max compact / min dense attributes
      const int szlen = 1 << (flags & 3);
      std::vector<uint8_t> pre = bytes_at(base_ + addr, p + szlen);
      const uint64_t chunk0 = rd(&pre[p], szlen);
      std::vector<std::pair<uint64_t, uint64_t>> blocks{{addr + p + szlen, chunk0}};
      const bool tracked = flags & 0x04;
      for (size_t bi = 0; bi < blocks.size(); ++bi) {
        std::vector<uint8_t> blk = bytes_at(base_ + blocks[bi].first, blocks[bi].second);
        // chunk 0's size counts the message bytes only; a continuation block is "OCHK" + messages + 4-byte checksum
        size_t q = bi == 0 ? 0 : 4;
        const size_t end = bi == 0 ? blk.size() : (blk.size() >= 4 ? blk.size() - 4 : 0);
        while (q + 4 + (tracked ? 2 : 0) <= end) {
          Msg m;
          m.type = blk[q];
          const uint64_t sz = rd(&blk[q + 1], 2);
          m.flags = blk[q + 3];
          q += 4 + (tracked ? 2 : 0);
          if (q + sz > end) break;
          m.data.assign(blk.begin() + q, blk.begin() + q + sz);
          q += sz;
          if (m.type == 0x0010) blocks.push_back({rd(m.data.data(), off_size_), rd(m.data.data() + off_size_, len_size_)});
          else if (m.type != 0) out.push_back(std::move(m));
        }
      }
      return out;
    }
    if (head[0] != 1) throw Error("object header version " + std::to_string(head[0]) + " not supported");
    const uint32_t nmsg = (uint32_t)rd(head + 2, 2);
    const uint64_t hsize = rd(head + 8, 4);
    std::vector<std::pair<uint64_t, uint64_t>> blocks{{addr + 16, hsize}};
    uint32_t seen = 0;
    for (size_t bi = 0; bi < blocks.size() && seen < nmsg; ++bi) {
      std::vector<uint8_t> blk = bytes_at(base_ + blocks[bi].first, blocks[bi].second);
      size_t q = 0;
      while (q + 8 <= blk.size() && seen < nmsg) {
        Msg m;
        m.type = (uint16_t)rd(&blk[q], 2);
        const uint64_t sz = rd(&blk[q + 2], 2);
        m.flags = blk[q + 4];
        q += 8;
        if (q + sz > blk.size()) throw Error("object header message overruns its block");
        m.data.assign(blk.begin() + q, blk.begin() + q + sz);
        q += sz;
        ++seen;
        if (m.type == 0x0010) blocks.push_back({rd(m.data.data(), off_size_), rd(m.data.data() + off_size_, len_size_)});
        else if (m.type != 0) out.push_back(std::move(m));
      }
    }
    return out;
  }

  // name -> object header address of every link in the root group
  const std::map<std::string, uint64_t>& links() {
    if (links_loaded_) return links_;
    links_loaded_ = true;
    uint64_t btree = root_btree_, heap = root_heap_;
    for (const Msg& m : messages(root_header_)) {
      if (m.type == 0x0011) {  // symbol table message
        btree = rd(m.data.data(), off_size_);
        heap = rd(m.data.data() + off_size_, off_size_);
      } else if (m.type == 0x0006) {  // link message (compact v2 group)
        const uint8_t* p = m.data.data();
        const uint8_t flags = p[1];
        size_t q = 2;
        uint8_t ltype = 0;
        if (flags & 0x08) ltype = p[q++];
        if (flags & 0x04) q += 8;  // creation order
        if (flags & 0x10) q += 1;  // character set
        const int nlen = 1 << (flags & 3);
        const uint64_t n = rd(p + q, nlen);
        q += nlen;
        const std::string name(reinterpret_cast<const char*>(p + q), (size_t)n);
        q += n;
        if (ltype == 0) links_[name] = rd(p + q, off_size_);
      } else if (m.type == 0x0002) {  // link info: dense storage lives in a fractal heap
        if (rd(m.data.data() + 2 + ((m.data[1] & 1) ? 8 : 0), off_size_) != kUndef)
          throw Error(path_ + ": the root group uses dense link storage (fractal heap), which is not supported");
      }
    }
    if (btree != kUndef && heap != kUndef) {
      // local heap: "HEAP" version reserved[3] data-size free-list-head data-address
      const std::vector<uint8_t> h = bytes_at(base_ + heap, 8 + 2 * (size_t)len_size_ + off_size_);
      if (std::memcmp(h.data(), "HEAP", 4)) throw Error(path_ + ": bad local heap signature");
      const uint64_t hsize = rd(&h[8], len_size_), haddr = rd(&h[8 + 2 * (size_t)len_size_], off_size_);
      heap_data_ = bytes_at(base_ + haddr, hsize);
      walk_group_btree(btree);
    }
    return links_;
  }

  void walk_group_btree(uint64_t addr) {
    const std::vector<uint8_t> hd = bytes_at(base_ + addr, 8 + 2 * (size_t)off_size_);
    if (std::memcmp(hd.data(), "TREE", 4) || hd[4] != 0) throw Error(path_ + ": bad group B-tree node");
    const uint8_t level = hd[5];
    const uint32_t used = (uint32_t)rd(&hd[6], 2);
    // keys are heap offsets (len_size_ bytes), children are addresses: key0 child0 key1 child1 ... key_used
    const std::vector<uint8_t> body = bytes_at(base_ + addr + hd.size(), (uint64_t)used * (len_size_ + off_size_) + len_size_);
    for (uint32_t i = 0; i < used; ++i) {
      const uint64_t child = rd(&body[(size_t)i * (len_size_ + off_size_) + len_size_], off_size_);
      if (level > 0) {
        walk_group_btree(child);
        continue;
      }
      const std::vector<uint8_t> sn = bytes_at(base_ + child, 8);
      if (std::memcmp(sn.data(), "SNOD", 4)) throw Error(path_ + ": bad symbol table node");
      const uint32_t nsym = (uint32_t)rd(&sn[6], 2);
      const size_t esz = 2 * (size_t)off_size_ + 8 + 16;
      const std::vector<uint8_t> ent = bytes_at(base_ + child + 8, nsym * esz);
      for (uint32_t e = 0; e < nsym; ++e) {
        const uint64_t noff = rd(&ent[e * esz], off_size_), oaddr = rd(&ent[e * esz + off_size_], off_size_);
        if (noff >= heap_data_.size()) throw Error(path_ + ": link name offset outside the local heap");
        links_[std::string(reinterpret_cast<const char*>(&heap_data_[noff]))] = oaddr;
      }
    }
  }

  // chunked layout: v1 B-tree of type 1; key = chunk size (4) filter mask (4) offsets ((rank+1) x 8)
  void read_chunks(const DatasetInfo& d, uint64_t addr, uint64_t row0, uint64_t nrows, uint64_t* out) {
    const std::vector<uint8_t> hd = bytes_at(base_ + addr, 8 + 2 * (size_t)off_size_);
    if (std::memcmp(hd.data(), "TREE", 4) || hd[4] != 1) throw Error(path_ + ": bad chunk B-tree node");
    const uint8_t level = hd[5];
    const uint32_t used = (uint32_t)rd(&hd[6], 2);
    const size_t rank = d.dims.size(), ksz = 8 + 8 * (rank + 1);
    const std::vector<uint8_t> body = bytes_at(base_ + addr + hd.size(), (uint64_t)used * (ksz + off_size_) + ksz);
    const uint64_t cols = d.dims[1], crow = d.chunk_dims[0], ccol = d.chunk_dims[1];
    for (uint32_t i = 0; i < used; ++i) {
      const uint8_t* key = &body[(size_t)i * (ksz + off_size_)];
      const uint64_t child = rd(key + ksz, off_size_);
      if (level > 0) {
        read_chunks(d, child, row0, nrows, out);
        continue;
      }
      const uint64_t csize = rd(key, 4), r0 = rd(key + 8, 8), c0 = rd(key + 16, 8);
      if (rd(key + 4, 4) != 0) throw Error(path_ + ": filtered chunk (compression) not supported");
      if (r0 >= row0 + nrows || r0 + crow <= row0) continue;
      if (csize != crow * ccol * 8) throw Error(path_ + ": unexpected chunk size");
      std::vector<uint64_t> chunk(crow * ccol);
      read_at(base_ + child, chunk.data(), csize);
      for (uint64_t r = std::max(r0, row0); r < std::min({r0 + crow, row0 + nrows, d.dims[0]}); ++r)
        for (uint64_t c = c0; c < std::min(c0 + ccol, cols); ++c) out[(r - row0) * cols + c] = chunk[(r - r0) * ccol + (c - c0)];
    }
  }

  std::string global_heap_object(uint64_t coll, uint32_t index) {
    const std::vector<uint8_t> h = bytes_at(base_ + coll, 8 + len_size_);
    if (std::memcmp(h.data(), "GCOL", 4)) throw Error(path_ + ": bad global heap signature");
    const uint64_t csize = rd(&h[8], len_size_);
    const std::vector<uint8_t> c = bytes_at(base_ + coll, csize);
    size_t q = 8 + len_size_;
    while (q + 8 + len_size_ <= c.size()) {
      const uint32_t idx = (uint32_t)rd(&c[q], 2);
      const uint64_t osz = rd(&c[q + 8], len_size_);
      if (idx == 0) break;
      if (idx == index) return std::string(reinterpret_cast<const char*>(&c[q + 8 + len_size_]), (size_t)osz);
      q += 8 + len_size_ + ((osz + 7) & ~7ull);
    }
    throw Error(path_ + ": global heap object not found");
  }

  void parse_attr(const Msg& m) {
    const uint8_t* p = m.data.data();
    const uint8_t ver = p[0];
    if (ver < 1 || ver > 3) return;
    const uint64_t nsz = rd(p + 2, 2), tsz = rd(p + 4, 2), ssz = rd(p + 6, 2);
    size_t q = 8 + (ver == 3 ? 1 : 0);
    auto pad = [&](uint64_t x) { return ver == 1 ? ((x + 7) & ~7ull) : x; };
    const std::string name(reinterpret_cast<const char*>(p + q));
    q += pad(nsz);
    const uint8_t* dt = p + q;
    q += pad(tsz);
    q += pad(ssz);  // scalar dataspace: nothing to read
    const uint8_t* data = p + q;
    const uint8_t cls = dt[0] & 0x0F;
    const uint32_t tsize = (uint32_t)rd(dt + 4, 4);
    Attr a;
    if (cls == 1 && tsize == 8) {
      a.kind = Attr::F64;
      std::memcpy(&a.f, data, 8);
    } else if (cls == 3) {
      a.kind = Attr::STR;
      a.s.assign(reinterpret_cast<const char*>(data), strnlen(reinterpret_cast<const char*>(data), tsize));
    } else if (cls == 9 && (dt[1] & 0x0F) == 1) {  // variable-length string: length, collection address, object index
      a.kind = Attr::STR;
      const uint64_t coll = rd(data + 4, off_size_);
      const uint32_t idx = (uint32_t)rd(data + 4 + off_size_, 4);
      a.s = coll == kUndef || coll == 0 ? std::string() : global_heap_object(coll, idx);
      while (!a.s.empty() && a.s.back() == '\0') a.s.pop_back();
    } else {
      return;
    }
    attrs_[name] = a;
  }

  std::ifstream f_;
  std::string path_;
  uint64_t size_ = 0, base_ = 0, root_header_ = kUndef, root_btree_ = kUndef, root_heap_ = kUndef;
  int off_size_ = 8, len_size_ = 8;
  bool links_loaded_ = false, attrs_loaded_ = false;
  std::map<std::string, uint64_t> links_;
  std::map<std::string, Attr> attrs_;
  std::vector<uint8_t> heap_data_;
};

// ------------------------------------------------------------------------------------------------------------ writer
// One call writes the whole file: 2-D u64 datasets (contiguous) + root attributes.
struct OutDataset {
  std::string name;
  uint64_t rows = 0, cols = 0;
  const uint64_t* data = nullptr;  // rows * cols values, row-major
};

class Writer {
 public:
  static void write(const std::string& path, const std::vector<OutDataset>& datasets,
                    const std::vector<std::pair<std::string, Attr>>& attrs) {
    if (datasets.empty() || datasets.size() > 8) throw Error("h5min writes 1..8 datasets per file (one symbol table node)");
    std::vector<OutDataset> ds = datasets;
    std::sort(ds.begin(), ds.end(), [](const OutDataset& a, const OutDataset& b) { return a.name < b.name; });  // SNOD entries are sorted by name

    // ---- local heap data segment: "" at offset 0, then the link names, each NUL-terminated and padded to 8
    std::vector<uint8_t> heap(8, 0);
    std::vector<uint64_t> name_off;
    for (const OutDataset& d : ds) {
      name_off.push_back(heap.size());
      heap.insert(heap.end(), d.name.begin(), d.name.end());
      heap.push_back(0);
      while (heap.size() % 8) heap.push_back(0);
    }
    const uint64_t heap_data_size = std::max<uint64_t>(heap.size() + 16, 88);  // room for the free block the heap must end with
    heap.resize(heap_data_size, 0);
    const uint64_t free_off = name_off.empty() ? 8 : (uint64_t)((name_off.back() + ds.back().name.size() + 1 + 7) & ~7ull);
    // free block: offset of next free block (1 = none: "last"), size of this block
    put(heap, free_off, 1, 8);
    put(heap, free_off + 8, heap_data_size - free_off, 8);

    // ---- global heap: one collection with the variable-length strings
    std::vector<uint8_t> gcol;
    std::vector<uint32_t> str_index(attrs.size(), 0);
    {
      uint32_t next = 1;
      std::vector<uint8_t> objs;
      for (size_t i = 0; i < attrs.size(); ++i) {
        if (attrs[i].second.kind != Attr::STR) continue;
        const std::string& s = attrs[i].second.s;
        str_index[i] = next;
        append(objs, next++, 2);
        append(objs, 1, 2);  // reference count
        append(objs, 0, 4);
        append(objs, s.size(), 8);
        objs.insert(objs.end(), s.begin(), s.end());
        while (objs.size() % 8) objs.push_back(0);
      }
      const uint64_t csize = std::max<uint64_t>(4096, (16 + objs.size() + 16 + 7) & ~7ull);
      gcol.insert(gcol.end(), {'G', 'C', 'O', 'L', 1, 0, 0, 0});
      append(gcol, csize, 8);
      gcol.insert(gcol.end(), objs.begin(), objs.end());
      // object 0: the free space, its size field counts the object header too
      append(gcol, 0, 2);
      append(gcol, 0, 2);
      append(gcol, 0, 4);
      append(gcol, csize - (16 + objs.size()), 8);
      gcol.resize(csize, 0);
    }

    // ---- addresses: superblock | root header | B-tree | SNOD | heap header | heap data | global heap | dataset headers | raw data
    const uint64_t a_root = 96;
    const std::vector<uint8_t> root_msgs_probe = root_messages(0, 0, attrs, str_index, 0);
    const uint64_t root_size = 16 + root_msgs_probe.size();
    const uint64_t a_btree = align8(a_root + root_size);
    const uint64_t btree_size = 24 + 2 * 16 * 8 + (2 * 16 + 1) * 8;
    const uint64_t a_snod = a_btree + btree_size;
    const uint64_t snod_size = 8 + 8 * 40;
    const uint64_t a_heap = a_snod + snod_size;
    const uint64_t a_heap_data = a_heap + 32;
    const uint64_t a_gcol = align8(a_heap_data + heap_data_size);
    uint64_t cur = a_gcol + gcol.size();
    std::vector<uint64_t> a_hdr(ds.size()), a_data(ds.size());
    std::vector<std::vector<uint8_t>> hdrs(ds.size());
    for (size_t i = 0; i < ds.size(); ++i) {
      a_hdr[i] = cur;
      hdrs[i] = dataset_header(ds[i], 0);
      cur = align8(cur + hdrs[i].size());
    }
    for (size_t i = 0; i < ds.size(); ++i) {
      a_data[i] = (ds[i].rows * ds[i].cols) != 0 ? cur : kUndef;
      cur += ds[i].rows * ds[i].cols * 8;
      hdrs[i] = dataset_header(ds[i], a_data[i]);
    }
    const uint64_t eof = cur;

    std::vector<uint8_t> out;
    // superblock version 0
    out.insert(out.end(), {0x89, 'H', 'D', 'F', '\r', '\n', 0x1a, '\n', 0, 0, 0, 0, 0, 8, 8, 0});
    append(out, 4, 2);    // group leaf node K
    append(out, 16, 2);   // group internal node K
    append(out, 0, 4);    // file consistency flags
    append(out, 0, 8);    // base address
    append(out, kUndef, 8);  // free-space info
    append(out, eof, 8);
    append(out, kUndef, 8);  // driver information block
    append(out, 0, 8);       // root symbol table entry: link name offset
    append(out, a_root, 8);
    append(out, 1, 4);  // cache type 1: scratch holds the B-tree and heap addresses
    append(out, 0, 4);
    append(out, a_btree, 8);
    append(out, a_heap, 8);
    // root group object header (version 1)
    const std::vector<uint8_t> rmsgs = root_messages(a_btree, a_heap, attrs, str_index, a_gcol);
    object_header_prefix(out, count_messages(rmsgs), rmsgs.size());
    out.insert(out.end(), rmsgs.begin(), rmsgs.end());
    out.resize(a_btree, 0);
    // group B-tree: one leaf entry
    out.insert(out.end(), {'T', 'R', 'E', 'E', 0, 0});
    append(out, 1, 2);
    append(out, kUndef, 8);
    append(out, kUndef, 8);
    append(out, 0, 8);  // key 0: the empty string
    append(out, a_snod, 8);
    append(out, name_off.back(), 8);  // key 1: the largest name in the child
    out.resize(a_snod, 0);
    // symbol table node
    out.insert(out.end(), {'S', 'N', 'O', 'D', 1, 0});
    append(out, ds.size(), 2);
    for (size_t i = 0; i < ds.size(); ++i) {
      append(out, name_off[i], 8);
      append(out, a_hdr[i], 8);
      append(out, 0, 4);
      append(out, 0, 4);
      append(out, 0, 8);
      append(out, 0, 8);
    }
    out.resize(a_heap, 0);
    // local heap
    out.insert(out.end(), {'H', 'E', 'A', 'P', 0, 0, 0, 0});
    append(out, heap_data_size, 8);
    append(out, free_off, 8);
    append(out, a_heap_data, 8);
    out.insert(out.end(), heap.begin(), heap.end());
    out.resize(a_gcol, 0);
    out.insert(out.end(), gcol.begin(), gcol.end());
    for (size_t i = 0; i < ds.size(); ++i) {
      out.resize(a_hdr[i], 0);
      out.insert(out.end(), hdrs[i].begin(), hdrs[i].end());
    }
    std::ofstream f(path, std::ios::binary | std::ios::trunc);
    if (!f) throw Error("cannot write " + path);
    f.write(reinterpret_cast<const char*>(out.data()), (std::streamsize)out.size());
    uint64_t pos = out.size();
    for (size_t i = 0; i < ds.size(); ++i) {
      const uint64_t bytes = ds[i].rows * ds[i].cols * 8;
      if (!bytes) continue;
      for (; pos < a_data[i]; ++pos) f.put(0);
      f.write(reinterpret_cast<const char*>(ds[i].data), (std::streamsize)bytes);
      pos += bytes;
    }
    if (!f) throw Error("short write on " + path);
  }

 private:
  static uint64_t align8(uint64_t x) { return (x + 7) & ~7ull; }
  static void append(std::vector<uint8_t>& v, uint64_t x, int n) {
    for (int i = 0; i < n; ++i) v.push_back((uint8_t)(x >> (8 * i)));
  }
  static void put(std::vector<uint8_t>& v, size_t at, uint64_t x, int n) {
    for (int i = 0; i < n; ++i) v[at + i] = (uint8_t)(x >> (8 * i));
  }
  // version-1 message: type, size (multiple of 8), flags, 3 reserved bytes, data
  static void message(std::vector<uint8_t>& v, uint16_t type, std::vector<uint8_t> data, uint8_t flags = 0) {
    while (data.size() % 8) data.push_back(0);
    append(v, type, 2);
    append(v, data.size(), 2);
    v.push_back(flags);
    v.insert(v.end(), {0, 0, 0});
    v.insert(v.end(), data.begin(), data.end());
  }
  static uint32_t count_messages(const std::vector<uint8_t>& msgs) {
    uint32_t n = 0;
    for (size_t q = 0; q + 8 <= msgs.size(); ++n) q += 8 + (msgs[q + 2] | (msgs[q + 3] << 8));
    return n;
  }
  static void object_header_prefix(std::vector<uint8_t>& v, uint32_t n_msgs, uint64_t bytes) {
    v.insert(v.end(), {1, 0});
    append(v, n_msgs, 2);
    append(v, 1, 4);      // object reference count
    append(v, bytes, 4);  // object header size
    append(v, 0, 4);      // pad the prefix to 16 bytes: messages are 8-byte aligned
  }
  static std::vector<uint8_t> pad8(std::vector<uint8_t> v) {
    while (v.size() % 8) v.push_back(0);
    return v;
  }
  static std::vector<uint8_t> scalar_space() { return {1, 0, 0, 0, 0, 0, 0, 0}; }
  static std::vector<uint8_t> f64_type() {
    std::vector<uint8_t> t{0x11, 0x20, 0x3F, 0x00};  // class 1 (floating point) version 1; LE, implied mantissa msb; sign at bit 63
    append(t, 8, 4);
    append(t, 0, 2);    // bit offset
    append(t, 64, 2);   // precision
    t.insert(t.end(), {52, 11, 0, 52});  // exponent location, exponent size, mantissa location, mantissa size
    append(t, 1023, 4);
    return t;
  }
  static std::vector<uint8_t> vlen_utf8_type() {
    std::vector<uint8_t> t{0x19, 0x01, 0x01, 0x00};  // class 9 (variable length) version 1; type = string, null-terminated, UTF-8
    append(t, 16, 4);
    t.insert(t.end(), {0x13, 0x00, 0x00, 0x00});  // base type: class 3 (string) version 1, null-terminated
    t[t.size() - 3] = 0x10;                        // ... character set UTF-8 (bits 4-7 of the first flag byte)
    append(t, 1, 4);
    return t;
  }
  static std::vector<uint8_t> u64_type() {
    std::vector<uint8_t> t{0x10, 0x00, 0x00, 0x00};  // class 0 (fixed point) version 1; LE, unsigned
    append(t, 8, 4);
    append(t, 0, 2);
    append(t, 64, 2);
    return t;
  }
  static std::vector<uint8_t> root_messages(uint64_t a_btree, uint64_t a_heap, const std::vector<std::pair<std::string, Attr>>& attrs,
                                            const std::vector<uint32_t>& str_index, uint64_t a_gcol) {
    std::vector<uint8_t> v, st;
    append(st, a_btree, 8);
    append(st, a_heap, 8);
    message(v, 0x0011, st);
    for (size_t i = 0; i < attrs.size(); ++i) {
      const std::string& name = attrs[i].first;
      const Attr& a = attrs[i].second;
      const std::vector<uint8_t> type = a.kind == Attr::F64 ? f64_type() : vlen_utf8_type(), space = scalar_space();
      std::vector<uint8_t> m{1, 0};
      append(m, name.size() + 1, 2);
      append(m, type.size(), 2);
      append(m, space.size(), 2);
      std::vector<uint8_t> nm(name.begin(), name.end());
      nm.push_back(0);
      for (const std::vector<uint8_t>& part : {pad8(nm), pad8(type), pad8(space)}) m.insert(m.end(), part.begin(), part.end());
      if (a.kind == Attr::F64) {
        uint64_t bits;
        std::memcpy(&bits, &a.f, 8);
        append(m, bits, 8);
      } else {
        append(m, a.s.size(), 4);
        append(m, a_gcol, 8);
        append(m, str_index[i], 4);
      }
      message(v, 0x000C, m);
    }
    return v;
  }
  static std::vector<uint8_t> dataset_header(const OutDataset& d, uint64_t a_data) {
    std::vector<uint8_t> v, space{1, 2, 0, 0, 0, 0, 0, 0};
    append(space, d.rows, 8);
    append(space, d.cols, 8);
    message(v, 0x0001, space);
    message(v, 0x0003, u64_type(), 1);  // constant message
    message(v, 0x0005, {2, 2, 2, 1, 0, 0, 0, 0});  // fill value v2 as libhdf5 writes it by default: allocate late, write if set, defined with size 0
    std::vector<uint8_t> layout{3, 1};  // version 3, contiguous
    append(layout, a_data, 8);
    append(layout, d.rows * d.cols * 8, 8);
    message(v, 0x0008, layout);
    std::vector<uint8_t> out;
    object_header_prefix(out, count_messages(v), v.size());
    out.insert(out.end(), v.begin(), v.end());
    return out;
  }
};

}  // namespace h5min
