// api.cu -- the extern "C" surface declared in include/hnsw_b200.h.
// Every entry point converts exceptions into hb_status codes + a thread-local
// message; nothing throws across the boundary.
#include <atomic>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "hb_internal.h"

#include <mutex>
#include <thread>
#include <vector>

namespace hb {
namespace {
thread_local std::string g_last_error;
}
void set_last_error(const std::string& m) { g_last_error = m; }
namespace {
std::atomic<uint64_t> g_launches{0};
}
void count_launches(uint64_t n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

namespace {

template <class T>
struct DevMem {
  T* p = nullptr;
  explicit DevMem(size_t n) { HB_CUDA(cudaMalloc(&p, (n ? n : 1) * sizeof(T))); }
  ~DevMem() { cudaFree(p); }
  DevMem(const DevMem&) = delete;
  DevMem& operator=(const DevMem&) = delete;
};

// Pageable host memory -> device through page-locked staging slots filled by several host threads.  The driver stages a
// pageable cudaMemcpy on one thread (~11 GB/s measured on the B200 box); T threads that copy 4 MB pieces into their
// own slots and issue their own DMAs keep the PCIe link busy.  One instance per device, kept for the process lifetime
// (pinning the 64 MB of slots costs more than a small upload).
class StagedUploader {
 public:
  static constexpr size_t kPiece = size_t(4) << 20;
  explicit StagedUploader(int device) : device_(device) {
    const unsigned hw = std::thread::hardware_concurrency();
    workers_ = (int)std::min(8u, std::max(1u, hw));
    HB_CUDA(cudaHostAlloc(&slots_, (size_t)workers_ * 2 * kPiece, cudaHostAllocDefault));
    streams_.resize(workers_);
    slot_ev_.resize((size_t)workers_ * 2);
    done_.resize(workers_);
    for (int t = 0; t < workers_; ++t) {
      HB_CUDA(cudaStreamCreateWithFlags(&streams_[t], cudaStreamNonBlocking));
      HB_CUDA(cudaEventCreateWithFlags(&done_[t], cudaEventDisableTiming));
      for (int s = 0; s < 2; ++s) HB_CUDA(cudaEventCreateWithFlags(&slot_ev_[2 * t + s], cudaEventDisableTiming));
    }
  }
  // Copies [h_src, h_src + bytes) to d_dst and makes `consumer` wait for it.  Returns when every piece has been issued.
  void upload(void* d_dst, const void* h_src, size_t bytes, cudaStream_t consumer) {
    std::lock_guard<std::mutex> lock(mu_);
    std::atomic<int> err{(int)cudaSuccess};
    const size_t part = ((bytes + workers_ - 1) / workers_ + 127) & ~size_t(127);
    std::vector<std::thread> th;
    for (int t = 0; t < workers_; ++t) {
      const size_t lo = std::min(bytes, part * t), hi = std::min(bytes, part * (t + 1));
      th.emplace_back([this, t, lo, hi, d_dst, h_src, &err] {
        auto ck = [&err](cudaError_t e) {
          if (e != cudaSuccess) {
            int ok = (int)cudaSuccess;
            err.compare_exchange_strong(ok, (int)e);
          }
        };
        ck(cudaSetDevice(device_));
        int slot = 0;
        for (size_t off = lo; off < hi; off += kPiece, slot ^= 1) {
          const size_t len = std::min(kPiece, hi - off);
          unsigned char* stage = static_cast<unsigned char*>(slots_) + ((size_t)2 * t + slot) * kPiece;
          ck(cudaEventSynchronize(slot_ev_[2 * t + slot]));  // the DMA that last read this slot (no-op when never recorded)
          std::memcpy(stage, static_cast<const unsigned char*>(h_src) + off, len);
          ck(cudaMemcpyAsync(static_cast<unsigned char*>(d_dst) + off, stage, len, cudaMemcpyHostToDevice, streams_[t]));
          ck(cudaEventRecord(slot_ev_[2 * t + slot], streams_[t]));
        }
        ck(cudaEventRecord(done_[t], streams_[t]));
      });
    }
    for (auto& x : th) x.join();
    HB_CUDA((cudaError_t)err.load());
    for (int t = 0; t < workers_; ++t) HB_CUDA(cudaStreamWaitEvent(consumer, done_[t], 0));
  }
  static StagedUploader& for_device(int device) {
    static std::mutex mu;
    static std::vector<std::unique_ptr<StagedUploader>> all;
    std::lock_guard<std::mutex> lock(mu);
    if ((size_t)device >= all.size()) all.resize(device + 1);
    if (!all[device]) all[device].reset(new StagedUploader(device));
    return *all[device];
  }

 private:
  int device_, workers_ = 1;
  void* slots_ = nullptr;
  std::vector<cudaStream_t> streams_;
  std::vector<cudaEvent_t> slot_ev_, done_;
  std::mutex mu_;
};

Index* as_index(const hb_index* h) {
  HB_REQUIRE(h != nullptr, HB_ERR_INVALID, "null index handle");
  return reinterpret_cast<Index*>(const_cast<hb_index*>(h));
}

}  // namespace
}  // namespace hb

using namespace hb;

extern "C" {

const char* hb_last_error(void) { return g_last_error.c_str(); }
int hb_version(void) { return HB_VERSION; }

uint64_t hb_kernel_launches(void) { return g_launches.load(std::memory_order_relaxed); }
double hb_index_last_search_ms(const hb_index* h) { return h ? reinterpret_cast<const Index*>(h)->last_search_ms : 0.0; }

int hb_device_count(int* count) {
  return guarded([&] {
    HB_REQUIRE(count != nullptr, HB_ERR_INVALID, "null argument");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess) {
      cudaGetLastError();
      n = 0;
    }
    *count = n;
  });
}

int hb_index_load_idx(const char* path, int device, int keep_host, hb_index** out) {
  return guarded([&] {
    HB_REQUIRE(path && out, HB_ERR_INVALID, "null argument");
    *out = nullptr;
    ensure_device(device);  // fail before parsing gigabytes when there is no GPU
    Attrs attrs;
    int algo = 0;
    int64_t ep = -1;
    std::vector<HostGraph> graphs;
    read_idx_file(path, attrs, algo, graphs, ep);
    Index* idx = index_from_host(algo, std::move(graphs), ep, device, keep_host != 0);
    const uint64_t size = idx->attrs.size;
    idx->attrs = attrs;
    idx->attrs.size = attrs.size ? attrs.size : size;
    *out = reinterpret_cast<hb_index*>(idx);
  });
}

int hb_index_from_parts(int algo, uint32_t n_graphs, const uint64_t* n_nodes, const uint64_t* const* sketches,
                        const uint64_t* const* down, const uint64_t* const* off, const void* const* ids, int id_bytes,
                        int64_t entry_point, int device, int keep_host, hb_index** out) {
  return guarded([&] {
    HB_REQUIRE(out && n_nodes && sketches, HB_ERR_INVALID, "null argument");
    *out = nullptr;
    HB_REQUIRE(n_graphs >= 1 && n_graphs <= 64, HB_ERR_INVALID, "n_graphs out of range");
    HB_REQUIRE(id_bytes == 4 || id_bytes == 8, HB_ERR_INVALID, "id_bytes must be 4 or 8");
    ensure_device(device);
    std::vector<HostGraph> graphs(n_graphs);
    for (uint32_t g = 0; g < n_graphs; ++g) {
      HostGraph& hg = graphs[g];
      hg.n = n_nodes[g];
      HB_REQUIRE(hg.n <= 0xFFFFFFFEull, HB_ERR_UNSUPPORTED, "more than 2^32-2 nodes in one graph");
      HB_REQUIRE(hg.n == 0 || sketches[g], HB_ERR_INVALID, "null sketch array");
      hg.sk.assign(sketches[g], sketches[g] + hg.n * 16);
      if (algo == HB_ALGO_BRUTEFORCE) {
        hg.off.assign(hg.n + 1, 0);
        continue;
      }
      HB_REQUIRE(off && off[g] && ids, HB_ERR_INVALID, "null adjacency arrays");
      hg.off.assign(off[g], off[g] + hg.n + 1);
      const uint64_t ne = hg.off[hg.n];
      HB_REQUIRE(ne == 0 || ids[g], HB_ERR_INVALID, "null id array");
      hg.ids.resize(ne);
      if (id_bytes == 4) {
        std::memcpy(hg.ids.data(), ids[g], ne * 4);
      } else {
        const uint64_t* wide = static_cast<const uint64_t*>(ids[g]);
        for (uint64_t j = 0; j < ne; ++j) {
          HB_REQUIRE(wide[j] < hg.n, HB_ERR_FORMAT, "neighbour id out of range");
          hg.ids[j] = (uint32_t)wide[j];
        }
      }
      if (g > 0) {
        HB_REQUIRE(down && down[g], HB_ERR_INVALID, "null down-link array");
        hg.down.assign(down[g], down[g] + hg.n);
      }
    }
    *out = reinterpret_cast<hb_index*>(index_from_host(algo, std::move(graphs), entry_point, device, keep_host != 0));
  });
}

int hb_index_save_idx(const hb_index* h, const char* path) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(path, HB_ERR_INVALID, "null path");
    HB_REQUIRE(idx->has_host, HB_ERR_INVALID, "index was created without keep_host; nothing to serialise");
    write_idx_file(path, idx->attrs, idx->algo, idx->host, idx->ep_file);
  });
}

int hb_index_get_attrs(const hb_index* h, int* format_size, uint64_t* size, double* buildtime, double* querytime, char* data,
                       uint64_t data_cap, char* params, uint64_t params_cap) {
  return guarded([&] {
    Index* idx = as_index(h);
    if (format_size) *format_size = idx->attrs.format_size ? 1 : 0;
    if (size) *size = idx->attrs.size ? idx->attrs.size : idx->n_base;
    if (buildtime) *buildtime = idx->attrs.buildtime;
    if (querytime) *querytime = idx->attrs.querytime;
    auto copy = [](const std::string& s, char* dst, uint64_t cap) {
      if (!dst || cap == 0) return;
      const size_t n = std::min<size_t>(s.size(), (size_t)cap - 1);
      std::memcpy(dst, s.data(), n);
      dst[n] = 0;
    };
    copy(idx->attrs.data, data, data_cap);
    copy(idx->attrs.params, params, params_cap);
  });
}

int hb_index_set_attrs(hb_index* h, double buildtime, const char* params) {
  return guarded([&] {
    Index* idx = as_index(h);
    idx->attrs.buildtime = buildtime;
    if (params) idx->attrs.params = params;
  });
}

int hb_index_compact(hb_index* h) {
  return guarded([&] { index_compact(*as_index(h)); });
}

int hb_index_set_format_size(hb_index* h, int format_size) {
  return guarded([&] { as_index(h)->attrs.format_size = format_size != 0; });
}

void hb_index_free(hb_index* h) {
  DeviceRestore restore;
  delete reinterpret_cast<Index*>(h);
}

uint64_t hb_index_size(const hb_index* h) { return h ? reinterpret_cast<const Index*>(h)->n_base : 0; }

int hb_index_get_info(const hb_index* h, hb_index_info* info) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(info, HB_ERR_INVALID, "null argument");
    info->algo = idx->algo;
    info->device = idx->device;
    info->n_graphs = idx->algo == HB_ALGO_BRUTEFORCE ? 0u : (uint32_t)idx->g_nodes.size();
    info->row_stride = idx->d_rowoff ? 0 : idx->stride;
    info->size = idx->n_base;
    info->n_nodes_total = idx->n_total;
    uint64_t e = 0;
    for (uint64_t x : idx->g_edges) e += x;
    info->n_edges_total = e;
    info->entry_point = idx->ep_file;
    info->device_bytes = idx->device_bytes;
    info->build_prunes = idx->build_prunes;
    info->build_unlinked = idx->build_unlinked;
    info->build_dropped_appends = idx->build_dropped_appends;
  });
}

int hb_index_graph_dims(const hb_index* h, uint32_t graph, uint64_t* n_nodes, uint64_t* n_edges) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(graph < idx->g_nodes.size(), HB_ERR_INVALID, "graph index out of range");
    if (n_nodes) *n_nodes = idx->g_nodes[graph];
    if (n_edges) *n_edges = idx->g_edges[graph];
  });
}

int hb_index_graph_export(const hb_index* h, uint32_t graph, uint64_t* sketches, uint64_t* down, uint64_t* off, uint64_t* ids) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(idx->has_host, HB_ERR_INVALID, "index was created without keep_host");
    HB_REQUIRE(graph < idx->host.size(), HB_ERR_INVALID, "graph index out of range");
    const HostGraph& g = idx->host[graph];
    if (sketches) std::memcpy(sketches, g.sk.data(), g.sk.size() * 8);
    if (down) for (size_t i = 0; i < g.down.size(); ++i) down[i] = g.down[i];
    if (off) std::memcpy(off, g.off.data(), g.off.size() * 8);
    if (ids) for (size_t i = 0; i < g.ids.size(); ++i) ids[i] = g.ids[i];
  });
}

int hb_knns_device(const hb_index* h, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                   void* d_dists, void* d_found, void* d_counters, void* stream) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(nq == 0 || (d_queries && d_keys && d_dists && d_found), HB_ERR_INVALID, "null device buffer");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (idx->algo == HB_ALGO_BRUTEFORCE) {
      // Indexes::Bruteforce: exact top-k; `ef` is ignored (bruteforce.rs:38).  Sorted (distance, key) order.
      HB_REQUIRE(k >= 1, HB_ERR_INVALID, "k out of range");
      HB_CUDA(cudaSetDevice(idx->device));
      if (nq == 0) return;
      uint64_t* packed = nullptr;
      HB_CUDA(cudaMallocAsync(&packed, nq * k * 8, s));
      try {
        ground_truth_device(idx->d_sk, idx->n_base, 0, d_queries, nq, k, packed, s);
        unpack_device(packed, nq * k, d_keys, d_dists, s);
      } catch (...) {
        cudaFreeAsync(packed, s);
        throw;
      }
      cudaFreeAsync(packed, s);
      const uint32_t nres = (uint32_t)std::min<uint64_t>(k, idx->n_base);
      std::vector<uint32_t> f(nq, nres);
      HB_CUDA(cudaMemcpyAsync(d_found, f.data(), nq * 4, cudaMemcpyHostToDevice, s));
      if (d_counters) HB_CUDA(cudaMemsetAsync(d_counters, 0, nq * 16, s));
      HB_CUDA(cudaStreamSynchronize(s));
      return;
    }
    knns_device(*idx, d_queries, nq, k, ef, flags, d_keys, d_dists, d_found, d_counters, s);
  });
}

int hb_knns_device_noalloc(const hb_index* h, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                           void* d_dists, void* d_found, void* stream) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(nq == 0 || (d_queries && d_keys && d_dists && d_found), HB_ERR_INVALID, "null device buffer");
    knns_device_noalloc(*idx, d_queries, nq, k, ef, flags, d_keys, d_dists, d_found, static_cast<cudaStream_t>(stream));
  });
}

#ifndef HB_HOST_COPY_THREADS_DEFAULT
#define HB_HOST_COPY_THREADS_DEFAULT 4u
#endif
#ifndef HB_HOST_STAGING_DEFAULT
#define HB_HOST_STAGING_DEFAULT 1  // HB_HOST_STAGING=0 restores the cudaMemcpyAsync bracket (profiles/r02aa_host_staging_threads.log)
#endif

namespace hb {
namespace {
constexpr uint64_t kHostStagingBytes = 64ull << 20;
// take an idle staging set from the index's pool (or make one) and grow it to the call's size
CallBuffers acquire_buffers(const Index& idx, uint64_t nq, uint32_t k, bool counters) {
  CallBuffers b;
  {
    std::lock_guard<std::mutex> lock(idx.pool_mu);
    if (!idx.pool.empty()) { b = idx.pool.back(); idx.pool.pop_back(); }
  }
  if (!b.stream) HB_CUDA(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
  if (b.cap_q < nq) {
    cudaFree(b.q); cudaFree(b.found); cudaFree(b.ctr);
    b.q = nullptr; b.found = nullptr; b.ctr = nullptr; b.cap_q = 0;
    HB_CUDA(cudaMalloc(&b.q, nq * 128));
    HB_CUDA(cudaMalloc(&b.found, nq * 4));
    HB_CUDA(cudaMalloc(&b.ctr, nq * sizeof(hb_counters)));
    b.cap_q = nq;
  }
  if (b.cap_out < nq * k) {
    cudaFree(b.keys); cudaFree(b.dists);
    b.keys = b.dists = nullptr; b.cap_out = 0;
    HB_CUDA(cudaMalloc(&b.keys, nq * k * 8));
    HB_CUDA(cudaMalloc(&b.dists, nq * k * 8));
    b.cap_out = nq * k;
  }
  (void)counters;
  return b;
}
// grow the set's page-locked host staging (hb_knns with pageable caller buffers)
void ensure_host_staging(CallBuffers& b, uint64_t nq, uint32_t k) {
  if (b.hcap_q < nq) {
    cudaFreeHost(b.hq); cudaFreeHost(b.hfound);
    b.hq = nullptr; b.hfound = nullptr; b.hcap_q = 0;
    HB_CUDA(cudaHostAlloc(&b.hq, nq * 128, cudaHostAllocMapped | cudaHostAllocPortable));
    HB_CUDA(cudaHostAlloc(&b.hfound, nq * 4, cudaHostAllocMapped | cudaHostAllocPortable));
    b.hcap_q = nq;
  }
  if (b.hcap_out < nq * k) {
    cudaFreeHost(b.hkeys); cudaFreeHost(b.hdists);
    b.hkeys = b.hdists = nullptr; b.hcap_out = 0;
    HB_CUDA(cudaHostAlloc(&b.hkeys, nq * k * 8, cudaHostAllocMapped | cudaHostAllocPortable));
    HB_CUDA(cudaHostAlloc(&b.hdists, nq * k * 8, cudaHostAllocMapped | cudaHostAllocPortable));
    b.hcap_out = nq * k;
  }
}

// A few persistent helper threads for host-side copies between pageable caller memory and page-locked staging: one
// thread moves ~12 GB/s, and the 3 MB a 10 k-query call touches are a fifth of its end-to-end time.
class HostCopyPool {
 public:
  struct Job { void* dst; const void* src; size_t bytes; };
  static HostCopyPool& get() {
    static HostCopyPool* pool = new HostCopyPool();  // never destroyed: no join at process exit
    return *pool;
  }
  // copies every job, split into pieces over the helpers and the calling thread; returns when all bytes have landed
  void copy(const Job* jobs, int n_jobs) {
    size_t total = 0;
    for (int i = 0; i < n_jobs; ++i) total += jobs[i].bytes;
    if (helpers_.empty() || total < (size_t(256) << 10)) {
      for (int i = 0; i < n_jobs; ++i) std::memcpy(jobs[i].dst, jobs[i].src, jobs[i].bytes);
      return;
    }
    std::lock_guard<std::mutex> call_lock(call_mu_);  // one copy at a time: concurrent callers queue here
    const size_t piece = std::max<size_t>(size_t(64) << 10, (total / ((helpers_.size() + 1) * 4) + 4095) & ~size_t(4095));
    size_t n_pieces = 0;
    for (;;) {
      // helpers enter work() only after raising active_ under mu_: with mu_ held and active_ == 0 none of them can be
      // reading the piece list or the counters, and none can start before the new generation is published
      std::unique_lock<std::mutex> lock(mu_);
      if (active_.load(std::memory_order_acquire) != 0) {
        lock.unlock();
        std::this_thread::yield();
        continue;
      }
      pieces_.clear();
      for (int i = 0; i < n_jobs; ++i)
        for (size_t off = 0; off < jobs[i].bytes; off += piece)
          pieces_.push_back({static_cast<unsigned char*>(jobs[i].dst) + off, static_cast<const unsigned char*>(jobs[i].src) + off,
                             std::min(piece, jobs[i].bytes - off)});
      n_pieces = pieces_.size();
      next_.store(0, std::memory_order_relaxed);
      done_.store(0, std::memory_order_relaxed);
      ++generation_;
      break;
    }
    cv_.notify_all();
    work();
    while (done_.load(std::memory_order_acquire) < n_pieces) std::this_thread::yield();
  }

 private:
  HostCopyPool() {
    const unsigned hw = std::thread::hardware_concurrency();
    const char* e = std::getenv("HB_HOST_COPY_THREADS");  // threads per copy including the caller (default 4)
    const unsigned want = e ? (unsigned)std::max(1, std::atoi(e)) : HB_HOST_COPY_THREADS_DEFAULT;
    const int n = (int)std::min(want - 1, hw > 1 ? hw - 1 : 0u);
    for (int i = 0; i < n; ++i) helpers_.emplace_back([this] { loop(); });
    for (auto& t : helpers_) t.detach();
  }
  void work() {
    for (;;) {
      const size_t i = next_.fetch_add(1, std::memory_order_relaxed);
      if (i >= pieces_.size()) break;
      std::memcpy(pieces_[i].dst, pieces_[i].src, pieces_[i].bytes);
      done_.fetch_add(1, std::memory_order_release);
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> lock(mu_);
        cv_.wait(lock, [&] { return generation_ != seen; });
        seen = generation_;
        active_.fetch_add(1, std::memory_order_acq_rel);  // under mu_: ordered after the caller's piece list
      }
      work();
      active_.fetch_sub(1, std::memory_order_acq_rel);
    }
  }
  std::vector<std::thread> helpers_;
  std::mutex call_mu_, mu_;
  std::condition_variable cv_;
  uint64_t generation_ = 0;
  std::vector<Job> pieces_;
  std::atomic<size_t> next_{0}, done_{0};
  std::atomic<int> active_{0};
};

void release_buffers(const Index& idx, const CallBuffers& b) {
  std::lock_guard<std::mutex> lock(idx.pool_mu);
  idx.pool.push_back(b);
}
}  // namespace
}  // namespace hb

int hb_knns(const hb_index* h, const uint64_t* queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, uint64_t* keys,
            uint64_t* dists, uint32_t* found, hb_counters* counters) {
  return guarded([&] {
    Index* idx = as_index(h);
    HB_REQUIRE(nq == 0 || (queries && keys && dists), HB_ERR_INVALID, "null host buffer");
    HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
    if (nq == 0) return;
    HB_CUDA(cudaSetDevice(idx->device));
    // Pinned (page-locked, mapped) caller buffers are used in place: the kernel reads each query's 128 bytes and writes
    // its k results across PCIe itself, overlapped with the other warps' searches, so no staging copy brackets the
    // launch.  hb_knns_device synchronises the stream before returning, which makes the writes visible to the host.
    if (idx->algo != HB_ALGO_BRUTEFORCE && found && !counters) {
      auto mapped = [](const void* p) -> void* {
        cudaPointerAttributes a{};
        if (cudaPointerGetAttributes(&a, p) != cudaSuccess) {
          cudaGetLastError();
          return nullptr;
        }
        return a.type == cudaMemoryTypeHost ? a.devicePointer : nullptr;
      };
      void *dq = mapped(queries), *dk = mapped(keys), *dd = mapped(dists), *df = mapped(found);
      if (dq && dk && dd && df) {
        CallBuffers b = acquire_buffers(*idx, 1, k, false);  // only its stream is used
        struct Release0 {
          const Index& i; const CallBuffers& b;
          ~Release0() { release_buffers(i, b); }
        } release0{*idx, b};
        int rc = hb_knns_device(h, dq, nq, k, ef, flags, dk, dd, df, nullptr, b.stream);
        if (rc != HB_OK) throw Error(rc, g_last_error);
        if (flags & HB_PAD_LIKE_CLI)
          for (uint64_t i = 0; i < nq; ++i)
            HB_REQUIRE(found[i] > 0, HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here, main.rs:472)");
        return;
      }
    }
    // Pageable caller buffers, batches whose staging stays modest: the same in-place scheme through the call's own
    // page-locked staging -- a threaded host copy in, the kernel reading and writing across PCIe, a threaded copy out;
    // no DMA brackets the launch.
    static const bool host_staging = [] { const char* e = std::getenv("HB_HOST_STAGING"); return (e ? std::atoi(e) : HB_HOST_STAGING_DEFAULT) != 0; }();
    if (host_staging && idx->algo != HB_ALGO_BRUTEFORCE && !counters && nq * (128 + 4 + 16ull * k) <= kHostStagingBytes) {
      CallBuffers b = acquire_buffers(*idx, 1, k, false);
      struct Release1 {
        const Index& i; CallBuffers& b;
        ~Release1() { release_buffers(i, b); }
      } release1{*idx, b};
      ensure_host_staging(b, nq, k);
      HostCopyPool& pool = HostCopyPool::get();
      const HostCopyPool::Job in{b.hq, queries, nq * 128};
      pool.copy(&in, 1);
      int rc = hb_knns_device(h, b.hq, nq, k, ef, flags, b.hkeys, b.hdists, b.hfound, nullptr, b.stream);
      if (rc != HB_OK) throw Error(rc, g_last_error);
      const HostCopyPool::Job out[3] = {{keys, b.hkeys, nq * k * 8}, {dists, b.hdists, nq * k * 8}, {found, b.hfound, found ? nq * 4 : 0}};
      pool.copy(out, 3);
      if (flags & HB_PAD_LIKE_CLI)
        for (uint64_t i = 0; i < nq; ++i)
          HB_REQUIRE(b.hfound[i] > 0, HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here, main.rs:472)");
      return;
    }
    CallBuffers b = acquire_buffers(*idx, nq, k, counters != nullptr);
    struct Release {
      const Index& i; const CallBuffers& b;
      ~Release() { release_buffers(i, b); }
    } release{*idx, b};
    // pinned caller memory makes these true asynchronous DMA transfers; pageable memory is staged by the driver
    HB_CUDA(cudaMemcpyAsync(b.q, queries, nq * 128, cudaMemcpyHostToDevice, b.stream));
    int rc = hb_knns_device(h, b.q, nq, k, ef, flags, b.keys, b.dists, b.found, counters ? (void*)b.ctr : nullptr, b.stream);
    if (rc != HB_OK) throw Error(rc, g_last_error);
    HB_CUDA(cudaMemcpyAsync(keys, b.keys, nq * k * 8, cudaMemcpyDeviceToHost, b.stream));
    HB_CUDA(cudaMemcpyAsync(dists, b.dists, nq * k * 8, cudaMemcpyDeviceToHost, b.stream));
    std::vector<uint32_t> f(nq);
    HB_CUDA(cudaMemcpyAsync(f.data(), b.found, nq * 4, cudaMemcpyDeviceToHost, b.stream));
    if (counters) HB_CUDA(cudaMemcpyAsync(counters, b.ctr, nq * sizeof(hb_counters), cudaMemcpyDeviceToHost, b.stream));
    HB_CUDA(cudaStreamSynchronize(b.stream));
    if (found) std::memcpy(found, f.data(), nq * 4);
    if (idx->algo == HB_ALGO_BRUTEFORCE) {
      // rows past found are UINT64_MAX from the packed form: zero them / pad like the graph path
      for (uint64_t i = 0; i < nq; ++i)
        for (uint32_t j = f[i]; j < k; ++j) {
          const bool pad = (flags & HB_PAD_LIKE_CLI) && f[i] > 0;
          keys[i * k + j] = pad ? keys[i * k] : 0;
          dists[i * k + j] = pad ? dists[i * k] : 0;
        }
    }
    if (flags & HB_PAD_LIKE_CLI)
      for (uint64_t i = 0; i < nq; ++i)
        HB_REQUIRE(f[i] > 0, HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here, main.rs:472)");
  });
}

int hb_search(const hb_index* h, const uint64_t* query, uint32_t k, uint32_t ef, uint32_t flags, uint64_t* keys, uint64_t* dists,
              uint32_t* found) {
  return hb_knns(h, query, 1, k, ef, flags, keys, dists, found, nullptr);
}

int hb_ground_truth_device_algo(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                                int algo, void* d_packed, void* stream) {
  return guarded([&] {
    HB_REQUIRE(nq == 0 || (d_queries && d_packed), HB_ERR_INVALID, "null device buffer");
    HB_REQUIRE(n == 0 || d_data, HB_ERR_INVALID, "null device buffer");
    ground_truth_device(d_data, n, key_offset, d_queries, nq, k, d_packed, static_cast<cudaStream_t>(stream), algo);
  });
}

int hb_ground_truth_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                           void* d_packed, void* stream) {
  return hb_ground_truth_device_algo(d_data, n, key_offset, d_queries, nq, k, HB_GT_AUTO, d_packed, stream);
}

int hb_topk_merge_device(const void* d_lists, uint32_t n_lists, uint64_t nq, uint32_t k, void* d_packed_out, void* stream) {
  return guarded([&] {
    HB_REQUIRE(nq == 0 || (d_lists && d_packed_out), HB_ERR_INVALID, "null device buffer");
    topk_merge_device(d_lists, n_lists, nq, k, d_packed_out, static_cast<cudaStream_t>(stream));
  });
}

int hb_unpack_device(const void* d_packed, uint64_t n, void* d_keys, void* d_dists, void* stream) {
  return guarded([&] {
    HB_REQUIRE(n == 0 || (d_packed && d_keys && d_dists), HB_ERR_INVALID, "null device buffer");
    unpack_device(d_packed, n, d_keys, d_dists, static_cast<cudaStream_t>(stream));
  });
}

int hb_pack_device(const void* d_keys, const void* d_dists, const void* d_found, uint64_t nq, uint32_t k, uint64_t key_offset,
                   void* d_packed, void* stream) {
  return guarded([&] {
    HB_REQUIRE(nq == 0 || (d_keys && d_dists && d_found && d_packed), HB_ERR_INVALID, "null device buffer");
    HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
    pack_device(d_keys, d_dists, d_found, nq, k, key_offset, d_packed, static_cast<cudaStream_t>(stream));
  });
}

int hb_ground_truth(const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k, int device, uint64_t* keys,
                    uint64_t* dists) {
  return hb_ground_truth_algo(data, n, queries, nq, k, device, HB_GT_AUTO, keys, dists);
}

int hb_ground_truth_algo(const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k, int device, int algo,
                         uint64_t* keys, uint64_t* dists) {
  return guarded([&] {
    HB_REQUIRE(nq == 0 || (queries && keys && dists), HB_ERR_INVALID, "null host buffer");
    HB_REQUIRE(n == 0 || data, HB_ERR_INVALID, "null host buffer");
    ensure_device(device);
    if (nq == 0) return;
    DevMem<uint64_t> dd(n * 16), dq(nq * 16), dp(nq * k), dk(nq * k), ds(nq * k);
    HB_CUDA(cudaMemcpy(dq.p, queries, nq * 128, cudaMemcpyHostToDevice));
    HB_CUDA(cudaDeviceSynchronize());  // the kernels below run on non-blocking streams, which the NULL stream does not order
    // Large inputs: the rows go up in slices (host threads staging through page-locked slots, StagedUploader) while the
    // previous slice is searched (the upload of 10 M rows takes longer than searching them); every slice yields an ascending list per query with global
    // keys, and the exact merge kernel combines them.  Small inputs: one copy, one search.
#ifndef HB_GT_SLICE_LOG2
#define HB_GT_SLICE_LOG2 21  // 2^21 rows = 256 MB of sketches per slice
#endif
    constexpr uint64_t kSliceRows = 1ull << HB_GT_SLICE_LOG2;
    const uint64_t n_slices = (n + kSliceRows - 1) / kSliceRows;
    if (n_slices >= 2 && nq >= 32 && k <= kSliceRows) {
      struct Streams {
        cudaStream_t comp = nullptr;
        ~Streams() {
          if (comp) cudaStreamDestroy(comp);
        }
      } st;
      HB_CUDA(cudaStreamCreateWithFlags(&st.comp, cudaStreamNonBlocking));
      StagedUploader& up = StagedUploader::for_device(device);
      DevMem<uint64_t> parts(n_slices * nq * k);
      DevMem<uint32_t> bound(nq);  // per-query bound on the k-th distance, carried from slice to slice
      HB_CUDA(cudaMemsetAsync(bound.p, 0xFF, nq * 4, st.comp));
      for (uint64_t i = 0; i < n_slices; ++i) {
        const uint64_t row0 = i * kSliceRows, rows = std::min(kSliceRows, n - row0);
        up.upload(dd.p + row0 * 16, data + row0 * 16, rows * 128, st.comp);  // st.comp waits for the slice
        ground_truth_device(dd.p + row0 * 16, rows, row0, dq.p, nq, k, parts.p + i * nq * k, st.comp, algo, bound.p);
      }
      topk_merge_device(parts.p, (uint32_t)n_slices, nq, k, dp.p, st.comp);
      unpack_device(dp.p, nq * k, dk.p, ds.p, st.comp);
      HB_CUDA(cudaStreamSynchronize(st.comp));
    } else {
      if (n) HB_CUDA(cudaMemcpy(dd.p, data, n * 128, cudaMemcpyHostToDevice));
      ground_truth_device(dd.p, n, 0, dq.p, nq, k, dp.p, nullptr, algo);
      unpack_device(dp.p, nq * k, dk.p, ds.p, nullptr);
    }
    HB_CUDA(cudaMemcpy(keys, dk.p, nq * k * 8, cudaMemcpyDeviceToHost));
    HB_CUDA(cudaMemcpy(dists, ds.p, nq * k * 8, cudaMemcpyDeviceToHost));
  });
}

int hb_synth_device(const hb_synth_params* p, uint64_t start_row, uint64_t n_rows, void* d_out, void* stream) {
  return guarded([&] {
    HB_REQUIRE(p && (n_rows == 0 || d_out), HB_ERR_INVALID, "null argument");
    synth_device(*p, start_row, n_rows, d_out, static_cast<cudaStream_t>(stream));
  });
}

int hb_build_hnsw(const uint64_t* data, uint64_t n, const hb_build_params* p, int device, int keep_host, hb_index** out) {
  return guarded([&] {
    HB_REQUIRE(out && p && (n == 0 || data), HB_ERR_INVALID, "null argument");
    *out = nullptr;
    *out = reinterpret_cast<hb_index*>(build_hnsw(data, n, *p, device, keep_host != 0));
  });
}

}  // extern "C"
