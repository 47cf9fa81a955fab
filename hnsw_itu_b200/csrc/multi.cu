// multi.cu -- multi-GPU entry points of the C ABI ("multi-GPU handled inside the call").
//
// The reference is one process on one host; its only scaling hook is the `--start/--len` row sub-range of `index` and
// `ground-truth` (cli/src/main.rs:596-602, 698-704, applied at main.rs:136-140).  Two partitionings follow from it:
//
//   * REPLICATE: every GPU holds the whole index; Index::knns (src/index/mod.rs:28-42) splits the batch into contiguous
//     slices, one per GPU, results land in disjoint slices of the caller's arrays in input order.  Queries are
//     independent: no data-path collective.  The index reaches the other GPUs by ncclBroadcast of the three device
//     arrays (sketches, adjacency rows, down-links) over NVLink -- no host bounce.
//   * SHARD: GPU r holds rows [start_r, start_r + len_r) -- what `-b start_r -l len_r` selects -- searches ITS rows for
//     ALL queries, and the per-shard ascending (distance, key) lists are exchanged with one ncclAllGather of nq*k
//     packed u64 per rank followed by the exact merge kernel.  The order is total (src/index/mod.rs:90-97), so the
//     merged list equals one Bruteforce::search + sort over all rows (bruteforce.rs:37-47, main.rs:738-741).
//
// hb_comm is one rank of a communicator.  It serves both launch styles: one process per GPU (hb_comm_init_rank with a
// unique id the host side distributes, e.g. over torch.distributed / MPI / a file) and one process driving N devices
// (hb_multi, which owns N ranks made by ncclCommInitAll and one worker thread per device).
//
// NCCL is resolved with dlopen at first use (libnccl.so.2; HB_NCCL_LIB overrides): the library has no link-time
// dependency on it and shares whichever NCCL the host process already loaded.
#include <dlfcn.h>
#include <nccl.h>

#include <condition_variable>
#include <cstring>
#include <functional>
#include <thread>

#include "hb_internal.h"

namespace hb {
namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

const NcclApi& nccl() {
  static NcclApi api;
  static std::once_flag once;
  static std::string err;
  std::call_once(once, [] {
    const char* names[] = {std::getenv("HB_NCCL_LIB"), "libnccl.so.2", "libnccl.so"};
    void* h = nullptr;
    for (const char* n : names)
      if (n && (h = dlopen(n, RTLD_NOW | RTLD_LOCAL))) break;
    if (!h) {
      err = std::string("NCCL not found (dlopen libnccl.so.2): ") + (dlerror() ? dlerror() : "");
      return;
    }
    auto sym = [&](const char* n) {
      void* p = dlsym(h, n);
      if (!p && err.empty()) err = std::string("NCCL symbol missing: ") + n;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommInitAll = reinterpret_cast<decltype(api.CommInitAll)>(sym("ncclCommInitAll"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
    api.GetVersion = reinterpret_cast<decltype(api.GetVersion)>(sym("ncclGetVersion"));
  });
  if (!err.empty()) throw Error(HB_ERR_UNSUPPORTED, err);
  return api;
}

#define HB_NCCL(expr)                                                                                      \
  do {                                                                                                     \
    ncclResult_t r_ = (expr);                                                                              \
    if (r_ != ncclSuccess) throw hb::Error(HB_ERR_CUDA, std::string(#expr) + ": " + nccl().GetErrorString(r_)); \
  } while (0)

}  // namespace

// one rank of a communicator, bound to one device
struct Comm {
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0, device = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};  // start, after search, after all-gather, after merge
  double phase_ms[3] = {0, 0, 0};
  void* gather = nullptr;  // (nranks, nq, k) packed lists
  void* local = nullptr;   // (nq, k) this rank's list
  uint64_t gather_cap = 0, local_cap = 0;
  ~Comm() {
    cudaSetDevice(device);
    if (comm) nccl().CommDestroy(comm);
    for (cudaEvent_t e : ev)
      if (e) cudaEventDestroy(e);
    cudaFree(gather);
    cudaFree(local);
  }
  void ensure(uint64_t nq, uint32_t k) {
    HB_CUDA(cudaSetDevice(device));
    for (cudaEvent_t& e : ev)
      if (!e) HB_CUDA(cudaEventCreate(&e));
    const uint64_t need = nq * k;
    if (local_cap < need) {
      cudaFree(local);
      local = nullptr;
      local_cap = 0;
      HB_CUDA(cudaMalloc(&local, std::max<uint64_t>(need, 1) * 8));
      local_cap = need;
    }
    if (gather_cap < need * nranks) {
      cudaFree(gather);
      gather = nullptr;
      gather_cap = 0;
      HB_CUDA(cudaMalloc(&gather, std::max<uint64_t>(need * nranks, 1) * 8));
      gather_cap = need * nranks;
    }
  }
  // local (nq, k) packed ascending list on every rank -> exact global (nq, k) list in d_out on every rank
  void exchange_and_merge(uint64_t nq, uint32_t k, void* d_out, cudaStream_t s) {
    HB_CUDA(cudaEventRecord(ev[1], s));
    if (nranks == 1) {
      HB_CUDA(cudaMemcpyAsync(d_out, local, nq * k * 8, cudaMemcpyDeviceToDevice, s));
      HB_CUDA(cudaEventRecord(ev[2], s));
    } else {
      HB_NCCL(nccl().AllGather(local, gather, nq * k, ncclUint64, comm, s));
      HB_CUDA(cudaEventRecord(ev[2], s));
      topk_merge_device(gather, (uint32_t)nranks, nq, k, d_out, s);
    }
    HB_CUDA(cudaEventRecord(ev[3], s));
  }
  void read_phases() {
    for (int i = 0; i < 3; ++i) {
      float ms = 0.f;
      if (cudaEventElapsedTime(&ms, ev[i], ev[i + 1]) != cudaSuccess) cudaGetLastError();
      phase_ms[i] = ms;
    }
  }
};

Comm* comm_init_rank(const ncclUniqueId& id, int nranks, int rank, int device) {
  HB_REQUIRE(nranks >= 1 && rank >= 0 && rank < nranks, HB_ERR_INVALID, "rank out of range");
  ensure_device(device);
  std::unique_ptr<Comm> c(new Comm());
  c->nranks = nranks;
  c->rank = rank;
  c->device = device;
  if (nranks > 1) HB_NCCL(nccl().CommInitRank(&c->comm, nranks, id, rank));
  return c.release();
}

// `ground-truth` over a row-sharded data set: every rank calls with its rows (first global key = key_offset) and the
// full query batch; d_out receives the global packed (nq, k) result on every rank.
void ground_truth_sharded(Comm& c, const void* d_shard, uint64_t n_local, uint64_t key_offset, const void* d_queries, uint64_t nq,
                          uint32_t k, void* d_out, cudaStream_t s, int algo) {
  if (nq == 0) return;
  c.ensure(nq, k);
  HB_CUDA(cudaEventRecord(c.ev[0], s));
  ground_truth_device(d_shard, n_local, key_offset, d_queries, nq, k, c.local, s, algo);
  c.exchange_and_merge(nq, k, d_out, s);
}

// Sharded HNSW (BASELINE config 5): `idx` was built over this rank's row range only (local keys); every rank searches
// all queries with the same ef, local keys are shifted by key_offset, lists are all-gathered and merged exactly.
void knns_sharded(Comm& c, const Index& idx, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef,
                  void* d_out, cudaStream_t s) {
  if (nq == 0) return;
  HB_REQUIRE(idx.device == c.device, HB_ERR_INVALID, "index and communicator rank live on different devices");
  c.ensure(nq, k);
  uint64_t *keys = nullptr, *dists = nullptr;
  uint32_t* found = nullptr;
  HB_CUDA(cudaMallocAsync(&keys, nq * k * 8, s));
  HB_CUDA(cudaMallocAsync(&dists, nq * k * 8, s));
  HB_CUDA(cudaMallocAsync(&found, nq * 4, s));
  try {
    HB_CUDA(cudaEventRecord(c.ev[0], s));
    knns_device(idx, d_queries, nq, k, ef, 0, keys, dists, found, nullptr, s);
    pack_device(keys, dists, found, nq, k, key_offset, c.local, s);
    c.exchange_and_merge(nq, k, d_out, s);
  } catch (...) {
    cudaFreeAsync(keys, s);
    cudaFreeAsync(dists, s);
    cudaFreeAsync(found, s);
    throw;
  }
  cudaFreeAsync(keys, s);
  cudaFreeAsync(dists, s);
  cudaFreeAsync(found, s);
}

// Replicate the index of rank `root` on every rank: metadata first (one small ncclBroadcast of a POD header), then
// the device arrays GPU-to-GPU.  Returns this rank's index (root gets `src` back).
struct ReplicaHeader {
  uint64_t n_total, n_base, device_bytes, n_graphs, adj_words, compact;
  int64_t ep_file;
  uint32_t stride, ep_gid;
  int32_t algo;
  uint32_t pad;
  uint64_t g_nodes[64], g_edges[64], g_base[64];
};

Index* index_broadcast(Comm& c, Index* src, int root, cudaStream_t s) {
  HB_REQUIRE(root >= 0 && root < c.nranks, HB_ERR_INVALID, "root out of range");
  if (c.nranks == 1) return src;
  HB_CUDA(cudaSetDevice(c.device));
  ReplicaHeader h{};
  if (c.rank == root) {
    HB_REQUIRE(src != nullptr, HB_ERR_INVALID, "the root rank must pass the index to replicate");
    HB_REQUIRE(src->device == c.device, HB_ERR_INVALID, "index and communicator rank live on different devices");
    HB_REQUIRE(src->g_nodes.size() <= 64, HB_ERR_UNSUPPORTED, "more than 64 graphs");
    h.n_total = src->n_total; h.n_base = src->n_base; h.device_bytes = src->device_bytes; h.n_graphs = src->g_nodes.size();
    h.adj_words = src->adj_words; h.compact = src->d_rowoff != nullptr;
    h.ep_file = src->ep_file; h.stride = src->stride; h.ep_gid = src->ep_gid; h.algo = src->algo;
    for (size_t g = 0; g < src->g_nodes.size(); ++g) {
      h.g_nodes[g] = src->g_nodes[g]; h.g_edges[g] = src->g_edges[g]; h.g_base[g] = src->g_base[g];
    }
  }
  ReplicaHeader* d_h = nullptr;
  HB_CUDA(cudaMalloc(&d_h, sizeof h));
  struct FreeH { void* p; ~FreeH() { cudaFree(p); } } free_h{d_h};
  HB_CUDA(cudaMemcpyAsync(d_h, &h, sizeof h, cudaMemcpyHostToDevice, s));
  HB_NCCL(nccl().Broadcast(d_h, d_h, sizeof h, ncclUint8, root, c.comm, s));
  HB_CUDA(cudaMemcpyAsync(&h, d_h, sizeof h, cudaMemcpyDeviceToHost, s));
  HB_CUDA(cudaStreamSynchronize(s));

  std::unique_ptr<Index> made;
  Index* idx = src;
  if (c.rank != root) {
    made.reset(new Index());
    idx = made.get();
    idx->device = c.device;
    idx->algo = h.algo;
    idx->attrs.algo = (uint32_t)h.algo;
    idx->attrs.size = h.n_base;
    idx->ep_file = h.ep_file;
    idx->n_total = h.n_total; idx->n_base = h.n_base; idx->stride = h.stride; idx->ep_gid = h.ep_gid;
    idx->device_bytes = h.device_bytes;
    idx->g_nodes.assign(h.g_nodes, h.g_nodes + h.n_graphs);
    idx->g_edges.assign(h.g_edges, h.g_edges + h.n_graphs);
    idx->g_base.assign(h.g_base, h.g_base + h.n_graphs);
    cudaDeviceProp prop;
    HB_CUDA(cudaGetDeviceProperties(&prop, c.device));
    idx->sm_count = prop.multiProcessorCount;
    HB_CUDA(cudaMalloc(&idx->d_sk, std::max<uint64_t>(h.n_total, 1) * 128));
    if (h.algo != HB_ALGO_BRUTEFORCE) {
      idx->adj_words = h.adj_words;
      HB_CUDA(cudaMalloc(&idx->d_adj, std::max<uint64_t>(h.adj_words * 4ull, 4)));
      if (h.compact) HB_CUDA(cudaMalloc(&idx->d_rowoff, (h.n_total + 1) * 4));
      HB_CUDA(cudaMalloc(&idx->d_down, std::max<uint64_t>(h.n_total - h.n_base, 1) * 4));
    }
  }
  // large arrays in pieces of <= 1 GiB (NCCL counts are size_t, but bounded pieces keep the staging bounded)
  auto bcast = [&](void* p, uint64_t bytes) {
    constexpr uint64_t kPiece = 1ull << 30;
    for (uint64_t off = 0; off < bytes; off += kPiece) {
      const uint64_t len = std::min(kPiece, bytes - off);
      HB_NCCL(nccl().Broadcast((char*)p + off, (char*)p + off, len, ncclUint8, root, c.comm, s));
    }
  };
  bcast(idx->d_sk, h.n_total * 128);
  if (h.algo != HB_ALGO_BRUTEFORCE) {
    bcast(idx->d_adj, h.adj_words * 4ull);
    if (h.compact) bcast(idx->d_rowoff, (h.n_total + 1) * 4);
    bcast(idx->d_down, (h.n_total - h.n_base) * 4);
  }
  HB_CUDA(cudaStreamSynchronize(s));
  made.release();
  return idx;
}


// ---- one process driving N devices ---------------------------------------------------------------------------------
namespace {

// a persistent host thread bound to one device; run() hands it a task, wait() returns its error (if any)
class Worker {
 public:
  explicit Worker(int device) : device_(device), th_([this] { loop(); }) {}
  ~Worker() {
    {
      std::lock_guard<std::mutex> l(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    th_.join();
  }
  void run(std::function<void()> f) {
    std::lock_guard<std::mutex> l(mu_);
    task_ = std::move(f);
    busy_ = true;
    err_ = nullptr;
    cv_.notify_all();
  }
  std::exception_ptr wait() {
    std::unique_lock<std::mutex> l(mu_);
    cv_.wait(l, [this] { return !busy_; });
    return err_;
  }

 private:
  void loop() {
    cudaSetDevice(device_);
    for (;;) {
      std::function<void()> f;
      {
        std::unique_lock<std::mutex> l(mu_);
        cv_.wait(l, [this] { return stop_ || (busy_ && task_); });
        if (stop_) return;
        f = std::move(task_);
        task_ = nullptr;
      }
      std::exception_ptr e;
      try {
        f();
      } catch (...) {
        e = std::current_exception();
      }
      {
        std::lock_guard<std::mutex> l(mu_);
        err_ = e;
        busy_ = false;
      }
      cv_.notify_all();
    }
  }
  int device_;
  std::mutex mu_;
  std::condition_variable cv_;
  std::function<void()> task_;
  bool busy_ = false, stop_ = false;
  std::exception_ptr err_;
  std::thread th_;  // last: started after the other members exist
};

}  // namespace

struct Multi {
  std::vector<int> devices;
  std::vector<std::unique_ptr<Comm>> comms;
  std::vector<std::unique_ptr<Worker>> workers;
  std::vector<cudaStream_t> streams;
  std::vector<Index*> idx;       // one per device: replicas or shards (owned unless borrowed[r])
  std::vector<bool> borrowed;
  std::vector<uint64_t> key_offset;  // SHARD: global key of the shard's first row
  int mode = HB_MULTI_REPLICATE;
  uint64_t n_rows = 0;  // Index::size(): all rows
  std::mutex call_mu;   // one collective call at a time (the ranks' streams and buffers are shared state)

  int ndev() const { return (int)devices.size(); }
  // run fn(r) on every device's worker thread; rethrows the first failure after all have finished
  void run_all(const std::function<void(int)>& fn) {
    for (int r = 0; r < ndev(); ++r) workers[r]->run([&fn, r] { fn(r); });
    std::exception_ptr first;
    for (int r = 0; r < ndev(); ++r) {
      std::exception_ptr e = workers[r]->wait();
      if (e && !first) first = e;
    }
    if (first) std::rethrow_exception(first);
  }
  void drop_indexes() {
    for (size_t r = 0; r < idx.size(); ++r)
      if (idx[r] && !borrowed[r]) delete idx[r];
    idx.assign(devices.size(), nullptr);
    borrowed.assign(devices.size(), false);
  }
  ~Multi() {
    drop_indexes();
    for (size_t r = 0; r < streams.size(); ++r) {
      cudaSetDevice(devices[r]);
      if (streams[r]) cudaStreamDestroy(streams[r]);
    }
    comms.clear();
    workers.clear();
  }
};

namespace {

// contiguous (start, len) ranges: range r = [r*n/parts, (r+1)*n/parts) -- the reference's `-b start -l len`
void split_range(uint64_t n, int parts, int r, uint64_t& start, uint64_t& len) {
  start = n / parts * r + std::min<uint64_t>(r, n % parts);
  len = n / parts + ((uint64_t)r < n % parts ? 1 : 0);
}

Multi* as_multi(const hb_multi* h) {
  HB_REQUIRE(h != nullptr, HB_ERR_INVALID, "null multi-device handle");
  return reinterpret_cast<Multi*>(const_cast<hb_multi*>(h));
}
Comm* as_comm(const hb_comm* h) {
  HB_REQUIRE(h != nullptr, HB_ERR_INVALID, "null communicator");
  return reinterpret_cast<Comm*>(const_cast<hb_comm*>(h));
}

template <class T>
struct DevArr {
  T* p = nullptr;
  explicit DevArr(size_t n) { HB_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); }
  ~DevArr() { cudaFree(p); }
  DevArr(const DevArr&) = delete;
  DevArr& operator=(const DevArr&) = delete;
};

// device 0's packed (nq, k) list -> the caller's host arrays; found = entries present, padding like Indexes::search
void packed_to_host(const void* d_packed, uint64_t nq, uint32_t k, uint32_t flags, uint64_t* keys, uint64_t* dists, uint32_t* found,
                    cudaStream_t s) {
  DevArr<uint64_t> dk(nq * k), dd(nq * k);
  unpack_device(d_packed, nq * k, dk.p, dd.p, s);
  HB_CUDA(cudaMemcpyAsync(keys, dk.p, nq * k * 8, cudaMemcpyDeviceToHost, s));
  HB_CUDA(cudaMemcpyAsync(dists, dd.p, nq * k * 8, cudaMemcpyDeviceToHost, s));
  HB_CUDA(cudaStreamSynchronize(s));
  for (uint64_t i = 0; i < nq; ++i) {
    uint32_t f = 0;
    while (f < k && keys[i * k + f] != ~0ull) ++f;
    if (found) found[i] = f;
    const bool pad = (flags & HB_PAD_LIKE_CLI) && f > 0;
    HB_REQUIRE(!(flags & HB_PAD_LIKE_CLI) || f > 0, HB_ERR_EMPTY_RESULT, "search returned no elements (the reference panics here, main.rs:472)");
    for (uint32_t j = f; j < k; ++j) {
      keys[i * k + j] = pad ? keys[i * k] : 0;
      dists[i * k + j] = pad ? dists[i * k] : 0;
    }
  }
}

}  // namespace
}  // namespace hb

using namespace hb;

extern "C" {

// ---- communicator (one rank = one device; any launch style) ---------------------------------------------------------
int hb_comm_unique_id(uint8_t id[HB_COMM_ID_BYTES]) {
  return guarded([&] {
    HB_REQUIRE(id != nullptr, HB_ERR_INVALID, "null argument");
    static_assert(sizeof(ncclUniqueId) <= HB_COMM_ID_BYTES, "ncclUniqueId does not fit");
    ncclUniqueId u;
    HB_NCCL(nccl().GetUniqueId(&u));
    std::memset(id, 0, HB_COMM_ID_BYTES);
    std::memcpy(id, &u, sizeof u);
  });
}

int hb_comm_init_rank(const uint8_t id[HB_COMM_ID_BYTES], int nranks, int rank, int device, hb_comm** out) {
  return guarded([&] {
    HB_REQUIRE(out && (id || nranks == 1), HB_ERR_INVALID, "null argument");
    *out = nullptr;
    ncclUniqueId u;
    std::memset(&u, 0, sizeof u);
    if (id) std::memcpy(&u, id, sizeof u);
    *out = reinterpret_cast<hb_comm*>(comm_init_rank(u, nranks, rank, device));
  });
}

void hb_comm_free(hb_comm* c) {
  DeviceRestore restore;
  delete reinterpret_cast<Comm*>(c);
}

int hb_comm_phase_ms(hb_comm* h, double ms[3]) {
  return guarded([&] {
    Comm* c = as_comm(h);
    HB_REQUIRE(ms != nullptr, HB_ERR_INVALID, "null argument");
    HB_CUDA(cudaSetDevice(c->device));
    c->read_phases();
    for (int i = 0; i < 3; ++i) ms[i] = c->phase_ms[i];
  });
}

int hb_ground_truth_sharded_device(hb_comm* h, const void* d_shard, uint64_t n_local, uint64_t key_offset, const void* d_queries,
                                   uint64_t nq, uint32_t k, void* d_packed, void* stream) {
  return guarded([&] {
    Comm* c = as_comm(h);
    HB_REQUIRE(nq == 0 || (d_queries && d_packed), HB_ERR_INVALID, "null device buffer");
    HB_REQUIRE(n_local == 0 || d_shard, HB_ERR_INVALID, "null device buffer");
    HB_CUDA(cudaSetDevice(c->device));
    ground_truth_sharded(*c, d_shard, n_local, key_offset, d_queries, nq, k, d_packed, static_cast<cudaStream_t>(stream), HB_GT_AUTO);
  });
}

int hb_knns_sharded_device(hb_comm* h, const hb_index* idx, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                           uint32_t ef, void* d_packed, void* stream) {
  return guarded([&] {
    Comm* c = as_comm(h);
    HB_REQUIRE(idx != nullptr, HB_ERR_INVALID, "null index handle");
    HB_REQUIRE(nq == 0 || (d_queries && d_packed), HB_ERR_INVALID, "null device buffer");
    HB_CUDA(cudaSetDevice(c->device));
    knns_sharded(*c, *reinterpret_cast<const Index*>(idx), key_offset, d_queries, nq, k, ef, d_packed, static_cast<cudaStream_t>(stream));
  });
}

int hb_index_broadcast(hb_comm* h, hb_index* src, int root, hb_index** out) {
  return guarded([&] {
    Comm* c = as_comm(h);
    HB_REQUIRE(out != nullptr, HB_ERR_INVALID, "null argument");
    *out = nullptr;
    *out = reinterpret_cast<hb_index*>(index_broadcast(*c, reinterpret_cast<Index*>(src), root, nullptr));
  });
}

// ---- one process, N devices --------------------------------------------------------------------------------------
int hb_multi_create(const int* devices, int ndev, hb_multi** out) {
  return guarded([&] {
    HB_REQUIRE(out && devices && ndev >= 1 && ndev <= 64, HB_ERR_INVALID, "bad device list");
    *out = nullptr;
    std::unique_ptr<Multi> m(new Multi());
    m->devices.assign(devices, devices + ndev);
    for (int r = 0; r < ndev; ++r) {
      ensure_device(devices[r]);
      for (int q = 0; q < r; ++q) HB_REQUIRE(devices[q] != devices[r], HB_ERR_INVALID, "a device is listed twice");
    }
    std::vector<ncclComm_t> raw(ndev, nullptr);
    if (ndev > 1) HB_NCCL(nccl().CommInitAll(raw.data(), ndev, devices));
    m->streams.assign(ndev, nullptr);
    for (int r = 0; r < ndev; ++r) {
      std::unique_ptr<Comm> c(new Comm());
      c->comm = raw[r];
      c->nranks = ndev;
      c->rank = r;
      c->device = devices[r];
      m->comms.push_back(std::move(c));
      HB_CUDA(cudaSetDevice(devices[r]));
      HB_CUDA(cudaStreamCreateWithFlags(&m->streams[r], cudaStreamNonBlocking));
      m->workers.emplace_back(new Worker(devices[r]));
    }
    m->idx.assign(ndev, nullptr);
    m->borrowed.assign(ndev, false);
    m->key_offset.assign(ndev, 0);
    *out = reinterpret_cast<hb_multi*>(m.release());
  });
}

void hb_multi_free(hb_multi* h) {
  DeviceRestore restore;
  delete reinterpret_cast<Multi*>(h);
}

int hb_multi_device_count(const hb_multi* h) { return h ? reinterpret_cast<const Multi*>(h)->ndev() : 0; }
uint64_t hb_multi_size(const hb_multi* h) { return h ? reinterpret_cast<const Multi*>(h)->n_rows : 0; }

// REPLICATE an index that lives on devices[0] (borrowed: the caller keeps owning `src`) onto the other devices
int hb_multi_replicate(hb_multi* h, const hb_index* src) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(src != nullptr, HB_ERR_INVALID, "null index handle");
    Index* s0 = reinterpret_cast<Index*>(const_cast<hb_index*>(src));
    HB_REQUIRE(s0->device == m->devices[0], HB_ERR_INVALID, "the index to replicate must live on the first device of the list");
    std::lock_guard<std::mutex> lock(m->call_mu);
    m->drop_indexes();
    m->mode = HB_MULTI_REPLICATE;
    m->n_rows = s0->n_base;
    m->idx[0] = s0;
    m->borrowed[0] = true;
    m->run_all([&](int r) { m->idx[r] = index_broadcast(*m->comms[r], r == 0 ? s0 : nullptr, 0, m->streams[r]); });
  });
}

// read_index (cli/src/main.rs:224-234) once, then replicate over NVLink
int hb_multi_load_idx(hb_multi* h, const char* path) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(path != nullptr, HB_ERR_INVALID, "null path");
    hb_index* first = nullptr;
    int rc = hb_index_load_idx(path, m->devices[0], 0, &first);
    if (rc != HB_OK) throw Error(rc, hb_last_error());
    rc = hb_multi_replicate(h, first);
    if (rc != HB_OK) {
      const std::string msg = hb_last_error();
      hb_index_free(first);
      std::lock_guard<std::mutex> lock(m->call_mu);
      m->idx[0] = nullptr;
      throw Error(rc, msg);
    }
    m->borrowed[0] = false;  // loaded here: owned here
  });
}

// SHARD: device r builds an HNSW over rows [start_r, start_r + len_r) of `data` (`index -b start_r -l len_r`)
int hb_multi_build_sharded(hb_multi* h, const uint64_t* data, uint64_t n, const hb_build_params* p) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(p && (n == 0 || data), HB_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lock(m->call_mu);
    m->drop_indexes();
    m->mode = HB_MULTI_SHARD;
    m->n_rows = n;
    m->run_all([&](int r) {
      uint64_t start, len;
      split_range(n, m->ndev(), r, start, len);
      m->key_offset[r] = start;
      m->idx[r] = build_hnsw(data + start * 16, len, *p, m->devices[r], false);
    });
  });
}

// SHARD from files: paths[r] is the .idx of shard r, key_offsets[r] the global key of its first row
int hb_multi_load_shards(hb_multi* h, const char* const* paths, const uint64_t* key_offsets) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(paths && key_offsets, HB_ERR_INVALID, "null argument");
    std::lock_guard<std::mutex> lock(m->call_mu);
    m->drop_indexes();
    m->mode = HB_MULTI_SHARD;
    m->run_all([&](int r) {
      hb_index* one = nullptr;
      const int rc = hb_index_load_idx(paths[r], m->devices[r], 0, &one);
      if (rc != HB_OK) throw Error(rc, hb_last_error());
      m->idx[r] = reinterpret_cast<Index*>(one);
      m->key_offset[r] = key_offsets[r];
    });
    m->n_rows = 0;
    for (Index* i : m->idx) m->n_rows += i->n_base;
  });
}

// Index::knns over all devices.  HOST buffers, results in input order.
int hb_multi_knns(hb_multi* h, const uint64_t* queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, uint64_t* keys,
                  uint64_t* dists, uint32_t* found) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(nq == 0 || (queries && keys && dists && found), HB_ERR_INVALID, "null host buffer");
    HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
    if (nq == 0) return;
    std::lock_guard<std::mutex> lock(m->call_mu);
    for (Index* i : m->idx) HB_REQUIRE(i != nullptr, HB_ERR_INVALID, "no index loaded on this multi-device handle");
    if (m->mode == HB_MULTI_REPLICATE) {
      // contiguous slices of the batch, disjoint slices of the outputs: no collective
      m->run_all([&](int r) {
        uint64_t start, len;
        split_range(nq, m->ndev(), r, start, len);
        if (!len) return;
        const int rc = hb_knns(reinterpret_cast<hb_index*>(m->idx[r]), queries + start * 16, len, k, ef, flags, keys + start * k,
                               dists + start * k, found + start, nullptr);
        if (rc != HB_OK) throw Error(rc, hb_last_error());
      });
      return;
    }
    // SHARD: every device searches its shard for all queries; all-gather + exact merge; device 0 returns the result
    std::vector<std::unique_ptr<DevArr<uint64_t>>> dq(m->ndev()), dout(m->ndev());
    m->run_all([&](int r) {
      cudaStream_t s = m->streams[r];
      dq[r].reset(new DevArr<uint64_t>(nq * 16));
      dout[r].reset(new DevArr<uint64_t>(nq * k));
      HB_CUDA(cudaMemcpyAsync(dq[r]->p, queries, nq * 128, cudaMemcpyHostToDevice, s));
      knns_sharded(*m->comms[r], *m->idx[r], m->key_offset[r], dq[r]->p, nq, k, ef, dout[r]->p, s);
      HB_CUDA(cudaStreamSynchronize(s));
      if (r == 0) packed_to_host(dout[0]->p, nq, k, flags, keys, dists, found, s);
    });
  });
}

// `ground-truth` with the rows sharded over the devices.  HOST buffers; keys/dists (nq, k), ascending, n < k pads
// with UINT64_MAX like hb_ground_truth.
int hb_multi_ground_truth(hb_multi* h, const uint64_t* data, uint64_t n, const uint64_t* queries, uint64_t nq, uint32_t k,
                          uint64_t* keys, uint64_t* dists) {
  return guarded([&] {
    Multi* m = as_multi(h);
    HB_REQUIRE(nq == 0 || (queries && keys && dists), HB_ERR_INVALID, "null host buffer");
    HB_REQUIRE(n == 0 || data, HB_ERR_INVALID, "null host buffer");
    HB_REQUIRE(k >= 1, HB_ERR_INVALID, "k out of range");
    if (nq == 0) return;
    std::lock_guard<std::mutex> lock(m->call_mu);
    m->run_all([&](int r) {
      cudaStream_t s = m->streams[r];
      uint64_t start, len;
      split_range(n, m->ndev(), r, start, len);
      DevArr<uint64_t> dd(len * 16), dq(nq * 16), dout(nq * k);
      if (len) HB_CUDA(cudaMemcpyAsync(dd.p, data + start * 16, len * 128, cudaMemcpyHostToDevice, s));
      HB_CUDA(cudaMemcpyAsync(dq.p, queries, nq * 128, cudaMemcpyHostToDevice, s));
      ground_truth_sharded(*m->comms[r], dd.p, len, start, dq.p, nq, k, dout.p, s, HB_GT_AUTO);
      HB_CUDA(cudaStreamSynchronize(s));
      if (r == 0) {
        DevArr<uint64_t> dk(nq * k), ds(nq * k);
        unpack_device(dout.p, nq * k, dk.p, ds.p, s);
        HB_CUDA(cudaMemcpyAsync(keys, dk.p, nq * k * 8, cudaMemcpyDeviceToHost, s));
        HB_CUDA(cudaMemcpyAsync(dists, ds.p, nq * k * 8, cudaMemcpyDeviceToHost, s));
        HB_CUDA(cudaStreamSynchronize(s));
      }
    });
  });
}

// device time (ms) of the last sharded call on the first device: [0] search, [1] all-gather, [2] merge
int hb_multi_phase_ms(hb_multi* h, double ms[3]) {
  return guarded([&] {
    Multi* m = as_multi(h);
    return (void)hb_comm_phase_ms(reinterpret_cast<hb_comm*>(m->comms[0].get()), ms);
  });
}

// the per-device index handle (replica or shard r), borrowed
const hb_index* hb_multi_index(const hb_multi* h, int r) {
  if (!h) return nullptr;
  const Multi* m = reinterpret_cast<const Multi*>(h);
  return (r >= 0 && r < m->ndev()) ? reinterpret_cast<const hb_index*>(m->idx[r]) : nullptr;
}

}  // extern "C"
