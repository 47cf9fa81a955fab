// hb_internal.h -- shared declarations of libhnswb200 (product code; never includes oracle/).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/hnsw_b200.h"

namespace hb {

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};

void set_last_error(const std::string& m);
void count_launches(uint64_t n = 1);  // every kernel launch of the library is counted (hb_kernel_launches)

#define HB_CUDA(expr)                                                                                     \
  do {                                                                                                    \
    cudaError_t e_ = (expr);                                                                              \
    if (e_ != cudaSuccess) {                                                                              \
      int c_ = (e_ == cudaErrorMemoryAllocation) ? HB_ERR_OOM                                             \
               : (e_ == cudaErrorNoDevice || e_ == cudaErrorInsufficientDriver || e_ == cudaErrorInvalidDevice) ? HB_ERR_NO_DEVICE \
                                                                                                        : HB_ERR_CUDA; \
      throw hb::Error(c_, std::string(#expr) + ": " + cudaGetErrorString(e_));                            \
    }                                                                                                     \
  } while (0)

#define HB_REQUIRE(cond, code, msg) \
  do {                              \
    if (!(cond)) throw hb::Error((code), (msg)); \
  } while (0)

// every extern "C" entry point runs its body through this: exceptions become hb_status codes + hb_last_error(), and the
// caller's current CUDA device is put back if the call changed it (the entry points cudaSetDevice to the index's /
// communicator's device; a host that keeps its own notion of "current device", like PyTorch, must not see that)
struct DeviceRestore {
  int prev = -1;
  DeviceRestore() {
    if (cudaGetDevice(&prev) != cudaSuccess) {
      cudaGetLastError();
      prev = -1;
    }
  }
  ~DeviceRestore() {
    int cur = -1;
    if (prev >= 0 && cudaGetDevice(&cur) == cudaSuccess && cur != prev) cudaSetDevice(prev);
  }
};

template <class F>
int guarded(F&& f) {
  DeviceRestore restore;
  try {
    f();
    return HB_OK;
  } catch (const Error& e) {
    set_last_error(e.what());
    return e.code;
  } catch (const std::bad_alloc&) {
    set_last_error("host allocation failed");
    return HB_ERR_OOM;
  } catch (const std::exception& e) {
    set_last_error(e.what());
    return HB_ERR_INTERNAL;
  } catch (...) {
    set_last_error("unknown error");
    return HB_ERR_INTERNAL;
  }
}

// ---- host-side graph arrays (one per graph; graph 0 = base) -----------------
struct HostGraph {
  uint64_t n = 0;
  std::vector<uint64_t> sk;    // n*16
  std::vector<uint64_t> down;  // n (upper layers only): index into the graph below
  std::vector<uint64_t> off;   // n+1
  std::vector<uint32_t> ids;   // canonical (file) order
};

// ResultAttrs, cli/src/main.rs:311-320
struct Attrs {
  bool format_size = true;
  std::string data = "hamming";
  uint64_t size = 0;
  uint32_t algo = 0;
  double buildtime = 0.0, querytime = 0.0;
  std::string params;
};

// scratch of ONE search call (control words, retry maps, the exact visited tables of the large-capacity kernel, timing
// events).  Pooled on the index: concurrent hb_knns* calls each take their own, so nothing serialises them
// (Index::search is callable from any thread, src/index/mod.rs:21-43).
struct SearchWorkspace {
  uint32_t vcap_log2 = 0;  // visited-table capacity per warp slot (log2, 32-bit words)
  uint32_t slots = 0;      // resident warp slots the tables were sized for
  uint32_t spill_cap = 0;
  uint32_t* vis = nullptr;     // slots << vcap_log2 words, all EMPTY between launches
  uint32_t* spill = nullptr;   // slots * spill_cap
  uint32_t* ctl = nullptr;     // [0] query counter, [1] overflow flag
  uint32_t* qstatus = nullptr; // per query (capacity qcap)
  uint32_t* qmap = nullptr;    // retry indirection (capacity qcap)
  uint64_t qcap = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  ~SearchWorkspace();
};

// device staging of one hb_knns call with host buffers (pooled on the index: no cudaMalloc on the hot path)
struct CallBuffers {
  uint64_t* q = nullptr;   // cap_q * 16
  uint64_t* keys = nullptr;  // cap_out
  uint64_t* dists = nullptr; // cap_out
  uint32_t* found = nullptr; // cap_q
  hb_counters* ctr = nullptr;  // cap_q
  uint64_t cap_q = 0, cap_out = 0;
  // page-locked, mapped host staging for callers with pageable buffers (hb_knns): the kernel reads the queries and
  // writes the results through these across PCIe, overlapped with the search, like it does for pinned caller memory
  uint64_t *hq = nullptr, *hkeys = nullptr, *hdists = nullptr;
  uint32_t* hfound = nullptr;
  uint64_t hcap_q = 0, hcap_out = 0;
  cudaStream_t stream = nullptr;
};

struct Index {
  int device = 0;
  int algo = HB_ALGO_HNSW;
  Attrs attrs;
  bool has_host = false;
  std::vector<HostGraph> host;  // kept when keep_host
  // dims (always kept)
  std::vector<uint64_t> g_nodes, g_edges, g_base;  // g_base[g] = first global id of graph g
  int64_t ep_file = -1;
  // device arrays: one id space over all graphs (base first)
  uint64_t n_total = 0, n_base = 0;
  uint32_t stride = 0;        // words per adjacency row, multiple of 32 (row = one or more 128-B lines)
  uint4* d_sk = nullptr;      // n_total rows of 128 B
  uint32_t* d_adj = nullptr;  // n_total * stride; word 0 = degree, then global neighbour ids
  // compact layout (index_compact): row gid starts at line d_rowoff[gid] (128-B lines) and spans ceil((deg + 1) / 32) lines
  uint32_t* d_rowoff = nullptr;
  uint64_t adj_words = 0;     // 32-bit words allocated in d_adj
  uint32_t* d_down = nullptr; // n_total - n_base: global id of the same point one graph below
  uint32_t ep_gid = 0;
  uint64_t device_bytes = 0;
  int sm_count = 0;
  uint64_t build_dropped_appends = 0, build_prunes = 0, build_unlinked = 0;  // builder.cu statistics
  mutable std::mutex pool_mu;
  mutable std::vector<CallBuffers> pool;  // idle staging sets
  mutable std::vector<std::unique_ptr<SearchWorkspace>> ws_pool;  // idle search workspaces (guarded by pool_mu)
  mutable double last_search_ms = 0.0;  // device time of the search kernel(s) of the last completed knns call on this index
  ~Index();
};

// idx_format.cpp
void read_idx_file(const char* path, Attrs& attrs, int& algo, std::vector<HostGraph>& graphs, int64_t& ep);
void write_idx_file(const char* path, const Attrs& attrs, int algo, const std::vector<HostGraph>& graphs, int64_t ep);

// index.cu
Index* index_from_host(int algo, std::vector<HostGraph>&& graphs, int64_t ep, int device, bool keep_host);
void ensure_device(int device);
void index_compact(Index& idx);  // fixed-stride rows -> line-granular rows + offsets (5x less HBM at M = 256)

// search.cu
void knns_device(const Index& idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                 void* d_dists, void* d_found, void* d_counters, cudaStream_t stream);
// the ef <= 256 path with its scratch in a static device array (no allocation of any kind; hb_knns_device_noalloc)
void knns_device_noalloc(const Index& idx, const void* d_queries, uint64_t nq, uint32_t k, uint32_t ef, uint32_t flags, void* d_keys,
                         void* d_dists, void* d_found, cudaStream_t stream);

// bruteforce.cu
// d_bound (optional, nq u32, 0xFFFFFFFF = none yet): per-query upper bound on the global k-th distance, read and
// tightened by the E2M1 kernel.  A caller that searches one data set slice by slice passes the same array to every
// call so that later slices start with the earlier slices' bound; the per-slice lists may then hold fewer than k rows
// (the rest cannot be in the global top-k) and the exact merge of the slices is unchanged.
void ground_truth_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                         void* d_packed, cudaStream_t stream, int algo = HB_GT_AUTO, uint32_t* d_bound = nullptr);
void ground_truth_scan_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                              void* d_packed, cudaStream_t stream);
// gt_tcgen05.cu
bool ground_truth_mma_supported(const void* d_data, const void* d_queries, uint64_t n, uint64_t nq, uint32_t k);
void ground_truth_mma_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                             void* d_packed, cudaStream_t stream);
// the same GEMM with E2M1 operands (tcgen05.mma kind::mxf4.block_scale, unit scales); gt_mxf4.cu
bool ground_truth_mxf4_supported(const void* d_data, const void* d_queries, uint64_t n, uint64_t nq, uint32_t k);
void ground_truth_mxf4_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                             void* d_packed, cudaStream_t stream, uint32_t* d_bound = nullptr);
void topk_merge_device(const void* d_lists, uint32_t n_lists, uint64_t nq, uint32_t k, void* d_out, cudaStream_t stream,
                       bool sorted = true, uint32_t width = 0);  // !sorted: lists of `width` (default k) slots in any order, kNone = none
void unpack_device(const void* d_packed, uint64_t n, void* d_keys, void* d_dists, cudaStream_t stream);
void pack_device(const void* d_keys, const void* d_dists, const void* d_found, uint64_t nq, uint32_t k, uint64_t key_offset,
                 void* d_packed, cudaStream_t stream);

// synth.cu
void synth_device(const hb_synth_params& p, uint64_t start, uint64_t n, void* d_out, cudaStream_t stream);

// multi.cu
struct Comm;

// builder.cu
Index* build_hnsw(const uint64_t* data, uint64_t n, const hb_build_params& p, int device, bool keep_host);

}  // namespace hb
