// index.cu -- device-resident index layout (north_star subsystem 1).
//
// Replaces the in-memory shape of HNSW<P> / NSW<P> (src/index/hnsw.rs:255-261,
// src/index/nsw.rs:324-328) and SimpleGraph (src/collections/simplegraph.rs:10-14:
// Vec<T> nodes + Vec<HashSet<usize>>) with a layout made for HBM:
//
//   * ONE id space over all graphs: base nodes are gid [0, N), layers[0] follows,
//     then layers[1] ...  Neighbour ids and down-links are stored as global ids,
//     so the search kernel runs the same inner loop on every layer and the
//     visited table never has to be cleared between layers.
//   * sketches: n_total rows of 128 B (8 x uint4), 128-B aligned -> one row =
//     4 full 32-B sectors, fetched by 8 lanes x LDG.128.  For the base graph this
//     is byte-identical to the file's node array.
//   * adjacency: fixed-stride rows ("ELL"), stride = multiple of 32 words so each
//     row starts on a 128-B line; word 0 = degree, words 1..deg = neighbour gids
//     in FILE ORDER (the canonical order).  A fixed stride removes the
//     CSR offset lookup -- one dependent HBM round trip per expansion -- and costs
//     only address space: bytes actually read are still 4*(deg+1).  At M = 256
//     (max degree M+1 = 257, nsw.rs:121-129) that is 1152 B/node = 11.5 GB for
//     10 M nodes out of 180 GB.
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>

#include "hb_internal.h"

#include <atomic>

namespace hb {

namespace {

// one warp per node: copy its CSR neighbour run into the fixed-stride row, adding the graph's gid base
__global__ void csr_to_rows_kernel(const uint64_t* __restrict__ off, const uint32_t* __restrict__ ids, uint64_t n,
                                   uint32_t gid_base, uint32_t stride, uint32_t* __restrict__ rows) {
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (warp >= n) return;
  const uint64_t b = off[warp], e = off[warp + 1];
  uint32_t* row = rows + warp * stride;
  if (lane == 0) row[0] = (uint32_t)(e - b);
  for (uint64_t j = b + lane; j < e; j += 32) row[1 + (j - b)] = ids[j] + gid_base;
}

__global__ void down_to_gid_kernel(const uint64_t* __restrict__ down, uint64_t n, uint32_t below_base, uint32_t* __restrict__ out) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = (uint32_t)down[i] + below_base;
}

// lines (of 32 words) a row needs: its degree word + deg neighbour ids
__global__ void row_lines_kernel(const uint32_t* __restrict__ rows, uint32_t stride, uint64_t n, uint32_t* __restrict__ lines) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) lines[i] = (rows[i * stride] + 1u + 31u) / 32u;
}
// one warp per node: copy the used lines of its fixed-stride row to its place in the compact array
__global__ void compact_rows_kernel(const uint32_t* __restrict__ rows, uint32_t stride, uint64_t n, const uint32_t* __restrict__ rowoff,
                                    uint32_t* __restrict__ out) {
  const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const uint32_t lane = threadIdx.x & 31;
  if (warp >= n) return;
  const uint32_t* src = rows + warp * stride;
  const uint32_t words = (rowoff[warp + 1] - rowoff[warp]) * 32u;
  uint32_t* dst = out + (size_t)rowoff[warp] * 32;
  for (uint32_t j = lane; j < words; j += 32) dst[j] = src[j];
}

template <class T>
struct DevBuf {
  T* p = nullptr;
  explicit DevBuf(size_t n) { HB_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T))); }
  ~DevBuf() { cudaFree(p); }
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
};

}  // namespace

void ensure_device(int device) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    throw Error(HB_ERR_NO_DEVICE, std::string("no CUDA device available (there is no CPU fallback): ") +
                                      (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0"));
  }
  HB_REQUIRE(device >= 0 && device < count, HB_ERR_INVALID, "device ordinal out of range");
  HB_CUDA(cudaSetDevice(device));
  // The stream-ordered allocations of the kernels' scratch (cudaMallocAsync) come from the device's default pool,
  // which by default returns its memory to the driver at every synchronisation: keep up to 4 GB cached so that
  // repeated calls do not pay for the allocation again.
  static std::atomic<uint64_t> configured{0};
  if (device < 64 && !(configured.load() >> device & 1)) {
    cudaMemPool_t pool = nullptr;
    if (cudaDeviceGetDefaultMemPool(&pool, device) == cudaSuccess) {
      uint64_t keep = 4ull << 30;
      cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
    }
    cudaGetLastError();
    configured.fetch_or(1ull << device);
  }
}

Index::~Index() {
  cudaSetDevice(device);
  cudaFree(d_sk);
  cudaFree(d_adj);
  cudaFree(d_rowoff);
  cudaFree(d_down);
  ws_pool.clear();
  for (CallBuffers& b : pool) {
    cudaFree(b.q); cudaFree(b.keys); cudaFree(b.dists); cudaFree(b.found); cudaFree(b.ctr);
    cudaFreeHost(b.hq); cudaFreeHost(b.hkeys); cudaFreeHost(b.hdists); cudaFreeHost(b.hfound);
    if (b.stream) cudaStreamDestroy(b.stream);
  }
}

SearchWorkspace::~SearchWorkspace() {
  cudaFree(vis);
  cudaFree(spill);
  cudaFree(ctl);
  cudaFree(qstatus);
  cudaFree(qmap);
  if (ev0) cudaEventDestroy(ev0);
  if (ev1) cudaEventDestroy(ev1);
}

Index* index_from_host(int algo, std::vector<HostGraph>&& graphs, int64_t ep, int device, bool keep_host) {
  HB_REQUIRE(algo >= HB_ALGO_BRUTEFORCE && algo <= HB_ALGO_HNSW, HB_ERR_INVALID, "unknown algo");
  HB_REQUIRE(!graphs.empty(), HB_ERR_INVALID, "an index needs at least the base graph");
  HB_REQUIRE(algo == HB_ALGO_HNSW || graphs.size() == 1, HB_ERR_INVALID, "only HNSW has upper layers");
  ensure_device(device);

  std::unique_ptr<Index> idx(new Index());
  idx->device = device;
  idx->algo = algo;
  idx->attrs.algo = (uint32_t)algo;
  idx->ep_file = ep;
  const size_t G = graphs.size();
  idx->g_nodes.resize(G);
  idx->g_edges.resize(G);
  idx->g_base.resize(G);

  // ---- validate and size --------------------------------------------------
  uint64_t total = 0, max_deg = 0;
  for (size_t g = 0; g < G; ++g) {
    HostGraph& hg = graphs[g];
    HB_REQUIRE(hg.sk.size() == hg.n * 16, HB_ERR_FORMAT, "sketch array size mismatch");
    if (hg.off.empty()) hg.off.assign((size_t)hg.n + 1, 0);
    HB_REQUIRE(hg.off.size() == hg.n + 1 && hg.off[0] == 0, HB_ERR_FORMAT, "offset array size mismatch");
    HB_REQUIRE(hg.off.back() == hg.ids.size(), HB_ERR_FORMAT, "offsets do not cover the id array");
    for (uint64_t v = 0; v < hg.n; ++v) {
      HB_REQUIRE(hg.off[v] <= hg.off[v + 1], HB_ERR_FORMAT, "offsets are not monotone");
      max_deg = std::max(max_deg, hg.off[v + 1] - hg.off[v]);
    }
    for (uint32_t id : hg.ids) HB_REQUIRE(id < hg.n, HB_ERR_FORMAT, "neighbour id out of range");
    if (g > 0) {
      HB_REQUIRE(hg.down.size() == hg.n, HB_ERR_FORMAT, "down-link array size mismatch");
      for (uint64_t d : hg.down) HB_REQUIRE(d < graphs[g - 1].n, HB_ERR_FORMAT, "down-link out of range");
    }
    idx->g_nodes[g] = hg.n;
    idx->g_edges[g] = hg.ids.size();
    idx->g_base[g] = total;
    total += hg.n;
  }
  HB_REQUIRE(total <= 0xFFFFFFFEull, HB_ERR_UNSUPPORTED, "more than 2^32-2 nodes over all graphs");
  HB_REQUIRE(max_deg <= 65535, HB_ERR_UNSUPPORTED, "degree above 65535");
  idx->n_total = total;
  idx->n_base = graphs[0].n;
  idx->attrs.size = graphs[0].n;
  if (algo != HB_ALGO_BRUTEFORCE && ep >= 0) {
    HB_REQUIRE((uint64_t)ep < graphs[G - 1].n, HB_ERR_FORMAT, "entry point out of range");
    idx->ep_gid = (uint32_t)(idx->g_base[G - 1] + (uint64_t)ep);
  }
  idx->stride = std::max<uint32_t>(64u, (uint32_t)(((max_deg + 1) + 31) / 32 * 32));  // >= 2 lines: the search loads two per expansion

  cudaDeviceProp prop;
  HB_CUDA(cudaGetDeviceProperties(&prop, device));
  idx->sm_count = prop.multiProcessorCount;

  // ---- upload ---------------------------------------------------------------
  HB_CUDA(cudaMalloc(&idx->d_sk, std::max<uint64_t>(total, 1) * 128));
  idx->device_bytes += total * 128;
  for (size_t g = 0; g < G; ++g)
    if (graphs[g].n)
      HB_CUDA(cudaMemcpy((char*)idx->d_sk + idx->g_base[g] * 128, graphs[g].sk.data(), graphs[g].n * 128, cudaMemcpyHostToDevice));

  if (algo != HB_ALGO_BRUTEFORCE) {
    const uint64_t adj_bytes = total * idx->stride * 4ull;
    HB_CUDA(cudaMalloc(&idx->d_adj, std::max<uint64_t>(adj_bytes, 4)));
    idx->device_bytes += adj_bytes;
    idx->adj_words = total * idx->stride;
    const uint64_t n_upper = total - idx->n_base;
    HB_CUDA(cudaMalloc(&idx->d_down, std::max<uint64_t>(n_upper, 1) * 4));
    idx->device_bytes += n_upper * 4;
    for (size_t g = 0; g < G; ++g) {
      HostGraph& hg = graphs[g];
      if (!hg.n) continue;
      DevBuf<uint64_t> d_off(hg.n + 1);
      DevBuf<uint32_t> d_ids(hg.ids.size());
      HB_CUDA(cudaMemcpy(d_off.p, hg.off.data(), (hg.n + 1) * 8, cudaMemcpyHostToDevice));
      if (!hg.ids.empty()) HB_CUDA(cudaMemcpy(d_ids.p, hg.ids.data(), hg.ids.size() * 4, cudaMemcpyHostToDevice));
      const uint64_t threads = hg.n * 32;
      csr_to_rows_kernel<<<(unsigned)((threads + 255) / 256), 256>>>(d_off.p, d_ids.p, hg.n, (uint32_t)idx->g_base[g], idx->stride,
                                                                     idx->d_adj + idx->g_base[g] * idx->stride);
    count_launches();
      HB_CUDA(cudaGetLastError());
      if (g > 0) {
        DevBuf<uint64_t> d_dn(hg.n);
        HB_CUDA(cudaMemcpy(d_dn.p, hg.down.data(), hg.n * 8, cudaMemcpyHostToDevice));
        down_to_gid_kernel<<<(unsigned)((hg.n + 255) / 256), 256>>>(d_dn.p, hg.n, (uint32_t)idx->g_base[g - 1],
                                                                   idx->d_down + (idx->g_base[g] - idx->n_base));
    count_launches();
        HB_CUDA(cudaGetLastError());
        HB_CUDA(cudaDeviceSynchronize());
      }
      HB_CUDA(cudaDeviceSynchronize());
    }
  }
  if (keep_host) {
    idx->host = std::move(graphs);
    idx->has_host = true;
  }
  static const bool compact = std::getenv("HB_INDEX_LAYOUT") && std::string(std::getenv("HB_INDEX_LAYOUT")) == "compact";
  if (compact) index_compact(*idx);
  return idx.release();
}

// Fixed-stride rows cost address space, not bandwidth -- but at M = 256 that is 1 408 B per node (16 GB for 10 M points)
// for ~190 B of ids.  The compact layout keeps each row's used 128-B lines back to back and finds a row through one
// 32-bit line offset per node: ~3 GB for the same index, at the price of one more dependent 4-byte read per expansion.
void index_compact(Index& idx) {
  if (idx.d_rowoff || idx.algo == HB_ALGO_BRUTEFORCE || idx.n_total == 0) return;
  HB_CUDA(cudaSetDevice(idx.device));
  const uint64_t n = idx.n_total;
  DevBuf<uint32_t> d_lines(n);
  row_lines_kernel<<<(unsigned)((n + 255) / 256), 256>>>(idx.d_adj, idx.stride, n, d_lines.p);
  count_launches();
  HB_CUDA(cudaGetLastError());
  std::vector<uint32_t> lines(n), off(n + 1);
  HB_CUDA(cudaMemcpy(lines.data(), d_lines.p, n * 4, cudaMemcpyDeviceToHost));
  uint64_t total = 0;
  for (uint64_t i = 0; i < n; ++i) {
    off[i] = (uint32_t)total;
    total += lines[i];
  }
  HB_REQUIRE(total < 0xFFFFFFFFull, HB_ERR_UNSUPPORTED, "compact adjacency above 2^32 lines");
  off[n] = (uint32_t)total;
  uint32_t *d_off = nullptr, *d_out = nullptr;
  HB_CUDA(cudaMalloc(&d_off, (n + 1) * 4));
  const uint64_t words = (total + 1) * 32;  // + one line: the search always loads two lines of a row
  cudaError_t e = cudaMalloc(&d_out, words * 4);
  if (e != cudaSuccess) {
    cudaFree(d_off);
    HB_CUDA(e);
  }
  HB_CUDA(cudaMemcpy(d_off, off.data(), (n + 1) * 4, cudaMemcpyHostToDevice));
  HB_CUDA(cudaMemset(d_out + total * 32, 0, 128));
  compact_rows_kernel<<<(unsigned)((n * 32 + 255) / 256), 256>>>(idx.d_adj, idx.stride, n, d_off, d_out);
  count_launches();
  HB_CUDA(cudaGetLastError());
  HB_CUDA(cudaDeviceSynchronize());
  idx.device_bytes -= idx.adj_words * 4;
  cudaFree(idx.d_adj);
  idx.d_adj = d_out;
  idx.d_rowoff = d_off;
  idx.adj_words = words;
  idx.device_bytes += words * 4 + (n + 1) * 4;
}

}  // namespace hb
