// bruteforce.cu -- exact k-nearest by (distance, key): the `ground-truth` path.
//
// Replaces Bruteforce::search (src/index/bruteforce.rs:37-47) + MinK::min_k
// (src/collections/mod.rs:52-77) + the CLI's per-row sort (cli/src/main.rs:738-741):
// for every query the k smallest (Hamming distance, key), ascending.  Because
// the order is total, any exact selection gives the reference's set, and the
// sorted list is the reference's CLI-observable output.
//
// This file is the CUDA-core XOR+POPC scan (each data row is loaded once per
// 16-query tile and compared against queries broadcast from shared memory;
// a per-query threshold = current k-th best filters almost every candidate
// with one compare).  The tcgen05 +-1 int8 GEMM formulation lives in
// gt_tcgen05.cu and shares the merge / unpack kernels below.
//
// Packed result element: (distance << 32) | key  -- one u64 compare == the
// reference's Ord (src/index/mod.rs:90-97); UINT64_MAX = "no element".
#include <algorithm>

#include "hb_internal.h"

namespace hb {
namespace {

constexpr uint32_t kFull = 0xFFFFFFFFu;
constexpr uint64_t kNone = ~0ull;
constexpr int kScanWarps = 8;

// insert v into the ascending list[0..k) (drops the last element); warp-collective
__device__ __forceinline__ void topk_insert(uint64_t* list, uint32_t k, uint64_t v, uint32_t lane) {
  uint32_t pos = 0;
  for (uint32_t base = 0; base < k; base += 32) {
    const uint32_t i = base + lane;
    const uint32_t m = __ballot_sync(kFull, (i < k) && (list[i] < v));
    pos += __popc(m);
    if (m != kFull) break;
  }
  const int hi = (int)k - 2;
  if (hi >= (int)pos) {
    for (int base = hi & ~31; base + 31 >= (int)pos; base -= 32) {
      const int i = base + (int)lane;
      const bool mv = i >= (int)pos && i <= hi;
      const uint64_t e = mv ? list[i] : 0ull;
      __syncwarp();
      if (mv) list[i + 1] = e;
      __syncwarp();
    }
  }
  if (lane == 0) list[pos] = v;
  __syncwarp();
}

// One warp = one (query tile, data chunk) work item.
//   smem per warp: qt queries (qt*8 uint4) + qt lists of k u64
__global__ void __launch_bounds__(kScanWarps * 32) scan_topk_kernel(const uint4* __restrict__ data, uint64_t n, uint64_t key_offset,
                                                                   const uint4* __restrict__ queries, uint32_t nq, uint32_t k,
                                                                   uint32_t qt, uint32_t n_qtiles, uint32_t n_chunks,
                                                                   uint64_t chunk_rows, uint64_t* __restrict__ partial) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const size_t per_warp = (size_t)qt * 128 + (size_t)qt * k * 8;
  uint4* qs = reinterpret_cast<uint4*>(smem_raw + wib * per_warp);
  uint64_t* lists = reinterpret_cast<uint64_t*>(smem_raw + wib * per_warp + (size_t)qt * 128);

  const uint64_t item = (uint64_t)blockIdx.x * kScanWarps + wib;
  if (item >= (uint64_t)n_qtiles * n_chunks) return;
  // query tile varies fastest so that concurrently running warps share data rows in L2
  const uint32_t tile = (uint32_t)(item % n_qtiles), chunk = (uint32_t)(item / n_qtiles);
  const uint32_t q0 = tile * qt, nqt = min(qt, nq - q0);
  for (uint32_t i = lane; i < nqt * 8; i += 32) qs[i] = __ldg(queries + (size_t)q0 * 8 + i);
  for (uint32_t i = lane; i < nqt * k; i += 32) lists[i] = kNone;
  __syncwarp();

  const uint64_t r0 = (uint64_t)chunk * chunk_rows, r1 = min(n, r0 + chunk_rows);
  for (uint64_t rb = r0; rb < r1; rb += 32) {
    const uint64_t r = rb + lane;
    const bool have = r < r1;
    uint4 row[8];
    const uint4* src = data + (have ? r : r0) * 8;
#pragma unroll
    for (int j = 0; j < 8; ++j) row[j] = __ldg(src + j);
    const uint64_t key = key_offset + r;
    for (uint32_t qi = 0; qi < nqt; ++qi) {
      uint32_t d = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const uint4 qv = qs[qi * 8 + j];
        d += __popc(row[j].x ^ qv.x) + __popc(row[j].y ^ qv.y) + __popc(row[j].z ^ qv.z) + __popc(row[j].w ^ qv.w);
      }
      const uint64_t cand = ((uint64_t)d << 32) | key;
      uint64_t* list = lists + (size_t)qi * k;
      uint32_t pm = __ballot_sync(kFull, have && cand < list[k - 1]);
      while (pm) {
        const int src_lane = __ffs(pm) - 1;
        const uint64_t v = __shfl_sync(kFull, cand, src_lane);
        if (v < list[k - 1]) topk_insert(list, k, v, lane);  // threshold may have moved since the ballot
        pm &= pm - 1;
      }
    }
  }
  __syncwarp();
  for (uint32_t i = lane; i < nqt * k; i += 32) {
    const uint32_t qi = i / k, j = i % k;
    partial[((size_t)chunk * nq + q0 + qi) * k + j] = lists[i];
  }
}

// one warp per query: exact merge of n_lists lists of `width` slots (ascending with width == k, or in any order when
// !sorted; kNone = no entry) into the k smallest
__global__ void __launch_bounds__(128) topk_merge_kernel(const uint64_t* __restrict__ lists, uint32_t n_lists, uint64_t nq, uint32_t k,
                                                         uint32_t sorted, uint32_t width, uint64_t* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  uint64_t* list = reinterpret_cast<uint64_t*>(smem_raw) + (size_t)wib * k;
  const uint64_t q = (uint64_t)blockIdx.x * 4 + wib;
  if (q >= nq) return;
  for (uint32_t i = lane; i < k; i += 32) list[i] = sorted ? lists[q * width + i] : kNone;
  __syncwarp();
  for (uint32_t l = sorted ? 1 : 0; l < n_lists; ++l) {
    const uint64_t* src = lists + ((size_t)l * nq + q) * width;
    for (uint32_t base = 0; base < width; base += 32) {
      const uint32_t i = base + lane;
      const uint64_t cand = (i < width) ? src[i] : kNone;
      uint32_t pm = __ballot_sync(kFull, cand < list[k - 1]);
      if (!pm) {
        if (sorted) break;  // src ascending: nothing further can qualify
        continue;
      }
      while (pm) {
        const int sl = __ffs(pm) - 1;
        const uint64_t v = __shfl_sync(kFull, cand, sl);
        if (v < list[k - 1]) topk_insert(list, k, v, lane);
        pm &= pm - 1;
      }
    }
  }
  __syncwarp();
  for (uint32_t i = lane; i < k; i += 32) out[q * k + i] = list[i];
}

__global__ void unpack_kernel(const uint64_t* __restrict__ packed, uint64_t n, uint64_t* __restrict__ keys, uint64_t* __restrict__ dists) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t v = packed[i];
  keys[i] = (v == kNone) ? kNone : (v & 0xFFFFFFFFull);
  dists[i] = (v == kNone) ? kNone : (v >> 32);
}

// (keys, dists, found) of a per-shard graph search -> packed ascending lists with global keys (entries past found = none)
__global__ void pack_kernel(const uint64_t* __restrict__ keys, const uint64_t* __restrict__ dists, const uint32_t* __restrict__ found,
                            uint64_t nq, uint32_t k, uint64_t key_offset, uint64_t* __restrict__ packed) {
  const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nq * k) return;
  const uint64_t q = i / k;
  const uint32_t j = (uint32_t)(i % k);
  packed[i] = (j < found[q]) ? ((dists[i] << 32) | (keys[i] + key_offset)) : kNone;
}

}  // namespace

void pack_device(const void* d_keys, const void* d_dists, const void* d_found, uint64_t nq, uint32_t k, uint64_t key_offset,
                 void* d_packed, cudaStream_t stream) {
  if (nq == 0) return;
  const uint64_t n = nq * k;
  pack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(static_cast<const uint64_t*>(d_keys), static_cast<const uint64_t*>(d_dists),
                                                               static_cast<const uint32_t*>(d_found), nq, k, key_offset,
                                                               static_cast<uint64_t*>(d_packed));
  count_launches();
  HB_CUDA(cudaGetLastError());
}

void topk_merge_device(const void* d_lists, uint32_t n_lists, uint64_t nq, uint32_t k, void* d_out, cudaStream_t stream, bool sorted,
                       uint32_t width) {
  if (width == 0) width = k;
  HB_REQUIRE(!sorted || width == k, HB_ERR_INVALID, "ascending lists have exactly k slots");
  HB_REQUIRE(k >= 1 && k <= HB_MAX_K, HB_ERR_INVALID, "k out of range");
  HB_REQUIRE(n_lists >= 1, HB_ERR_INVALID, "need at least one list");
  if (nq == 0) return;
  const size_t smem = (size_t)4 * k * 8;
  HB_CUDA(cudaFuncSetAttribute(topk_merge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  topk_merge_kernel<<<(unsigned)((nq + 3) / 4), 128, smem, stream>>>(static_cast<const uint64_t*>(d_lists), n_lists, nq, k,
                                                                     sorted ? 1u : 0u, width, static_cast<uint64_t*>(d_out));
    count_launches();
  HB_CUDA(cudaGetLastError());
}

void unpack_device(const void* d_packed, uint64_t n, void* d_keys, void* d_dists, cudaStream_t stream) {
  if (n == 0) return;
  unpack_kernel<<<(unsigned)((n + 255) / 256), 256, 0, stream>>>(static_cast<const uint64_t*>(d_packed), n,
                                                                 static_cast<uint64_t*>(d_keys), static_cast<uint64_t*>(d_dists));
    count_launches();
  HB_CUDA(cudaGetLastError());
}

void ground_truth_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                         void* d_packed, cudaStream_t stream, int algo, uint32_t* d_bound) {
  HB_REQUIRE(algo == HB_GT_AUTO || algo == HB_GT_SCAN || algo == HB_GT_MMA || algo == HB_GT_MXF4, HB_ERR_INVALID, "unknown ground-truth algorithm");
  if (algo == HB_GT_AUTO)
    algo = (nq >= 32 && n >= 4096 && ground_truth_mxf4_supported(d_data, d_queries, n, nq, k)) ? HB_GT_MXF4 : HB_GT_SCAN;
  if (algo == HB_GT_MXF4)
    ground_truth_mxf4_device(d_data, n, key_offset, d_queries, nq, k, d_packed, stream, d_bound);
  else if (algo == HB_GT_MMA)
    ground_truth_mma_device(d_data, n, key_offset, d_queries, nq, k, d_packed, stream);
  else
    ground_truth_scan_device(d_data, n, key_offset, d_queries, nq, k, d_packed, stream);
}

void ground_truth_scan_device(const void* d_data, uint64_t n, uint64_t key_offset, const void* d_queries, uint64_t nq, uint32_t k,
                              void* d_packed, cudaStream_t stream) {
  HB_REQUIRE(k >= 1 && k <= 1024, HB_ERR_UNSUPPORTED, "ground truth supports 1 <= k <= 1024");
  HB_REQUIRE(key_offset + n <= 0xFFFFFFFFull, HB_ERR_UNSUPPORTED, "keys must stay below 2^32");
  HB_REQUIRE(nq <= 0xFFFFFFF0ull, HB_ERR_UNSUPPORTED, "too many queries in one call");
  if (nq == 0) return;
  int dev = 0, sms = 0;
  HB_CUDA(cudaGetDevice(&dev));
  HB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  // queries per tile: as many as fit ~12 KB of lists per warp, at most 16
  uint32_t qt = std::max<uint32_t>(1, std::min<uint32_t>(16, (12u << 10) / (k * 8)));
  const uint32_t n_qtiles = (uint32_t)((nq + qt - 1) / qt);
  const uint64_t want_items = (uint64_t)sms * 64;
  uint64_t n_chunks = std::max<uint64_t>(1, (want_items + n_qtiles - 1) / n_qtiles);
  n_chunks = std::min<uint64_t>(n_chunks, std::max<uint64_t>(1, (n + 4095) / 4096));
  n_chunks = std::min<uint64_t>(n_chunks, 1024);
  uint64_t chunk_rows = std::max<uint64_t>(1, (n + n_chunks - 1) / n_chunks);
  chunk_rows = (chunk_rows + 31) / 32 * 32;
  n_chunks = std::max<uint64_t>(1, (n + chunk_rows - 1) / chunk_rows);

  uint64_t* partial = nullptr;
  HB_CUDA(cudaMallocAsync(&partial, n_chunks * nq * k * 8, stream));
  const size_t smem = (size_t)kScanWarps * ((size_t)qt * 128 + (size_t)qt * k * 8);
  HB_CUDA(cudaFuncSetAttribute(scan_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const uint64_t items = (uint64_t)n_qtiles * n_chunks;
  scan_topk_kernel<<<(unsigned)((items + kScanWarps - 1) / kScanWarps), kScanWarps * 32, smem, stream>>>(
      static_cast<const uint4*>(d_data), n, key_offset, static_cast<const uint4*>(d_queries), (uint32_t)nq, k, qt, n_qtiles,
      (uint32_t)n_chunks, chunk_rows, partial);
    count_launches();
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    try {
      topk_merge_device(partial, (uint32_t)n_chunks, nq, k, d_packed, stream);
    } catch (...) {
      cudaFreeAsync(partial, stream);
      throw;
    }
  }
  cudaFreeAsync(partial, stream);
  HB_CUDA(e);
}

}  // namespace hb
