// synth.cu -- synthetic 1024-bit sketches of laion2B-hammingv2 SHAPE, generated on
// the device (bench / test input only; not part of the reference's path).
//
// SimHash of a latent mixture, computed in exact integer arithmetic so that this
// kernel, the C oracle and numpy agree bit for bit (see DESIGN.md "synthetic
// data"): row i draws a cluster, z = centre_scale*centre + noise_scale*noise with
// components that are sums of 8 signed bytes of a splitmix64 stream, and bit j of
// the sketch is [<plane_j, z> > 0].
#include "hb_internal.h"

namespace hb {
namespace {

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t stream_key(uint64_t seed, uint64_t stream) {
  return mix64(seed ^ mix64(stream + 0x632BE59BD9B4E019ull));
}
__host__ __device__ __forceinline__ uint64_t hash_at(uint64_t key, uint64_t idx) {
  return mix64(key + (idx + 1) * 0x9E3779B97F4A7C15ull);
}
__host__ __device__ __forceinline__ int32_t gauss8(uint64_t r) {
  int32_t s = 4;
#pragma unroll
  for (int b = 0; b < 8; ++b) s += (int32_t)(int8_t)(uint8_t)(r >> (8 * b));
  return s;
}

__global__ void planes_kernel(uint64_t key, uint32_t count, int16_t* __restrict__ planes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) planes[i] = (int16_t)gauss8(hash_at(key, i));
}

constexpr int kRowsPerBlock = 16;

// block = 16 rows x 16 words; z of the 16 rows staged in shared memory
__global__ void __launch_bounds__(256) synth_kernel(hb_synth_params p, uint64_t kc, uint64_t kr0, uint64_t kr1,
                                                    const int16_t* __restrict__ planes, uint64_t start, uint64_t n,
                                                    uint64_t* __restrict__ out) {
  extern __shared__ int32_t z[];  // kRowsPerBlock * L
  const uint32_t L = p.latent_dim;
  const uint64_t row0 = (uint64_t)blockIdx.x * kRowsPerBlock;
  for (uint32_t e = threadIdx.x; e < kRowsPerBlock * L; e += blockDim.x) {
    const uint32_t r = e / L, t = e % L;
    const uint64_t i = start + row0 + r;
    int32_t v = 0;
    if (row0 + r < n) {
      const uint64_t cl = hash_at(kr0, i) % p.n_clusters;
      v = (int32_t)p.centre_scale * gauss8(hash_at(kc, cl * L + t)) + (int32_t)p.noise_scale * gauss8(hash_at(kr1, i * L + t));
    }
    z[e] = v;
  }
  __syncthreads();
  const uint32_t r = threadIdx.x >> 4, wd = threadIdx.x & 15;
  if (row0 + r >= n) return;
  const int32_t* zr = z + r * L;
  uint64_t word = 0;
  for (uint32_t b = 0; b < 64; ++b) {
    const uint32_t pj = (wd * 64 + b) % p.distinct_planes;
    const int16_t* a = planes + (size_t)pj * L;
    long long s = 0;
    for (uint32_t t = 0; t < L; ++t) s += (long long)((int32_t)a[t] * zr[t]);
    if (s > 0) word |= 1ull << b;
  }
  out[(row0 + r) * 16 + wd] = word;
}

}  // namespace

void synth_device(const hb_synth_params& p, uint64_t start, uint64_t n, void* d_out, cudaStream_t stream) {
  HB_REQUIRE(p.latent_dim >= 1 && p.latent_dim <= 512, HB_ERR_INVALID, "latent_dim out of range");
  HB_REQUIRE(p.n_clusters >= 1, HB_ERR_INVALID, "n_clusters must be positive");
  HB_REQUIRE(p.distinct_planes >= 1 && p.distinct_planes <= 1024, HB_ERR_INVALID, "distinct_planes out of range");
  // |plane| <= 1028, |z| <= (cs + ns) * 1028: each product must fit int32
  HB_REQUIRE((uint64_t)(p.centre_scale + p.noise_scale) * 1028ull * 1028ull < (1ull << 31), HB_ERR_INVALID, "scales too large");
  if (n == 0) return;
  int16_t* planes = nullptr;
  const uint32_t count = p.distinct_planes * p.latent_dim;
  HB_CUDA(cudaMallocAsync(&planes, (size_t)count * 2, stream));
  planes_kernel<<<(count + 255) / 256, 256, 0, stream>>>(stream_key(p.seed_planes, 0), count, planes);
    count_launches();
  const uint64_t blocks = (n + kRowsPerBlock - 1) / kRowsPerBlock;
  synth_kernel<<<(unsigned)blocks, 256, (size_t)kRowsPerBlock * p.latent_dim * 4, stream>>>(
      p, stream_key(p.seed_centres, 0), stream_key(p.seed_rows, 0), stream_key(p.seed_rows, 1), planes, start, n,
      static_cast<uint64_t*>(d_out));
    count_launches();
  cudaError_t e = cudaGetLastError();
  cudaFreeAsync(planes, stream);
  HB_CUDA(e);
}

}  // namespace hb
