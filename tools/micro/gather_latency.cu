// Microbenchmark (tools/, not product): latency of a dependent chain of 32-row x 128-B random gathers (the search
// kernel's access pattern: 8 lanes x LDG.128 per row, 8 loads in flight per lane) as a function of the region size,
// with 1 warp on the whole GPU (unloaded) and with 24 warps on every SM (loaded).  Answers: how much of the gather
// latency is address translation (region >> TLB reach) rather than DRAM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void chase(const uint4* __restrict__ base, uint64_t n_rows, int steps, uint64_t* out, long long* cycles) {
  const uint32_t lane = threadIdx.x & 31, g = lane >> 3, s = lane & 7;
  uint64_t state = 0x9E3779B97F4A7C15ull * (blockIdx.x * blockDim.x + threadIdx.x / 32 + 1);
  uint32_t acc = 0;
  const long long t0 = clock64();
  for (int it = 0; it < steps; ++it) {
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      uint32_t z = (uint32_t)state + 0x9E3779B1u * (uint32_t)(4 * j + g + 1);
      z ^= z >> 16; z *= 0x7FEB352Du; z ^= z >> 15; z *= 0x846CA68Bu; z ^= z >> 16;
      v[j] = __ldg(base + (uint64_t)__umulhi(z, (uint32_t)n_rows) * 8 + s);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) acc += __popc(v[j].x) + __popc(v[j].y) + __popc(v[j].z) + __popc(v[j].w);
    acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 1); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 2); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 4);
    acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 8); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 16);
    state = (uint32_t)state * 747796405u + acc;  // next addresses depend on the loaded data
  }
  const long long t1 = clock64();
  if (lane == 0) { out[blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32] = state; if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0; }
}
int main() {
  const uint64_t max_bytes = 32ull << 30;
  uint4* buf; cudaMalloc(&buf, max_bytes); cudaMemset(buf, 0x5A, max_bytes);
  uint64_t* out; cudaMalloc(&out, 1 << 20); long long* cyc; cudaMallocManaged(&cyc, 8);
  const int steps = 2000;
  for (uint64_t mb : {32ull, 128ull, 512ull, 1280ull, 4096ull, 14000ull, 32000ull}) {
    const uint64_t rows = (mb << 20) / 128;
    for (int loaded = 0; loaded < 2; ++loaded) {
      dim3 grid(loaded ? 148 * 6 : 1), block(loaded ? 128 : 32);
      chase<<<grid, block>>>(buf, rows, 50, out, cyc); cudaDeviceSynchronize();
      cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
      cudaEventRecord(e0); chase<<<grid, block>>>(buf, rows, steps, out, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      const double gbs = (double)grid.x * (block.x / 32) * steps * 32 * 128 / (ms * 1e-3) / 1e9;
      printf("region %6llu MB  %s: %8.0f cycles/step  %7.1f ns/step  %8.1f GB/s\n", (unsigned long long)mb, loaded ? "24 warps/SM" : "1 warp      ",
             (double)*cyc / steps, ms * 1e6 / steps, gbs);
    }
  }
  return 0;
}
