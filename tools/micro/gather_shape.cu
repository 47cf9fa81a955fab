// Microbenchmark (tools/, not product): unloaded latency of one dependent gather step as a function of how the rows
// are requested.  mode 0: J x LDG.128 per lane, 8 lanes per row (rows = 4 J);  mode 1: cp.async 16 B into shared memory,
// same mapping, then LDS;  mode 2: one row per LDG.32 instruction (32 lanes x 4 B), rows = J instructions.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t hash32(uint32_t z) { z ^= z >> 16; z *= 0x7FEB352Du; z ^= z >> 15; z *= 0x846CA68Bu; z ^= z >> 16; return z; }
template <int J, int MODE>
__global__ void chase(const uint4* __restrict__ base, uint32_t n_rows, int steps, uint64_t* out, long long* cycles) {
  __shared__ __align__(128) uint4 stage[4][32 * 8];
  __shared__ __align__(8) uint64_t mbar[4];
  uint32_t phase = 0;
  if (MODE == 3) {
    if ((threadIdx.x & 31) == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&mbar[threadIdx.x >> 5])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
  }
  const uint32_t lane = threadIdx.x & 31, wib = threadIdx.x >> 5, g = lane >> 3, s = lane & 7;
  uint32_t state = 0x9E3779B1u * (blockIdx.x * 4 + wib + 1), acc = 0;
  const long long t0 = clock64();
  for (int it = 0; it < steps; ++it) {
    if (MODE == 0) {
      uint4 v[J];
#pragma unroll
      for (int j = 0; j < J; ++j) v[j] = __ldg(base + (uint64_t)__umulhi(hash32(state + 0x9E3779B1u * (4 * j + g + 1)), n_rows) * 8 + s);
#pragma unroll
      for (int j = 0; j < J; ++j) acc += __popc(v[j].x) + __popc(v[j].y) + __popc(v[j].z) + __popc(v[j].w);
    } else if (MODE == 1) {
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const uint4* src = base + (uint64_t)__umulhi(hash32(state + 0x9E3779B1u * (4 * j + g + 1)), n_rows) * 8 + s;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&stage[wib][(4 * j + g) * 8 + s]);
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src));
      }
      asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
      __syncwarp();
#pragma unroll
      for (int j = 0; j < J; ++j) { const uint4 v = stage[wib][(4 * j + g) * 8 + s]; acc += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w); }
    } else if (MODE == 3) {
      // TMA 1-D bulk copies: one 128-B row per lane and instruction (J x 32 rows per step), completion on an mbarrier
      const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&mbar[wib]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(J * 32 * 128) : "memory");
      __syncwarp();
#pragma unroll
      for (int j = 0; j < J; ++j) {
        const uint4* src = base + (uint64_t)__umulhi(hash32(state + 0x9E3779B1u * (32 * j + lane + 1)), n_rows) * 8;
        const uint32_t dst = (uint32_t)__cvta_generic_to_shared(&stage[wib][((32 * j + lane) % 32) * 8]);
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], 128, [%2];" ::"r"(dst), "l"(src), "r"(bar) : "memory");
      }
      uint32_t done = 0;
      while (!done) asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar), "r"(phase) : "memory");
      phase ^= 1u;
#pragma unroll
      for (int j = 0; j < 8; ++j) { const uint4 v = stage[wib][(4 * j + g) * 8 + s]; acc += __popc(v.x) + __popc(v.y) + __popc(v.z) + __popc(v.w); }
    } else {
      uint32_t v[J];
#pragma unroll
      for (int j = 0; j < J; ++j) v[j] = __ldg(reinterpret_cast<const uint32_t*>(base + (uint64_t)__umulhi(hash32(state + 0x9E3779B1u * (j + 1)), n_rows) * 8) + lane);
#pragma unroll
      for (int j = 0; j < J; ++j) acc += __popc(v[j]);
    }
    acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 1); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 2); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 4);
    acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 8); acc += __shfl_xor_sync(0xFFFFFFFFu, acc, 16);
    state = state * 747796405u + acc;
  }
  const long long t1 = clock64();
  if (lane == 0) { out[blockIdx.x * 4 + wib] = state; if (blockIdx.x == 0 && threadIdx.x == 0) *cycles = t1 - t0; }
}
template <int J, int MODE>
void run(const uint4* buf, uint32_t rows, uint64_t* out, long long* cyc, const char* name) {
  for (int loaded = 0; loaded < 2; ++loaded) {
    dim3 grid(loaded ? 148 * 6 : 1), block(loaded ? 128 : 32);
    chase<J, MODE><<<grid, block>>>(buf, rows, 50, out, cyc); cudaDeviceSynchronize();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int steps = 2000;
    cudaEventRecord(e0); chase<J, MODE><<<grid, block>>>(buf, rows, steps, out, cyc); cudaEventRecord(e1); cudaDeviceSynchronize();
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const int nrows = MODE == 2 ? J : MODE == 3 ? 32 * J : 4 * J;
    printf("%-28s rows/step %2d %s: %6.0f cycles/step %7.1f GB/s\n", name, nrows, loaded ? "24 warps/SM" : "1 warp     ", (double)*cyc / steps,
           (double)grid.x * (block.x / 32) * steps * nrows * 128 / (ms * 1e-3) / 1e9);
  }
}
int main() {
  const uint64_t bytes = 14ull << 30;
  uint4* buf; cudaMalloc(&buf, bytes); cudaMemset(buf, 0x5A, bytes);
  uint64_t* out; cudaMalloc(&out, 1 << 20); long long* cyc; cudaMallocManaged(&cyc, 8);
  const uint32_t rows = (uint32_t)(bytes / 128);
  run<1, 0>(buf, rows, out, cyc, "LDG.128 8 lanes/row");
  run<2, 0>(buf, rows, out, cyc, "LDG.128 8 lanes/row");
  run<4, 0>(buf, rows, out, cyc, "LDG.128 8 lanes/row");
  run<8, 0>(buf, rows, out, cyc, "LDG.128 8 lanes/row");
  run<2, 1>(buf, rows, out, cyc, "cp.async 16B 8 lanes/row");
  run<8, 1>(buf, rows, out, cyc, "cp.async 16B 8 lanes/row");
  run<1, 3>(buf, rows, out, cyc, "cp.async.bulk 128B row/lane");
  run<1, 2>(buf, rows, out, cyc, "LDG.32 one row/instr");
  run<8, 2>(buf, rows, out, cyc, "LDG.32 one row/instr");
  run<32, 2>(buf, rows, out, cyc, "LDG.32 one row/instr");
  return 0;
}
