// gt_peak.cu -- measured tensor-pipe peak for the ground-truth roofline (hb_tensor_peak).
//
// MEASURED_PEAKS.json carries only a bf16 cuBLAS figure; the ground-truth kernels run tcgen05.mma kind::mxf4.block_scale
// (gt_mxf4.cu) and kind::i8 (gt_tcgen05.cu).  This probe issues the SAME instruction on the SAME tile shape and shared-
// memory layout (M = 128 x N = 240 x K = 64 E2M1, resp. M = 128 x N = 256 x K = 32 int8; K-major, 128-byte swizzle,
// two TMEM accumulators) from one elected thread per SM, with nothing else in the kernel: no operand expansion, no TMA,
// no epilogue.  Its Top/s is what the tensor pipe sustains for this instruction stream and is the denominator of
// `ground_truth.roofline` in bench.py.
#include <algorithm>

#include "gt_common.cuh"
#include "hb_internal.h"

namespace hb {
namespace {
using namespace gt;
constexpr auto mbar_wait = gt::mbar_wait<0>;

constexpr int kM = 128;
constexpr int kRing = 4;         // commit groups in flight
constexpr int kGroupMmas = 16;   // MMAs per commit group = one 1024-bit tile of the real kernels
constexpr uint32_t kSfCol = 496;

template <int KIND>  // 0 = mxf4 (N = 240, K = 64), 1 = i8 (N = 256, K = 32)
struct Shape {
  static constexpr int N = KIND == 0 ? 240 : 256;
  static constexpr int K = KIND == 0 ? 64 : 32;
  static constexpr uint32_t idesc = KIND == 0
      ? ((1u << 7) | (1u << 10) | (1u << 23) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kM >> 4) << 24))
      : ((2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(kM >> 4) << 24));
  static constexpr uint32_t a_bytes = kM * 128, b_bytes = 256 * 128;
  static constexpr uint32_t smem = a_bytes + b_bytes + 64 + 1024;
};

template <int KIND>
__global__ void __launch_bounds__(128, 1) tensor_peak_kernel(uint32_t groups) {
  using S = Shape<KIND>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  unsigned char* sm = smem_raw + (base - raw);
  const uint32_t sA = base, sB = base + S::a_bytes, sBar = sB + S::b_bytes;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sm + S::a_bytes + S::b_bytes + 48);
  const uint32_t warp = threadIdx.x >> 5;

  // operands: a fixed pseudo-random bit pattern (the values do not change the instruction's throughput)
  for (uint32_t i = threadIdx.x; i < (S::a_bytes + S::b_bytes) / 4; i += blockDim.x)
    reinterpret_cast<uint32_t*>(sm)[i] = KIND == 0 ? (0x2A2A2A2Au ^ ((i * 0x9E3779B1u) & 0x88888888u)) : (((i * 0x9E3779B1u) >> 7) & 0x01010101u);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (threadIdx.x == 0) {
    for (int r = 0; r < kRing; ++r) mbar_init(sBar + 8 * r, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  if (KIND == 0) {  // unit E8M0 scale factors in columns 496..511 of all 128 lanes
    const uint32_t one = 0x7F7F7F7Fu;
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1, %1};" ::"r"(
            tmem + ((warp * 32u) << 16) + kSfCol),
        "r"(one)
        : "memory");
    asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");

  if (warp == 0) {
    const uint64_t da = umma_desc(sA), db = umma_desc(sB);
    for (uint32_t g = 0; g < groups; ++g) {
      const uint32_t slot = g % kRing;
      if (g >= (uint32_t)kRing) mbar_wait(sBar + 8 * slot, ((g / kRing) - 1u) & 1u);  // the group that used this slot has retired
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (elect_one_sync()) {
        const uint32_t tmem_d = tmem + (g & 1u) * 256u;
#pragma unroll
        for (int j = 0; j < kGroupMmas; ++j) {  // +32 B inside the swizzle atom per K step = +2 in the encoded address
          if (KIND == 0) umma_mxf4(tmem_d, da + 2 * (j & 3), db + 2 * (j & 3), S::idesc, tmem + kSfCol, j != 0);
          else umma_i8(tmem_d, da + 2 * (j & 3), db + 2 * (j & 3), S::idesc, j != 0);
        }
        umma_commit(sBar + 8 * slot);
      }
      __syncwarp();
    }
    // drain: wait for the last kRing groups
    for (uint32_t g = groups > (uint32_t)kRing ? groups - kRing : 0; g < groups; ++g) mbar_wait(sBar + 8 * (g % kRing), (g / kRing) & 1u);
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
  }
}

template <int KIND>
double run_peak(int sm_count, cudaStream_t s) {
  using S = Shape<KIND>;
  HB_CUDA(cudaFuncSetAttribute(tensor_peak_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::smem));
  cudaEvent_t e0, e1;
  HB_CUDA(cudaEventCreate(&e0));
  HB_CUDA(cudaEventCreate(&e1));
  double best = 0.0;
  const uint32_t groups = 20000;  // x 16 MMAs x ~120 cycles = ~20 ms
  for (int rep = 0; rep < 4; ++rep) {  // first pass warms up
    HB_CUDA(cudaEventRecord(e0, s));
    tensor_peak_kernel<KIND><<<sm_count, 128, S::smem, s>>>(groups);
    count_launches();
    HB_CUDA(cudaGetLastError());
    HB_CUDA(cudaEventRecord(e1, s));
    HB_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    HB_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    const double ops = 2.0 * kM * S::N * S::K * (double)kGroupMmas * groups * sm_count;
    if (rep) best = std::max(best, ops / (ms * 1e-3) / 1e12);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return best;
}

}  // namespace
}  // namespace hb

using namespace hb;

extern "C" int hb_tensor_peak(int device, int kind, double* tops) {
  return guarded([&] {
    HB_REQUIRE(tops != nullptr && (kind == 0 || kind == 1), HB_ERR_INVALID, "kind: 0 = mxf4 (E2M1), 1 = i8");
    ensure_device(device);
    cudaDeviceProp prop;
    HB_CUDA(cudaGetDeviceProperties(&prop, device));
    *tops = kind == 0 ? run_peak<0>(prop.multiProcessorCount, nullptr) : run_peak<1>(prop.multiProcessorCount, nullptr);
  });
}
